from .dummy_datamodule import DummyDataModule
from .dummy_datamodule_config import DummyDataModuleConfig
from .dummy_dataset import DummyDataset
