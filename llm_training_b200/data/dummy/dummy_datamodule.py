"""(reference: src/llm_training/data/dummy/dummy_datamodule.py:7-20)"""
from ..base_datamodule import BaseDataModule, DatasetDict
from .dummy_datamodule_config import DummyDataModuleConfig
from .dummy_dataset import DummyDataset


class DummyDataModule(BaseDataModule):
    config: DummyDataModuleConfig

    def __init__(self, config: DummyDataModuleConfig) -> None:
        super().__init__(config)

    def prepare_data(self) -> None: ...

    def setup(self, stage: str | None = None) -> None:
        if self.trainer is not None:   # every rank generates from the same base seed
            self.config.base_seed = self.trainer.strategy.broadcast(self.config.base_seed)
        super().setup(stage)

    def load_data(self) -> DatasetDict:
        return {'train': DummyDataset(self.config)}
