"""Data layer (reference: src/llm_training/data/__init__.py:1-8)."""
from .base_datacollator import BaseDataCollator
from .base_datamodule import BaseDataModule
from .base_datamodule_config import BaseDataModuleConfig
from .dummy import DummyDataModule, DummyDataModuleConfig, DummyDataset
from .hf_based import HFBasedDataModule, HFBasedDataModuleConfig
from .instruction_tuning import (InstructionTuningDataCollator, InstructionTuningDataModule,
                                 InstructionTuningDataModuleConfig)
from .pre_training import PreTrainingDataCollator, PreTrainingDataModule, PreTrainingDataModuleConfig
from .preference_tuning import (PreferenceTuningDataCollator, PreferenceTuningDataModule,
                                PreferenceTuningDataModuleConfig)
from .resumable_dataloader import ResumableDataLoader
