"""Trainer runtime, exposed under the reference's import path ``llm_training.lightning``
(reference: src/llm_training/lightning/__init__.py:1-4)."""
from .callbacks import (Callback, ExtraConfig, LearningRateMonitor, ModelCheckpoint, OutputRedirection,
                        SaveConfigCallback, StepTimeMonitor, TQDMProgressBar, TrainingTimeEstimator)
from .loggers import CSVLogger, WandbLogger
from .module import LightningDataModule, LightningModule
from .strategy import DeepSpeedStrategy, FSDP2Precision, FSDP2Strategy, SingleDeviceStrategy, Strategy
from .trainer import Trainer, seed_everything
from .cli import HFTokenizer, LightningCLI
