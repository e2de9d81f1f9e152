from .instruction_tuning_datacollator import InstructionTuningDataCollator
from .instruction_tuning_datamodule import InstructionTuningDataModule
from .instruction_tuning_datamodule_config import (InstructionTuningDataModuleConfig, OverlongHandlingMethod,
                                                   PackingMethod)
