from .base import SingleDeviceStrategy, Strategy
from .deepspeed import DeepSpeedStrategy
from .fsdp2 import FSDP2Precision, FSDP2Strategy

__all__ = ['Strategy', 'SingleDeviceStrategy', 'FSDP2Strategy', 'FSDP2Precision', 'DeepSpeedStrategy']
