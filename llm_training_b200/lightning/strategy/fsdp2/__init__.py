from .fsdp2_precision import FSDP2Precision, MixedPrecisionPolicy, OffloadPolicy
from .fsdp2_strategy import FSDP2Strategy
