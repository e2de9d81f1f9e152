"""Import location kept from the reference (src/llm_training/lightning/strategy/fsdp2/fsdp2_precision.py)."""
from ..precision import CPUOffloadPolicy, FSDP2Precision, MixedPrecisionPolicy, OffloadPolicy  # noqa: F401
