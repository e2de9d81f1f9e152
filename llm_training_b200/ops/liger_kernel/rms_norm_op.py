"""Same entry point as the reference's Liger wrapper (ops/liger_kernel/rms_norm_op.py:7-19)."""
from ..rms_norm_op import rms_norm

__all__ = ['rms_norm']
