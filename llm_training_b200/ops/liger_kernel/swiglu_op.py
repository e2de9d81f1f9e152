"""Same entry points as the reference's Liger wrappers (ops/liger_kernel/swiglu_op.py:6-39)."""
from ..swiglu_op import silu_mul, swiglu

__all__ = ['silu_mul', 'swiglu']
