"""Same entry point as the reference's Liger wrapper (ops/liger_kernel/rope_op.py:7-21)."""
from ..rope_op import apply_rope

__all__ = ['apply_rope']
