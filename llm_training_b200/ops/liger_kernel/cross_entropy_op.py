"""Same entry points as the reference's Liger wrappers (ops/liger_kernel/cross_entropy_op.py:12-54)."""
from ..cross_entropy_op import cross_entropy, fused_linear_cross_entropy

__all__ = ['cross_entropy', 'fused_linear_cross_entropy']
