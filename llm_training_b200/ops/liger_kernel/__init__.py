"""``llm_training.ops.liger_kernel``: the reference routes these names to Liger's Triton kernels
(src/llm_training/ops/liger_kernel/*.py).  Here they are the hand-written sm_100a kernels of this repository (CPU
tensors take the fp32 PyTorch restatements), kept under the same import path so user code keeps working."""
from .cross_entropy_op import cross_entropy, fused_linear_cross_entropy
from .rms_norm_op import rms_norm
from .rope_op import apply_rope
from .swiglu_op import silu_mul, swiglu

__all__ = ['cross_entropy', 'fused_linear_cross_entropy', 'rms_norm', 'apply_rope', 'silu_mul', 'swiglu']
