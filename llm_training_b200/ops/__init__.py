"""Hot-path ops.  Each op has (a) an sm_100a kernel in ``csrc/kernels`` used for CUDA bf16 tensors and
(b) a plain-PyTorch fp32 restatement in :mod:`.reference` used on CPU and as the test oracle.

Public names follow the reference (src/llm_training/ops/__init__.py:1-4 plus
ops/liger_kernel/__init__.py and ops/attention_op.py).
"""
from .attention_op import (flash_attention_forward, get_unpad_data, packed_attention,
                           prepare_4d_causal_attention_mask)
from .cross_entropy_op import (cross_entropy, fused_linear_cross_entropy,
                               fused_linear_logprob, shift_labels)
from .embedding_op import embedding, neftune_noise_scale
from .fused_blocks import norm_multi_linear, swiglu_down_linear
from .linear_op import linear, multi_linear
from .rms_norm_op import rms_norm
from .rope_op import apply_rope, rope_qk_packed, rotate_half
from .rope_utils import ROPE_INIT_FUNCTIONS, RoPEConfig
from .segments import segment_bounds
from .swiglu_op import silu_mul, silu_mul_fused, swiglu
