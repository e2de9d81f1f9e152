"""Rotary position embedding, rotate-half convention
(reference: src/llm_training/ops/rope_op.py:4-20; cos/sin cache at models/llama/llama_model.py:367-412)."""
import torch

from . import _lib
from .reference import apply_rope_ref, rotate_half  # noqa: F401  (re-exported)


def apply_rope(q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor
               ) -> tuple[torch.Tensor, torch.Tensor]:
    """Reference-layout API: q/k ``[B, H, S, D]``, cos/sin ``[B, S, D]``. Out of place."""
    return apply_rope_ref(q, k, cos, sin)


class _RopeQKFn(torch.autograd.Function):
    """In place on the q|k columns of a packed ``[T, (Hq + Hkv [+ Hkv]) * D]`` projection output."""

    @staticmethod
    def forward(ctx, qkv: torch.Tensor, pos: torch.Tensor, cos_t: torch.Tensor, sin_t: torch.Tensor,
                n_rot_heads: int, head_dim: int) -> torch.Tensor:
        tokens, width = qkv.shape
        _lib.check(_lib.lib().llmt_rope_inplace(_lib.ptr(qkv), _lib.ptr(pos), _lib.ptr(cos_t), _lib.ptr(sin_t),
                                                _lib.c_int(tokens), _lib.c_int(n_rot_heads), _lib.c_int(head_dim),
                                                _lib.c_i64(qkv.stride(0)), _lib.c_int(0), _lib.stream()), 'rope')
        ctx.mark_dirty(qkv)
        ctx.save_for_backward(pos, cos_t, sin_t)
        ctx.meta = (n_rot_heads, head_dim)
        return qkv

    @staticmethod
    def backward(ctx, d: torch.Tensor):
        pos, cos_t, sin_t = ctx.saved_tensors
        n_rot_heads, head_dim = ctx.meta
        d = d.clone(memory_format=torch.contiguous_format)   # never rotate the caller's gradient in place
        tokens = d.shape[0]
        _lib.check(_lib.lib().llmt_rope_inplace(_lib.ptr(d), _lib.ptr(pos), _lib.ptr(cos_t), _lib.ptr(sin_t),
                                                _lib.c_int(tokens), _lib.c_int(n_rot_heads), _lib.c_int(head_dim),
                                                _lib.c_i64(d.stride(0)), _lib.c_int(1), _lib.stream()), 'rope_bwd')
        return d, None, None, None, None, None


def rope_qk_packed(qkv: torch.Tensor, pos: torch.Tensor, cos_table: torch.Tensor, sin_table: torch.Tensor,
                   n_rot_heads: int, head_dim: int) -> torch.Tensor:
    """Rotate the first ``n_rot_heads`` heads of every row of ``qkv [T, H_total*D]``.

    ``cos_table``/``sin_table`` are fp32 ``[max_pos, D/2]``; ``pos`` is int32 ``[T]``.
    """
    if _lib.use_kernels(qkv):
        return _RopeQKFn.apply(qkv, pos, cos_table, sin_table, n_rot_heads, head_dim)
    t, width = qkv.shape
    heads_total = width // head_dim
    x = qkv.view(t, heads_total, head_dim)
    cos = cos_table[pos.long()]
    sin = sin_table[pos.long()]
    cos = torch.cat((cos, cos), dim=-1)
    sin = torch.cat((sin, sin), dim=-1)
    rot, _ = apply_rope_ref(x[:, :n_rot_heads], x[:, :0], cos, sin)
    return torch.cat((rot, x[:, n_rot_heads:]), dim=1).reshape(t, width)
