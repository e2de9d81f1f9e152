"""Building blocks shared by the decoder-only models (Llama, Phi-3).

All blocks work on the PACKED token layout: activations are ``[T, hidden]`` with ``T = B * S`` and a
:class:`PackedContext` carries what the reference recomputes per layer (mask -> cu_seqlens with a
host sync, attention_op.py:375-412; cos/sin gather with ``.item()``, llama_model.py:389-412):
document bounds, int32 positions and the fp32 RoPE tables — built once per batch, on device.
"""
import math
from dataclasses import dataclass
from typing import Any

import torch
import torch.nn.functional as F
import torch.utils.checkpoint
from torch import nn

from ..ops import (packed_attention, prepare_4d_causal_attention_mask, rms_norm, rope_qk_packed,
                   segment_bounds)
from ..ops.rope_utils import ROPE_INIT_FUNCTIONS, RoPEConfig


@dataclass
class PackedContext:
    batch_size: int
    seq_len: int
    seg_start: torch.Tensor          # int32 [T]
    seg_end: torch.Tensor            # int32 [T]
    positions: torch.Tensor          # int32 [T]
    cos: torch.Tensor                # fp32 [P, D/2] (attention scaling folded in)
    sin: torch.Tensor
    attention_mask: torch.Tensor | None = None   # original [B, S] segment ids (eager / sdpa paths)
    tp: Any = None                   # parallel.TensorParallelContext or None

    @property
    def tokens(self) -> int:
        return self.batch_size * self.seq_len


class RMSNorm(nn.Module):
    """``weight * x / rms(x)`` in fp32 (reference: LlamaRMSNorm, llama_model.py:271-286)."""

    def __init__(self, hidden_size: int, eps: float) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.variance_epsilon = eps

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        return rms_norm(hidden_states, self.weight, self.variance_epsilon)

    def extra_repr(self) -> str:
        return f'{tuple(self.weight.shape)}, eps={self.variance_epsilon}'


class RotaryEmbedding(nn.Module):
    """fp32 cos/sin tables ``[P, D/2]`` for all six RoPE variants.

    The table is sized from static information only — ``max(max_position_embeddings, seq_len)`` rounded
    up to 4096 like the reference (llama_model.py:371) — so no device->host sync is needed.  Position
    ids must therefore be smaller than that bound (true for every collator in ``llm_training.data``).
    """

    def __init__(self, head_dim: int, rope_theta: float, max_position_embeddings: int,
                 rope_scaling: dict[str, Any] | None, original_max_position_embeddings: int | None = None) -> None:
        super().__init__()
        scaling = dict(rope_scaling) if rope_scaling is not None else None
        self.rope_type = 'default'
        if scaling is not None:
            self.rope_type = scaling.get('rope_type', scaling.get('type', 'default'))
        self.max_position_embeddings = max_position_embeddings
        # LongRoPE measures "long" against the pre-extension context length
        self.original_max_seq_len = original_max_position_embeddings or max_position_embeddings
        cfg_scaling = None
        if scaling is not None:
            cfg_scaling = {k: v for k, v in scaling.items() if k not in ('type', 'rope_type')}
            if self.rope_type == 'longrope' and 'factor' not in cfg_scaling:
                cfg_scaling['factor'] = max_position_embeddings / self.original_max_seq_len
        self.rope_config = RoPEConfig(
            type=self.rope_type, base=rope_theta, dim=head_dim,
            max_position_embeddings=self.original_max_seq_len if self.rope_type == 'longrope' else max_position_embeddings,
            scaling_config=cfg_scaling)
        self.rope_init_fn = ROPE_INIT_FUNCTIONS[self.rope_type]
        self._cached_len = 0
        self._cached_regime: Any = None
        self._cos: torch.Tensor | None = None
        self._sin: torch.Tensor | None = None
        self.attention_scaling = 1.0

    def tables(self, seq_len: int, device: torch.device) -> tuple[torch.Tensor, torch.Tensor]:
        need = max(self.max_position_embeddings, seq_len)
        need = math.ceil(need / 4096) * 4096
        # sequence-length dependent variants: the frequency set itself changes with the regime
        if self.rope_type == 'dynamic':
            regime = need if seq_len > self.max_position_embeddings else 0
        elif self.rope_type == 'longrope':
            regime = seq_len > self.original_max_seq_len
        else:
            regime = None
        if (self._cos is None or self._cos.device != device or need > self._cached_len
                or regime != self._cached_regime):
            arg_len = seq_len if self.rope_type in ('dynamic', 'longrope') else None
            inv_freq, self.attention_scaling = self.rope_init_fn(self.rope_config, device, arg_len)
            t = torch.arange(need, device=device, dtype=torch.float32)
            freqs = torch.outer(t, inv_freq.to(device=device, dtype=torch.float32))
            self._cos = (freqs.cos() * self.attention_scaling).contiguous()
            self._sin = (freqs.sin() * self.attention_scaling).contiguous()
            self._cached_len, self._cached_regime = need, regime
        return self._cos, self._sin

    @torch.no_grad()
    def forward(self, x: torch.Tensor, position_ids: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        """Reference-shaped output: cos/sin ``[B, S, D]`` gathered at ``position_ids`` (llama_model.py:389-412)."""
        cos, sin = self.tables(position_ids.shape[-1], x.device)
        c, s = cos[position_ids.long()], sin[position_ids.long()]
        return torch.cat((c, c), -1).to(x.dtype), torch.cat((s, s), -1).to(x.dtype)


def build_packed_context(rotary: RotaryEmbedding, batch_size: int, seq_len: int, device: torch.device,
                         attention_mask: torch.Tensor | None, position_ids: torch.Tensor | None,
                         tp: Any = None) -> PackedContext:
    seg_start, seg_end = segment_bounds(attention_mask, batch_size, seq_len, device)
    if position_ids is None:
        positions = torch.arange(seq_len, device=device, dtype=torch.int32).repeat(batch_size)
    else:
        positions = position_ids.to(device=device, dtype=torch.int32).reshape(-1)
        if positions.numel() == seq_len and batch_size > 1:
            positions = positions.repeat(batch_size)
    cos, sin = rotary.tables(seq_len, device)
    return PackedContext(batch_size, seq_len, seg_start, seg_end, positions.contiguous(), cos, sin,
                         attention_mask=attention_mask, tp=tp)


def _eager_or_sdpa_attention(qkv: torch.Tensor, ctx: PackedContext, hq: int, hkv: int, d: int, impl: str,
                             window: int | None, dropout_p: float, training: bool) -> torch.Tensor:
    """4-D-mask attention (reference: LlamaAttention / LlamaSdpaAttention, llama_model.py:434-744)."""
    b, s = ctx.batch_size, ctx.seq_len
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    q = q.reshape(b, s, hq, d).transpose(1, 2)
    k = k.reshape(b, s, hkv, d).transpose(1, 2).repeat_interleave(hq // hkv, dim=1)
    v = v.reshape(b, s, hkv, d).transpose(1, 2).repeat_interleave(hq // hkv, dim=1)
    mask = prepare_4d_causal_attention_mask(ctx.attention_mask, (b, s), q, 0, window)
    p = dropout_p if training else 0.0
    if impl == 'sdpa':
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=p)
    else:
        w = torch.matmul(q, k.transpose(2, 3)) / math.sqrt(d) + mask
        w = F.softmax(w, dim=-1, dtype=torch.float32).to(v.dtype)
        w = F.dropout(w, p=p, training=training)
        o = torch.matmul(w, v)
    return o.transpose(1, 2).reshape(b * s, hq * d)


def dense_core_attention(query_states: torch.Tensor, key_states: torch.Tensor, value_states: torch.Tensor,
                         attention_mask: torch.Tensor | None, dropout_p: float = 0.0, training: bool = False,
                         impl: str = 'eager') -> torch.Tensor:
    """``[B, H, S, D]`` q / k / v (KV heads already repeated) + additive ``[B, 1, S, S]`` mask -> ``[B, S, H*D]``: the
    contract of the reference's ``core_attention_forward`` (llama_model.py:506-537, phi3_model.py:436-480).  The
    training path of this repository never materialises these tensors (``attention_core`` below)."""
    p = dropout_p if training else 0.0
    if impl == 'sdpa':
        o = F.scaled_dot_product_attention(query_states, key_states, value_states, attn_mask=attention_mask, dropout_p=p)
    else:
        w = torch.matmul(query_states, key_states.transpose(2, 3)) / math.sqrt(query_states.shape[-1])
        if attention_mask is not None:
            w = w + attention_mask
        w = F.softmax(w, dim=-1, dtype=torch.float32).to(value_states.dtype)
        w = F.dropout(w, p=p, training=training)
        o = torch.matmul(w, value_states)
    b, h, s, d = o.shape
    return o.transpose(1, 2).reshape(b, s, h * d)


def attention_core(qkv: torch.Tensor, ctx: PackedContext, *, num_heads: int, num_kv_heads: int, head_dim: int,
                   impl: str, window: int | None = None, dropout_p: float = 0.0, training: bool = False,
                   selective_recompute: bool = False) -> torch.Tensor:
    """RoPE + attention on the fused ``q|k|v`` projection output ``[T, (Hq+2Hkv)*D]`` -> ``[T, Hq*D]``."""
    qkv = rope_qk_packed(qkv, ctx.positions, ctx.cos, ctx.sin, num_heads + num_kv_heads, head_dim)
    use_kernel_path = impl == 'flash_attention_2'
    p_drop = dropout_p if training else 0.0
    # one seed per call, drawn on the host: activation recompute re-runs `core` with the same mask
    seed = int(torch.randint(0, 2 ** 31 - 1, (1,)).item()) if (use_kernel_path and p_drop > 0.0) else None

    def core(t: torch.Tensor) -> torch.Tensor:
        if use_kernel_path:
            return packed_attention(t, ctx.seg_start, ctx.seg_end, num_heads, num_kv_heads, head_dim, None, window,
                                    dropout_p=p_drop, seed=seed)
        return _eager_or_sdpa_attention(t, ctx, num_heads, num_kv_heads, head_dim,
                                        'eager' if impl == 'flash_attention_2' else impl, window, dropout_p, training)

    if selective_recompute and training and torch.is_grad_enabled():
        return torch.utils.checkpoint.checkpoint(core, qkv, use_reentrant=False)
    return core(qkv)


def repeat_kv(hidden_states: torch.Tensor, n_rep: int) -> torch.Tensor:
    """``[B, Hkv, S, D] -> [B, Hkv * n_rep, S, D]`` (reference: llama_model.py:430-431).  Only the eager / SDPA
    fallbacks need it; the packed attention kernels read each KV head once for its whole query group."""
    if n_rep == 1:
        return hidden_states
    b, h, s, d = hidden_states.shape
    return hidden_states[:, :, None].expand(b, h, n_rep, s, d).reshape(b, h * n_rep, s, d)
