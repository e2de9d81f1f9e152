"""Plain-PyTorch fp32 restatements of every hot op.

These are (a) the numerical oracle for the sm_100a kernels in ``csrc/kernels`` and (b) the code
path used on CPU tensors (unit tests, gloo multi-process tests).  They follow the math of the
reference's torch ops — src/llm_training/ops/rms_norm_op.py:4-14, rope_op.py:4-20,
swiglu_op.py:5-29, cross_entropy_op.py:4-8, attention_op.py:286-372 — restated for the packed
token layout used throughout this framework.
"""
import math

import torch
import torch.nn.functional as F


def rms_norm_ref(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    xf = x.float()
    rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (xf * rstd * weight.float()).to(x.dtype)


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    half = x.shape[-1] // 2
    return torch.cat((-x[..., half:], x[..., :half]), dim=-1)


def apply_rope_ref(q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor,
                   head_dim_axis_layout: str = 'bhsd') -> tuple[torch.Tensor, torch.Tensor]:
    """``bhsd``: q/k are [B, H, S, D], cos/sin [B, S, D] (reference layout).
    ``thd``: q/k are [T, H, D], cos/sin [T, D] (packed layout)."""
    unsq = 1
    cos_ = cos.unsqueeze(unsq).float()
    sin_ = sin.unsqueeze(unsq).float()
    qf, kf = q.float(), k.float()
    q_out = qf * cos_ + rotate_half(qf) * sin_
    k_out = kf * cos_ + rotate_half(kf) * sin_
    return q_out.to(q.dtype), k_out.to(k.dtype)


def silu_mul_ref(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    return (F.silu(gate.float()) * up.float()).to(gate.dtype)


def cross_entropy_ref(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100,
                      reduction: str = 'mean') -> torch.Tensor:
    return F.cross_entropy(logits.float(), labels, ignore_index=ignore_index, reduction=reduction)


def linear_cross_entropy_ref(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                             ignore_index: int = -100) -> torch.Tensor:
    logits = hidden.float() @ weight.float().t()
    return cross_entropy_ref(logits, labels, ignore_index)


def token_logprob_ref(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                      ignore_index: int = -100) -> torch.Tensor:
    """log p(label_t | ...) per token, 0 where ``labels == ignore_index``
    (reference: src/llm_training/lms/dpo/dpo.py:110-114 before the per-sequence reduction)."""
    logits = hidden.float() @ weight.float().t()
    logp = torch.log_softmax(logits, dim=-1)
    safe = labels.clamp_min(0)
    picked = logp.gather(-1, safe.unsqueeze(-1)).squeeze(-1)
    return picked * labels.ne(ignore_index)


def attention_mask_from_bounds(seg_start: torch.Tensor, window: int | None = None) -> torch.Tensor:
    """Boolean [T, T] mask, True = may attend. Row i sees j with seg_start[i] <= j <= i."""
    t = seg_start.numel()
    idx = torch.arange(t, device=seg_start.device)
    allowed = (idx[None, :] <= idx[:, None]) & (idx[None, :] >= seg_start[:, None])
    if window is not None and window > 0:
        allowed &= idx[None, :] >= (idx[:, None] - window)      # `window` keys back + the token itself (HF / flash-attn)
    return allowed


def dropout_keep_mask(seed: int, num_heads: int, tokens: int, drop_thr: int, device: torch.device | str = 'cpu'
                      ) -> torch.Tensor:
    """Boolean ``[Hq, T, T]`` keep-mask of attention dropout, bit-identical to the kernels' (csrc/kernels/common.cuh
    ``dropout_bits``): element (head, q, kv) is dropped when byte ``kv % 4`` of the 32-bit hash of
    ``(seed, head, q, kv // 4)`` is below ``drop_thr``."""
    m = 0xffffffff
    h = torch.arange(num_heads, dtype=torch.int64, device=device).view(-1, 1, 1)
    q = torch.arange(tokens, dtype=torch.int64, device=device).view(1, -1, 1)
    kv = torch.arange(tokens, dtype=torch.int64, device=device).view(1, 1, -1)
    x = (seed + h * 0xC2B2AE3D + q * 0x9E3779B1 + (kv // 4) * 0x85EBCA77) & m
    x = x ^ (x >> 16)
    x = (x * 0x7feb352d) & m
    x = x ^ (x >> 15)
    x = (x * 0x846ca68b) & m
    x = x ^ (x >> 16)
    byte = (x >> (8 * (kv % 4))) & 0xff
    return byte >= drop_thr


def attention_ref(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, seg_start: torch.Tensor,
                  scale: float | None = None, window: int | None = None,
                  softcap: float = 0.0, dropout_keep: torch.Tensor | None = None,
                  dropout_p: float = 0.0) -> torch.Tensor:
    """Packed causal attention with per-token document lower bounds.

    q: [T, Hq, D]; k, v: [T, Hkv, D]; seg_start: int [T]. Returns [T, Hq, D].  ``dropout_keep``: boolean
    ``[Hq, T, T]`` (True = keep), applied to the normalised probabilities with the usual ``1 / (1 - p)`` scaling.
    """
    t, hq, d = q.shape
    hkv = k.shape[1]
    scale = (1.0 / math.sqrt(d)) if scale is None else scale
    qf = q.float().transpose(0, 1)                                   # [Hq, T, D]
    kf = k.float().transpose(0, 1).repeat_interleave(hq // hkv, 0)   # [Hq, T, D]
    vf = v.float().transpose(0, 1).repeat_interleave(hq // hkv, 0)
    s = torch.matmul(qf, kf.transpose(1, 2)) * scale
    if softcap and softcap > 0:
        s = softcap * torch.tanh(s / softcap)
    allowed = attention_mask_from_bounds(seg_start, window)
    s = s.masked_fill(~allowed[None], float('-inf'))
    p = torch.softmax(s, dim=-1)
    if dropout_keep is not None:
        p = p * dropout_keep.to(p.dtype) / (1.0 - dropout_p)
    o = torch.matmul(p, vf)                                          # [Hq, T, D]
    return o.transpose(0, 1).to(q.dtype)


def adamw_ref(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, *, lr: float,
              beta1: float, beta2: float, eps: float, weight_decay: float, step: int,
              grad_scale: float = 1.0) -> None:
    """In-place decoupled AdamW on fp32 master ``p`` (torch.optim.AdamW semantics)."""
    gf = g.float() * grad_scale
    p.mul_(1.0 - lr * weight_decay)
    m.mul_(beta1).add_(gf, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(gf, gf, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
