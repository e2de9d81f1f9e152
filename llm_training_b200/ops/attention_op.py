"""Attention: packed varlen causal attention (the hot path) plus the 4-D-mask utilities of the
reference for the ``eager`` / ``sdpa`` implementations.

Reference: src/llm_training/ops/attention_op.py — ``prepare_4d_causal_attention_mask`` (:317-372),
packed-sequence mask (:286-314), segment ids -> cu_seqlens (:375-412), ``flash_attention_forward``
(:538-654).  The flash-attn 2.8 CUDA kernels it calls are replaced by ``csrc/kernels/attention.cu``.
"""
import ctypes
import math
import os

import torch

from . import _lib
from .reference import attention_mask_from_bounds, attention_ref, dropout_keep_mask
from .segments import segment_bounds


class _PackedAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv: torch.Tensor, seg_start: torch.Tensor, seg_end: torch.Tensor, hq: int, hkv: int,
                head_dim: int, scale: float, window: int, softcap: float, drop_thr: int, seed: int,
                deterministic: bool = False) -> torch.Tensor:
        L = _lib.lib()
        tokens = qkv.shape[0]
        out = torch.empty((tokens, hq * head_dim), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((hq, tokens), dtype=torch.float32, device=qkv.device)
        _lib.check(L.llmt_attn_fwd_ex(_lib.ptr(qkv), _lib.ptr(out), _lib.ptr(lse), _lib.ptr(seg_start),
                                      _lib.c_int(tokens), _lib.c_int(hq), _lib.c_int(hkv), _lib.c_int(head_dim),
                                      _lib.c_i64(qkv.stride(0)), _lib.c_f32(scale), _lib.c_int(window),
                                      _lib.c_f32(softcap), _lib.c_int(drop_thr), ctypes.c_uint32(seed),
                                      _lib.stream()), 'attn_fwd')
        ctx.save_for_backward(qkv, out, lse, seg_start, seg_end)
        ctx.meta = (hq, hkv, head_dim, scale, window, softcap, drop_thr, seed, deterministic)
        return out

    @staticmethod
    def backward(ctx, dout: torch.Tensor):
        qkv, out, lse, seg_start, seg_end = ctx.saved_tensors
        hq, hkv, head_dim, scale, window, softcap, drop_thr, seed, deterministic = ctx.meta
        L = _lib.lib()
        tokens = qkv.shape[0]
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        delta = torch.empty((hq, tokens), dtype=torch.float32, device=qkv.device)
        # fp32 dQ accumulator [T, Hq*D] (scratch), converted into dqkv by the kernel's final pass
        dq_acc = torch.zeros((tokens, hq * head_dim), dtype=torch.float32, device=qkv.device)
        # deterministic: per (head, 64-query block) turn counters, so the kv blocks add into dq_acc in a fixed order
        order = torch.empty(hq * ((tokens + 63) // 64), dtype=torch.int32, device=qkv.device) if deterministic else None
        # few kv heads on this GPU (tensor-parallel shard): several CTAs share each kv block and add fp32 dK / dV tiles
        dkv_acc = None
        if not deterministic and L.llmt_attn_bwd_splits(_lib.c_int(tokens), _lib.c_int(hkv)) > 1:
            dkv_acc = torch.empty((tokens, 2 * hkv * head_dim), dtype=torch.float32, device=qkv.device)
        _lib.check(L.llmt_attn_bwd_ex(_lib.ptr(qkv), _lib.ptr(out), _lib.ptr(dout), _lib.ptr(lse), _lib.ptr(delta),
                                      _lib.ptr(dq_acc), _lib.ptr(dqkv), _lib.ptr(seg_start), _lib.ptr(seg_end),
                                      _lib.c_int(tokens), _lib.c_int(hq), _lib.c_int(hkv), _lib.c_int(head_dim),
                                      _lib.c_i64(qkv.stride(0)), _lib.c_f32(scale), _lib.c_int(window),
                                      _lib.c_f32(softcap), _lib.c_int(drop_thr), ctypes.c_uint32(seed),
                                      _lib.ptr(order), _lib.ptr(dkv_acc), _lib.stream()), 'attn_bwd')
        return (dqkv,) + (None,) * 11


def deterministic_default() -> bool:
    """The reference reads ``FLASH_ATTENTION_DETERMINISTIC=1`` when the caller passes ``deterministic=None``
    (attention_op.py:590-591); ``torch.use_deterministic_algorithms(True)`` (what ``Trainer(deterministic=True)`` sets)
    asks for the same thing."""
    return os.environ.get('FLASH_ATTENTION_DETERMINISTIC', '0') == '1' or torch.are_deterministic_algorithms_enabled()


def dropout_threshold(dropout_p: float) -> int:
    """The kernels drop an attention probability when its random byte is below this threshold: p is quantised to a
    multiple of 1/256 (flash-attn quantises the same way)."""
    thr = int(round(float(dropout_p) * 256))
    if not 0 <= thr < 256:
        raise ValueError(f'attention dropout p={dropout_p} out of range [0, 255/256]')
    return thr


def packed_attention(qkv: torch.Tensor, seg_start: torch.Tensor, seg_end: torch.Tensor, num_heads: int,
                     num_kv_heads: int, head_dim: int, scale: float | None = None,
                     window: int | None = None, dropout_p: float = 0.0, softcap: float = 0.0,
                     seed: int | None = None, deterministic: bool | None = None) -> torch.Tensor:
    """Causal attention over a packed token stream.

    ``qkv`` is ``[T, (Hq + 2*Hkv) * D]`` laid out ``q heads | k heads | v heads`` per row (the output
    of the fused QKV projection, RoPE already applied).  Token ``t`` attends keys ``j`` with
    ``seg_start[t] <= j <= t`` (and ``j >= t - window`` if a sliding window is set: ``window`` keys back plus the
    token itself, the meaning of flash-attn's ``window_size=(window, window)`` and of the reference's eager mask,
    attention_op.py:160-161,588).  ``softcap`` > 0 replaces the scaled scores by ``softcap * tanh(score / softcap)``;
    ``dropout_p`` drops attention probabilities (the caller passes 0 outside training) with a mask that is a pure
    function of ``(seed, head, query, key)``, regenerated by the backward kernel — the arguments of the reference's
    dispatcher (attention_op.py:545-551).  ``deterministic`` (default: ``deterministic_default()``) makes the backward
    bit-reproducible: the kv blocks add their dQ contributions in ascending order instead of in arrival order (the
    forward, dK and dV are reproducible either way).  Returns ``[T, Hq*D]``.
    """
    scale = (1.0 / math.sqrt(head_dim)) if scale is None else scale
    w = int(window) if window else 0
    thr = dropout_threshold(dropout_p) if dropout_p else 0
    if thr and seed is None:
        seed = int(torch.randint(0, 2 ** 31 - 1, (1,)).item())       # host-side RNG: no device sync
    seed = int(seed or 0) & 0xffffffff
    if _lib.use_kernels(qkv):
        # the kernels count the visible keys INCLUDING the query's own position
        return _PackedAttentionFn.apply(qkv.contiguous(), seg_start, seg_end, num_heads, num_kv_heads, head_dim,
                                        scale, w + 1 if w else 0, float(softcap or 0.0), thr, seed,
                                        deterministic_default() if deterministic is None else bool(deterministic))
    t = qkv.shape[0]
    q, k, v = qkv.split([num_heads * head_dim, num_kv_heads * head_dim, num_kv_heads * head_dim], dim=-1)
    keep = dropout_keep_mask(seed, num_heads, t, thr, qkv.device) if thr else None
    o = attention_ref(q.reshape(t, num_heads, head_dim), k.reshape(t, num_kv_heads, head_dim),
                      v.reshape(t, num_kv_heads, head_dim), seg_start, scale, w or None, float(softcap or 0.0),
                      keep, thr / 256.0)
    return o.reshape(t, num_heads * head_dim)


# ---------------------------------------------------------------------------------------------------
# 4-D additive masks (eager / sdpa implementations, and the reference's public helper)
# ---------------------------------------------------------------------------------------------------
def prepare_4d_causal_attention_mask(attention_mask: torch.Tensor | None, input_shape: tuple[int, int],
                                     inputs_embeds: torch.Tensor, past_key_values_length: int = 0,
                                     sliding_window: int | None = None) -> torch.Tensor:
    """Additive ``[B, 1, S, S]`` mask (0 = attend, dtype-min = blocked): causal, block-diagonal over the
    documents of a pack (segment ids), pad keys blocked.  Vectorised: no Python loop over tokens
    (reference loops over every token, attention_op.py:286-301)."""
    assert past_key_values_length == 0, 'training-only framework: no KV cache'
    bsz, seq = input_shape
    device, dtype = inputs_embeds.device, inputs_embeds.dtype
    start, _ = segment_bounds(attention_mask, bsz, seq, device)
    start = start.view(bsz, seq) - (torch.arange(bsz, device=device, dtype=torch.int32) * seq)[:, None]
    idx = torch.arange(seq, device=device)
    allowed = (idx[None, None, :] <= idx[None, :, None]) & (idx[None, None, :] >= start[:, :, None])
    if sliding_window:
        allowed &= idx[None, None, :] >= (idx[None, :, None] - sliding_window)
    if attention_mask is not None:
        allowed &= attention_mask.to(device).bool()[:, None, :]
        # fully masked (pad) query rows attend to themselves so softmax stays finite
        allowed |= torch.eye(seq, dtype=torch.bool, device=device)[None] & ~attention_mask.to(device).bool()[:, :, None]
    mask = torch.zeros((bsz, 1, seq, seq), dtype=dtype, device=device)
    return mask.masked_fill(~allowed[:, None], torch.finfo(dtype).min)


def get_unpad_data(attention_mask: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor, int]:
    """(indices of non-pad tokens, cu_seqlens int32 over documents, max document length).
    API parity with the reference's ``_get_unpad_data`` (attention_op.py:388-412); the training path
    does not use it (it would force a host sync)."""
    bsz, seq = attention_mask.shape
    start, end = segment_bounds(attention_mask, bsz, seq, attention_mask.device)
    flat = attention_mask.reshape(-1)
    keep = flat != 0
    indices = torch.nonzero(keep, as_tuple=False).flatten()
    idx = torch.arange(bsz * seq, device=attention_mask.device, dtype=torch.int32)
    is_doc_start = (start == idx) & keep
    lengths = (end - start)[is_doc_start]
    cu = torch.nn.functional.pad(torch.cumsum(lengths, 0, dtype=torch.int32), (1, 0))
    return indices, cu, int(lengths.max().item()) if lengths.numel() else 0


# ---------------------------------------------------------------------------------------------------
# helpers of the reference's public surface (attention_op.py:286-314, 375-385, 488-536), vectorised
# ---------------------------------------------------------------------------------------------------
def get_sequence_indices(attention_mask: torch.Tensor) -> list[tuple[int, int, int]]:
    """``(row, first, last)`` (inclusive) of every run of equal segment ids, row by row, in order of appearance.
    One pass of tensor ops instead of the reference's Python loop over every token."""
    mask = attention_mask if attention_mask.dim() == 2 else attention_mask.view(1, -1)
    rows, seq = mask.shape
    if seq == 0:
        return []
    change = torch.ones((rows, seq), dtype=torch.bool, device=mask.device)
    change[:, 1:] = mask[:, 1:] != mask[:, :-1]
    starts = change.nonzero()                                    # sorted by (row, column)
    r, first = starts[:, 0], starts[:, 1]
    last = torch.empty_like(first)
    last[:-1] = first[1:] - 1
    last[-1] = seq - 1
    row_end = torch.ones_like(r, dtype=torch.bool)
    row_end[:-1] = r[1:] != r[:-1]
    last[row_end] = seq - 1
    return [(int(a), int(b), int(c)) for a, b, c in zip(r.tolist(), first.tolist(), last.tolist())]


def update_4d_causal_mask_for_packed_sequence(causal_mask: torch.Tensor,
                                              sequence_indices: list[tuple[int, int, int]],
                                              inplace: bool = False) -> torch.Tensor:
    """Block everything a document's queries could see before the document's first token."""
    causal_mask = causal_mask if inplace else causal_mask.clone()
    min_dtype = torch.finfo(causal_mask.dtype).min
    for i, s, e in sequence_indices:
        causal_mask[i, :, s:e + 1, :s] = min_dtype
    return causal_mask


def get_max_seqlen_in_batch(attention_mask: torch.Tensor) -> torch.Tensor:
    """Lengths (int32) of the documents of a segment-id mask, ordered by row then by segment id; padding (id 0) is
    not a document.  Same result as the reference's loop over ``max(segment id)`` without the per-id reductions."""
    mask = attention_mask if attention_mask.dim() == 2 else attention_mask.view(1, -1)
    n_ids = int(mask.max()) if mask.numel() else 0
    if n_ids <= 0:
        return torch.zeros(0, dtype=torch.int32, device=mask.device)
    ids = torch.arange(1, n_ids + 1, device=mask.device)
    counts = (mask.unsqueeze(-1) == ids).sum(dim=1).flatten()    # [rows, n_ids] -> row-major
    return counts[counts > 0].to(torch.int32)


def prepare_fa2_from_position_ids(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                  position_ids: torch.Tensor):
    """Flatten ``[B, S, H, D]`` q / k / v and derive ``cu_seqlens`` from the places where ``position_ids`` restart at
    0 (every document of a pack restarts its positions).  Returns ``(q, k, v, indices_q, (cu_q, cu_k),
    (max_q, max_k))`` like the reference, for callers that drive a varlen attention kernel themselves."""
    query = query.reshape(-1, query.size(-2), query.size(-1))
    key = key.reshape(-1, key.size(-2), key.size(-1))
    value = value.reshape(-1, value.size(-2), value.size(-1))
    flat = position_ids.flatten()
    indices_q = torch.arange(flat.size(0), device=flat.device, dtype=torch.int32)
    cu = torch.cat((indices_q[flat == 0], torch.tensor([flat.size(0)], device=flat.device, dtype=torch.int32)))
    max_len = flat.max() + 1
    return query, key, value, indices_q, (cu, cu), (max_len, max_len)


def flash_attention_forward(query_states: torch.Tensor, key_states: torch.Tensor, value_states: torch.Tensor,
                            attention_mask: torch.Tensor | None, query_length: int, is_causal: bool = True,
                            dropout: float = 0.0, softmax_scale: float | None = None,
                            sliding_window: int | None = None, softcap: float | None = None,
                            deterministic: bool | None = None, **_) -> torch.Tensor:
    """Reference-compatible entry point: q/k/v ``[B, S, H, D]``, ``attention_mask [B, S]`` of segment ids.
    Routes to the packed kernel; segment ids are always honoured (fixes SURVEY Appendix B2)."""
    assert is_causal, 'only causal attention is implemented'
    bsz, seq, hq, d = query_states.shape
    hkv = key_states.shape[2]
    start, end = segment_bounds(attention_mask, bsz, seq, query_states.device)
    qkv = torch.cat((query_states.reshape(bsz * seq, hq * d), key_states.reshape(bsz * seq, hkv * d),
                     value_states.reshape(bsz * seq, hkv * d)), dim=-1)
    out = packed_attention(qkv, start, end, hq, hkv, d, softmax_scale, sliding_window, dropout_p=dropout,
                           softcap=softcap or 0.0, deterministic=deterministic)
    return out.view(bsz, seq, hq, d)
