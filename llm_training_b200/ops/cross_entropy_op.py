"""Cross-entropy family: ``shift_labels``, ``cross_entropy``, ``fused_linear_cross_entropy`` and the
per-token log-prob used by DPO / ORPO.

Reference: src/llm_training/ops/cross_entropy_op.py:4-8 (shift_labels),
ops/liger_kernel/cross_entropy_op.py:10-54 (CE / fused linear-CE wrappers),
lms/dpo/dpo.py:73-114 and lms/orpo/orpo.py:61-93 (log-prob gather).

CUDA path: logits are produced in bf16 by the tcgen05 GEMM in row chunks, reduced by
``csrc/kernels/cross_entropy.cu`` and turned into dlogits IN PLACE, so neither fp32 logits
(reference: ``logits.float()``, lms/clm/clm.py:147) nor a [T, V] log-softmax tensor ever exist.
Up to ``_CHUNK_ROWS`` tokens (the usual micro-batch) the bf16 logits chunk is simply kept until backward: the
upstream gradient is folded into the per-row scale of the in-place dlogits kernel, dH = dlogits W comes from one GEMM
and dW = dlogits^T H goes through ``ops.linear_op.wgrad``, i.e. straight into the sharding engine's flat gradient
buffer when it allows that — no dW temporary, no scaling pass, no AccumulateGrad pass over [V, d].
Longer inputs run chunk by chunk with the backward GEMMs inside the forward call (the gradient of a scalar loss
w.r.t. the logits does not depend on the upstream gradient except for a scale), which lets one chunk buffer be
recycled; the saved dH / dW are scaled in backward.
"""
import torch
import torch.nn.functional as F

from . import _lib
from .gemm_op import gemm
from .linear_op import wgrad
from .reference import cross_entropy_ref, token_logprob_ref


def shift_labels(labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """labels[t] <- labels[t+1]; last position of every row <- ignore_index."""
    shifted = torch.empty_like(labels)
    shifted[..., :-1] = labels[..., 1:]
    shifted[..., -1] = ignore_index
    return shifted


# ---------------------------------------------------------------------------------------------------
# plain CE on materialised logits (API parity; the models use the fused form below)
# ---------------------------------------------------------------------------------------------------
class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, labels: torch.Tensor, ignore_index: int) -> torch.Tensor:
        L = _lib.lib()
        rows, vocab = logits.shape
        lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
        tgt = torch.empty_like(lse)
        out2 = torch.empty(2, dtype=torch.float32, device=logits.device)
        _lib.check(L.llmt_ce_fwd(_lib.ptr(logits), _lib.ptr(labels), _lib.ptr(lse), _lib.ptr(tgt), _lib.c_int(rows),
                                 _lib.c_int(vocab), _lib.c_i64(logits.stride(0)), _lib.c_i64(ignore_index),
                                 _lib.c_f32(0.0), _lib.stream()), 'ce_fwd')
        _lib.check(L.llmt_ce_reduce(_lib.ptr(lse), _lib.ptr(tgt), _lib.ptr(labels), _lib.ptr(out2), _lib.c_int(rows),
                                    _lib.c_i64(ignore_index), _lib.stream()), 'ce_reduce')
        ctx.save_for_backward(logits, labels, lse, out2)
        ctx.ignore_index = ignore_index
        return out2[0] / out2[1].clamp_min(1.0)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        logits, labels, lse, out2 = ctx.saved_tensors
        rows, vocab = logits.shape
        dlogits = logits.clone()
        scale = (grad_out.float() / out2[1].clamp_min(1.0)).expand(rows).contiguous()
        _lib.check(_lib.lib().llmt_ce_bwd_inplace(_lib.ptr(dlogits), _lib.ptr(labels), _lib.ptr(lse), _lib.ptr(scale),
                                                  _lib.c_int(rows), _lib.c_int(vocab), _lib.c_i64(dlogits.stride(0)),
                                                  _lib.c_i64(ctx.ignore_index), _lib.c_f32(0.0), _lib.stream()),
                   'ce_bwd')
        return dlogits, None, None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100,
                  reduction: str = 'mean') -> torch.Tensor:
    assert reduction == 'mean'
    if logits.dim() == 3:
        logits = logits.flatten(end_dim=1)
        labels = labels.flatten(end_dim=1)
    if _lib.use_kernels(logits) and logits.stride(-1) == 1 and logits.stride(0) % 8 == 0:
        return _CrossEntropyFn.apply(logits, labels.contiguous(), ignore_index)
    return cross_entropy_ref(logits, labels, ignore_index)


# ---------------------------------------------------------------------------------------------------
# fused linear + CE / log-prob
# ---------------------------------------------------------------------------------------------------
_CHUNK_ROWS = 8192      # 8192 x 128256 bf16 logits = 2.1 GB; one chunk for the usual 8k-token micro-batch


def _chunked_logits_pass(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int,
                         row_scale_fn, need_grads: bool):
    """Shared driver.  For each row chunk: logits = H W^T (bf16) -> lse/tgt -> (optional) in-place
    dlogits -> dH chunk, dW accumulation.  ``row_scale_fn(lse, tgt, labels_chunk)`` returns the fp32
    per-row factor applied to (softmax - onehot)."""
    L = _lib.lib()
    rows, dim = hidden.shape
    vocab = weight.shape[0]
    dev = hidden.device
    lse = torch.empty(rows, dtype=torch.float32, device=dev)
    tgt = torch.empty_like(lse)
    chunk = min(_CHUNK_ROWS, rows)
    vpad = (vocab + 7) // 8 * 8
    buf = torch.empty((chunk, vpad), dtype=torch.bfloat16, device=dev)
    dh = torch.empty_like(hidden) if need_grads else None
    single = rows <= chunk
    # several chunks accumulate dW in fp32; a single chunk writes the bf16 gradient directly
    dw = None
    if need_grads:
        dw = torch.empty((vocab, dim), dtype=weight.dtype, device=dev) if single \
            else torch.zeros((vocab, dim), dtype=torch.float32, device=dev)
    for r0 in range(0, rows, chunk):
        r1 = min(rows, r0 + chunk)
        n = r1 - r0
        logits = buf[:n, :vocab]
        gemm(hidden[r0:r1], weight, out=logits)
        lab = labels[r0:r1]
        _lib.check(L.llmt_ce_fwd(_lib.ptr(logits), _lib.ptr(lab), _lib.ptr(lse[r0:r1]), _lib.ptr(tgt[r0:r1]),
                                 _lib.c_int(n), _lib.c_int(vocab), _lib.c_i64(vpad), _lib.c_i64(ignore_index),
                                 _lib.c_f32(0.0), _lib.stream()), 'ce_fwd')
        if need_grads:
            scale = row_scale_fn(r0, r1)
            _lib.check(L.llmt_ce_bwd_inplace(_lib.ptr(logits), _lib.ptr(lab), _lib.ptr(lse[r0:r1]), _lib.ptr(scale),
                                             _lib.c_int(n), _lib.c_int(vocab), _lib.c_i64(vpad),
                                             _lib.c_i64(ignore_index), _lib.c_f32(0.0), _lib.stream()), 'ce_bwd')
            gemm(logits, weight, b_mn_major=True, out=dh[r0:r1])                      # dH = dlogits @ W
            gemm(logits, hidden[r0:r1], a_mn_major=True, b_mn_major=True, out=dw,     # dW (+)= dlogits^T @ H
                 residual=None if single else dw)
    return lse, tgt, dh, dw


class _FusedLinearCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int):
        need_grads = hidden.requires_grad or weight.requires_grad
        rows = hidden.shape[0]
        ctx.deferred = rows <= _CHUNK_ROWS
        if ctx.deferred:
            L = _lib.lib()
            vocab = weight.shape[0]
            vpad = (vocab + 7) // 8 * 8
            buf = torch.empty((rows, vpad), dtype=torch.bfloat16, device=hidden.device)
            logits = buf[:, :vocab]
            gemm(hidden, weight, out=logits)
            lse = torch.empty(rows, dtype=torch.float32, device=hidden.device)
            tgt = torch.empty_like(lse)
            out2 = torch.empty(2, dtype=torch.float32, device=hidden.device)      # [sum of token losses, valid tokens]
            _lib.check(L.llmt_ce_fwd(_lib.ptr(logits), _lib.ptr(labels), _lib.ptr(lse), _lib.ptr(tgt),
                                     _lib.c_int(rows), _lib.c_int(vocab), _lib.c_i64(vpad), _lib.c_i64(ignore_index),
                                     _lib.c_f32(0.0), _lib.stream()), 'ce_fwd')
            _lib.check(L.llmt_ce_reduce(_lib.ptr(lse), _lib.ptr(tgt), _lib.ptr(labels), _lib.ptr(out2),
                                        _lib.c_int(rows), _lib.c_i64(ignore_index), _lib.stream()), 'ce_reduce')
            if need_grads:
                ctx.save_for_backward(hidden, weight, labels, lse, buf, out2)
                ctx.weights = (weight,)
                ctx.ignore_index = ignore_index
                ctx.consumed = False
            return out2[0] / out2[1].clamp_min(1.0)
        n_valid = labels.ne(ignore_index).sum().clamp_min(1).float()
        inv = (1.0 / n_valid)

        def row_scale(r0, r1):
            return inv.expand(r1 - r0).contiguous()

        lse, tgt, dh, dw = _chunked_logits_pass(hidden, weight, labels, ignore_index, row_scale, need_grads)
        valid = labels.ne(ignore_index)
        loss = ((lse - tgt) * valid).sum() * inv
        if need_grads:
            ctx.save_for_backward(dh, dw)
        ctx.weight_dtype = weight.dtype
        return loss

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        if ctx.deferred:
            hidden, weight, labels, lse, buf, out2 = ctx.saved_tensors
            if ctx.consumed:
                raise RuntimeError('fused_linear_cross_entropy: the logits were turned into their gradient in place by '
                                   'the first backward pass; a second backward through the same loss is not supported')
            ctx.consumed = True
            rows, vocab = hidden.shape[0], weight.shape[0]
            logits = buf[:, :vocab]
            scale = (grad_out.float() / out2[1].clamp_min(1.0)).expand(rows).contiguous()
            _lib.check(_lib.lib().llmt_ce_bwd_inplace(_lib.ptr(logits), _lib.ptr(labels), _lib.ptr(lse),
                                                      _lib.ptr(scale), _lib.c_int(rows), _lib.c_int(vocab),
                                                      _lib.c_i64(buf.shape[1]), _lib.c_i64(ctx.ignore_index),
                                                      _lib.c_f32(0.0), _lib.stream()), 'ce_bwd')
            dh = gemm(logits, weight, b_mn_major=True) if ctx.needs_input_grad[0] else None     # dlogits @ W
            dw = wgrad(logits, hidden, ctx.weights)[0] if ctx.needs_input_grad[1] else None     # dlogits^T @ H
            return dh, dw, None, None
        dh, dw = ctx.saved_tensors
        g = grad_out.float()
        return (dh * g.to(dh.dtype)), (dw * g.to(dw.dtype)).to(ctx.weight_dtype), None, None


def fused_linear_cross_entropy(hidden_states: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                               ignore_index: int = -100, reduction: str = 'mean') -> torch.Tensor:
    """mean CE of ``hidden_states @ weight.T`` against ``labels`` without materialising [T, V] logits."""
    assert reduction == 'mean'
    if hidden_states.dim() == 3:
        hidden_states = hidden_states.flatten(end_dim=1)
        labels = labels.flatten(end_dim=1)
    if _lib.use_kernels(hidden_states) and weight.dtype == torch.bfloat16:
        return _FusedLinearCEFn.apply(hidden_states.contiguous(), weight.contiguous(), labels.contiguous(),
                                      ignore_index)
    logits = F.linear(hidden_states, weight)
    return cross_entropy_ref(logits, labels, ignore_index)


class _FusedLinearLogProbFn(torch.autograd.Function):
    """Per-token log p(label).  Backward needs the upstream per-token gradient, so the logits are
    recomputed chunk by chunk in backward (one extra lm_head GEMM instead of a saved [T, V] tensor)."""

    @staticmethod
    def forward(ctx, hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int):
        lse, tgt, _, _ = _chunked_logits_pass(hidden, weight, labels, ignore_index, None, False)
        ctx.save_for_backward(hidden, weight, labels)
        ctx.ignore_index = ignore_index
        return (tgt - lse) * labels.ne(ignore_index)

    @staticmethod
    def backward(ctx, dlogp: torch.Tensor):
        hidden, weight, labels = ctx.saved_tensors
        neg = (-dlogp.float()).contiguous()     # d/dlogits of logp = onehot - softmax = -(softmax - onehot)

        def row_scale(r0, r1):
            return neg[r0:r1].contiguous()

        _, _, dh, dw = _chunked_logits_pass(hidden, weight, labels, ctx.ignore_index, row_scale, True)
        return dh, dw.to(weight.dtype), None, None


def fused_linear_logprob(hidden_states: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                         ignore_index: int = -100) -> torch.Tensor:
    """``log_softmax(hidden @ weight.T)[labels]`` per token (0 at ignored positions); same shape as labels."""
    shape = labels.shape
    h2 = hidden_states.reshape(-1, hidden_states.shape[-1])
    l2 = labels.reshape(-1)
    if _lib.use_kernels(h2) and weight.dtype == torch.bfloat16:
        out = _FusedLinearLogProbFn.apply(h2.contiguous(), weight.contiguous(), l2.contiguous(), ignore_index)
    else:
        out = token_logprob_ref(h2, weight, l2, ignore_index)
    return out.view(shape)
