"""``y = x W^T (+ b) (+ residual)`` — every projection of the models goes through here.

CUDA/bf16: forward, dgrad and wgrad are three launches of the tcgen05 GEMM
(``csrc/kernels/gemm.cu``), with the residual add fused into the forward epilogue and the
weight-gradient accumulation (``main_grad += dY^T X``) fused into the wgrad epilogue.
Replaces the cuBLAS calls behind ``nn.Linear`` in the reference.
"""
import torch
import torch.nn.functional as F

from . import _lib
from .gemm_op import gemm


def _as_2d(t: torch.Tensor) -> torch.Tensor:
    t2 = t.reshape(-1, t.shape[-1])
    if t2.stride(-1) != 1 or (t2.shape[0] > 1 and t2.stride(0) % 8 != 0) or t2.data_ptr() % 16 != 0:
        t2 = t2.contiguous()
    return t2


def complete_lazy_gather(weights) -> None:
    """ZeRO-3 on NVLink leaves a layer's q|k|v rows to be gathered by the GEMM that consumes them
    (``FlatParamEngine.first_gemm``); any other consumer calls this first."""
    for w in weights:
        lazy = getattr(w, '_llmt_lazy_unit', None)
        if lazy is not None and lazy[1].lazy_heads[lazy[2]]['pending']:
            lazy[0].complete_head(lazy[1], lazy[2])


def wgrad(dy2: torch.Tensor, x2: torch.Tensor, weights: tuple[torch.Tensor, ...]) -> list:
    """dW_i = dY_i^T X for row-stacked weights.  When the sharding engine has given every weight a gradient view
    inside its flat buffer and marked it ``_llmt_direct_grad`` (weights used once per step), the GEMM epilogue
    accumulates straight into that buffer (C += A^T B) and the engine is notified — no temporary dW, no
    AccumulateGrad pass.  Otherwise returns the gradients for autograd to accumulate."""
    direct = all(getattr(w, '_llmt_direct_grad', False) and w.grad is not None and w.grad.dtype == dy2.dtype
                 for w in weights)
    if direct:
        target = weights[0].grad if len(weights) == 1 else _adjacent_rows_view(tuple(w.grad for w in weights))
        if target is not None and target.is_contiguous():
            # first gradient of the step for these weights: the engine skipped zeroing their range of the flat buffer
            # (`_llmt_grad_fresh`), so the GEMM overwrites instead of accumulating — one pass over dW less on each side
            fresh = all(getattr(w, '_llmt_grad_fresh', False) for w in weights)
            if not fresh:
                for w in weights:
                    if getattr(w, '_llmt_grad_fresh', False):   # mixed group: make the stale ranges well defined
                        w.grad.zero_()
            gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=target, residual=None if fresh else target)
            for w in weights:
                w._llmt_grad_fresh = False
                w._llmt_grad_ready(w)
            return [None] * len(weights)
    dw = gemm(dy2, x2, a_mn_major=True, b_mn_major=True)
    return [dw] if len(weights) == 1 else list(dw.split([w.shape[0] for w in weights], dim=0))


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, residual: torch.Tensor | None) -> torch.Tensor:
        x2 = _as_2d(x)
        w = _as_2d(weight)
        r2 = _as_2d(residual) if residual is not None else None
        y = gemm(x2, w, residual=r2)
        ctx.save_for_backward(x2, w)
        ctx.x_shape = x.shape
        ctx.has_residual = residual is not None
        ctx.weights = (weight,)
        return y if x.dim() == 2 else y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2, w = ctx.saved_tensors
        dy2 = _as_2d(dy)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, b_mn_major=True).view(ctx.x_shape)        # [T,out] @ [out,in]
        if ctx.needs_input_grad[1]:
            dw = wgrad(dy2, x2, ctx.weights)[0]                         # [T,out]^T @ [T,in]
        dres = dy if ctx.has_residual else None
        return dx, dw, dres


def linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None,
           residual: torch.Tensor | None = None) -> torch.Tensor:
    complete_lazy_gather((weight,))
    if _lib.use_kernels(x) and weight.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and weight.shape[0] % 8 == 0:
        y = _LinearFn.apply(x, weight, residual)
        if bias is not None:
            y = y + bias
        return y
    y = F.linear(x, weight, bias)
    return y if residual is None else y + residual


# ---------------------------------------------------------------------------------------------------
# several projections of the same input as ONE GEMM (q|k|v, gate|up)
# ---------------------------------------------------------------------------------------------------
def _adjacent_rows_view(weights: tuple[torch.Tensor, ...]) -> torch.Tensor | None:
    """If the weights are consecutive row blocks of one buffer (the flat parameter layout built by
    ``parallel.ParamStore``), return the fused ``[sum(out_i), in]`` view without copying."""
    first = weights[0]
    if not all(w.is_contiguous() and w.dim() == 2 and w.shape[1] == first.shape[1] and w.dtype == first.dtype
               for w in weights):
        return None
    expected = first.data_ptr()
    for w in weights:
        if w.data_ptr() != expected or w.untyped_storage().data_ptr() != first.untyped_storage().data_ptr():
            return None
        expected += w.numel() * w.element_size()
    rows = sum(w.shape[0] for w in weights)
    return torch.as_strided(first.detach(), (rows, first.shape[1]), (first.shape[1], 1))


class _MultiLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, *weights: torch.Tensor) -> torch.Tensor:
        fused = _adjacent_rows_view(weights)
        if fused is None:
            fused = torch.cat([w.detach() for w in weights], dim=0)
        x2 = _as_2d(x)
        y = gemm(x2, fused)
        ctx.save_for_backward(x2, fused)
        ctx.x_shape = x.shape
        ctx.splits = [w.shape[0] for w in weights]
        ctx.weights = weights
        # no view for the (usual) 2-D packed case: RoPE rotates this tensor in place
        return y if x.dim() == 2 else y.view(*x.shape[:-1], fused.shape[0])

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2, fused = ctx.saved_tensors
        dy2 = _as_2d(dy)
        dx = gemm(dy2, fused, b_mn_major=True).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dws: list = [None] * len(ctx.splits)
        if any(ctx.needs_input_grad[1:]):
            dws = wgrad(dy2, x2, ctx.weights)
        return (dx, *dws)


def multi_linear(x: torch.Tensor, weights: list[torch.Tensor], biases: list[torch.Tensor | None] | None = None
                 ) -> torch.Tensor:
    """``concat_i(x @ W_i^T (+ b_i))`` along the last dim, computed as a single GEMM on CUDA."""
    complete_lazy_gather(weights)
    if _lib.use_kernels(x) and all(w.dtype == torch.bfloat16 for w in weights) and x.shape[-1] % 8 == 0:
        y = _MultiLinearFn.apply(x, *weights)
    else:
        y = F.linear(x, torch.cat(list(weights), dim=0))
    if biases is not None and any(b is not None for b in biases):
        full = torch.cat([b if b is not None else torch.zeros(w.shape[0], dtype=y.dtype, device=y.device)
                          for w, b in zip(weights, biases)])
        y = y + full
    return y
