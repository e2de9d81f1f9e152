"""Block-level fusions that trade a cheap elementwise recompute for activation memory.

``norm_multi_linear``   RMSNorm followed by one or several projections of the same input (q|k|v, gate|up)
``swiglu_down_linear``  SiLU(gate)*up followed by the down projection (+ residual)

The reference saves the normalised activations and the SwiGLU product for backward (Liger ops feeding
``nn.Linear``; models/llama/llama_model.py:283,427).  Here only the GEMM *inputs that cannot be rebuilt cheaply*
are saved — the layer input ``x`` + ``rstd`` and the ``gate|up`` GEMM output — and the bandwidth-bound
RMSNorm forward kernel is re-run in backward; the SwiGLU product is rebuilt by the SiLU·mul BACKWARD kernel in the
pass that already reads ``gate|up``.  Per LLaMA-3-8B layer at 8192 tokens this drops
352 MB of saved activations (2 x 64 MB normed inputs + 224 MB SwiGLU output) for ~0.1 ms of recompute, which is
what lets the 8B model train on ONE 180 GB GPU without layer-level activation checkpointing.
"""
import torch

from . import _lib
from .gemm_op import gemm
from .linear_op import _adjacent_rows_view, _as_2d, complete_lazy_gather, linear, multi_linear, wgrad
from .rms_norm_op import rms_norm, rmsnorm_bwd_raw, rmsnorm_fwd_raw
from .swiglu_op import silu_mul_bwd_act_raw, silu_mul_bwd_raw, silu_mul_fused, silu_mul_fwd_raw


class _NormMultiLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, norm_weight: torch.Tensor, eps: float, *weights: torch.Tensor):
        x2 = _as_2d(x)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        nw = norm_weight if norm_weight.dtype == torch.bfloat16 else norm_weight.to(torch.bfloat16)
        xn, rstd = rmsnorm_fwd_raw(x2, nw, eps)
        fused = _adjacent_rows_view(weights)
        lazy = getattr(weights[0], '_llmt_lazy_unit', None)
        head = lazy[1].lazy_heads.get(lazy[2]) if lazy is not None else None
        if fused is not None and head is not None and head['pending'] and head['n'] == fused.shape[0] \
                and x2.shape[0] % 128 == 0:
            # ZeRO-3 on NVLink: this GEMM is the first consumer of the layer's q|k|v (or gate|up) rows and gathers
            # them itself
            y = lazy[0].first_gemm(xn, lazy[1], lazy[2])
        else:
            complete_lazy_gather(weights)
            if fused is None:
                fused = torch.cat([w.detach() for w in weights], dim=0)
            y = gemm(xn, fused)
        ctx.save_for_backward(x2, nw, rstd, fused)
        ctx.eps = eps
        ctx.x_shape = x.shape
        ctx.splits = [w.shape[0] for w in weights]
        ctx.weights = weights
        ctx.norm_dtype = norm_weight.dtype
        ctx.norm_param = norm_weight
        ctx.set_materialize_grads(False)
        # second output: `x` itself, for the residual connection around this block.  Its gradient then comes back to
        # THIS node together with dy, and the norm-backward kernel adds it while it writes dx (instead of autograd
        # summing the two branches of x in a separate pass over [T, d]).
        return (y if x.dim() == 2 else y.view(*x.shape[:-1], fused.shape[0])), x

    @staticmethod
    def backward(ctx, dy: torch.Tensor | None, dres: torch.Tensor | None):
        x2, nw, rstd, fused = ctx.saved_tensors
        if dy is None:                                                      # only the pass-through output was used
            return (dres, None, None, *([None] * len(ctx.splits)))
        dy2 = _as_2d(dy)
        xn, _ = rmsnorm_fwd_raw(x2, nw, ctx.eps, want_rstd=False)          # recompute (bandwidth-bound, cheap)
        dws: list = [None] * len(ctx.splits)
        if any(ctx.needs_input_grad[3:]):
            dws = wgrad(dy2, xn, ctx.weights)
        del xn
        dxn = gemm(dy2, fused, b_mn_major=True)
        r2 = None
        if dres is not None:
            r2 = dres.reshape(x2.shape)
            if not r2.is_contiguous() or r2.dtype != x2.dtype:
                r2 = r2.to(x2.dtype).contiguous()
        dx, dnw = rmsnorm_bwd_raw(dxn, x2, nw, rstd, dres=r2,
                                  weight=ctx.norm_param if ctx.needs_input_grad[1] else None)
        return (dx.view(ctx.x_shape), None if dnw is None else dnw.to(ctx.norm_dtype), None, *dws)


def norm_multi_linear(x: torch.Tensor, norm_weight: torch.Tensor, eps: float, weights: list[torch.Tensor],
                      return_residual: bool = False) -> torch.Tensor | tuple[torch.Tensor, torch.Tensor]:
    """``concat_i(rms_norm(x) @ W_i^T)``; the normalised activations are not kept for backward.

    ``return_residual=True`` also returns ``x`` for the residual connection around the block (``h = x + f(norm(x))``):
    use the returned tensor as the residual and the two gradient branches of ``x`` are summed inside the norm-backward
    kernel."""
    if (_lib.use_kernels(x) and all(w.dtype == torch.bfloat16 for w in weights) and x.shape[-1] % 8 == 0
            and x.shape[-1] <= 16384):
        y, x_pass = _NormMultiLinearFn.apply(x, norm_weight, eps, *weights)
        return (y, x_pass) if return_residual else y
    y = multi_linear(rms_norm(x, norm_weight, eps), weights)        # completes a pending lazy gather itself
    return (y, x) if return_residual else y


class _SwiGLUDownFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gate_up: torch.Tensor, w_down: torch.Tensor, residual: torch.Tensor | None):
        gu = _as_2d(gate_up)
        if not gu.is_contiguous():
            gu = gu.contiguous()
        act = silu_mul_fwd_raw(gu)
        w = _as_2d(w_down)
        r2 = _as_2d(residual) if residual is not None else None
        y = gemm(act, w, residual=r2)
        ctx.save_for_backward(gu, w)
        ctx.weights = (w_down,)
        ctx.shape = gate_up.shape
        ctx.has_residual = residual is not None
        return y if gate_up.dim() == 2 else y.view(*gate_up.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        gu, w = ctx.saved_tensors
        dy2 = _as_2d(dy)
        dw = None
        dact = gemm(dy2, w, b_mn_major=True)
        if ctx.needs_input_grad[1]:
            # one pass over gate | up gives both d(gate | up) and the recomputed SwiGLU product for the wgrad GEMM
            dgu, act = silu_mul_bwd_act_raw(dact, gu)
            del dact
            dw = wgrad(dy2, act, ctx.weights)[0]
            del act
        else:
            dgu = silu_mul_bwd_raw(dact, gu)
        return dgu.view(ctx.shape), dw, (dy if ctx.has_residual else None)


def swiglu_down_linear(gate_up: torch.Tensor, w_down: torch.Tensor, residual: torch.Tensor | None = None
                       ) -> torch.Tensor:
    """``(silu(gate) * up) @ W_down^T (+ residual)`` with ``gate_up[..., :F] | gate_up[..., F:]``."""
    f = gate_up.shape[-1] // 2
    if _lib.use_kernels(gate_up) and w_down.dtype == torch.bfloat16 and f % 8 == 0:
        return _SwiGLUDownFn.apply(gate_up, w_down, residual)
    return linear(silu_mul_fused(gate_up), w_down, residual=residual)
