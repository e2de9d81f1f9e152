"""RMSNorm (reference: src/llm_training/ops/rms_norm_op.py:4-14 and the Liger wrapper
ops/liger_kernel/rms_norm_op.py:7-19)."""
import torch

from . import _lib
from .reference import rms_norm_ref


def rmsnorm_fwd_raw(x2: torch.Tensor, w: torch.Tensor, eps: float, want_rstd: bool = True
                    ) -> tuple[torch.Tensor, torch.Tensor | None]:
    """Kernel launch only (no autograd): x2 is contiguous bf16 [rows, cols], w bf16 [cols]."""
    rows, cols = x2.shape
    y = torch.empty_like(x2)
    rstd = torch.empty(rows, dtype=torch.float32, device=x2.device) if want_rstd else None
    _lib.check(_lib.lib().llmt_rmsnorm_fwd(_lib.ptr(x2), _lib.ptr(w), _lib.ptr(y), _lib.ptr(rstd), _lib.c_int(rows),
                                           _lib.c_int(cols), _lib.c_f32(eps), _lib.stream()), 'rmsnorm_fwd')
    return y, rstd


def _direct_grad_target(weight: torch.Tensor | None) -> torch.Tensor | None:
    """The weight's view inside the sharding engine's flat gradient buffer when the kernel may write its dw straight
    there (``FlatParamEngine.enable_direct_grad_accumulation``; same contract as ``ops.linear_op.wgrad``)."""
    if weight is None or not getattr(weight, '_llmt_direct_grad', False):
        return None
    g = weight.grad
    if g is None or not g.is_contiguous() or g.dtype not in (torch.bfloat16, torch.float32):
        return None
    return g


def rmsnorm_bwd_raw(dy2: torch.Tensor, x2: torch.Tensor, w: torch.Tensor, rstd: torch.Tensor,
                    dres: torch.Tensor | None = None, weight: torch.Tensor | None = None
                    ) -> tuple[torch.Tensor, torch.Tensor | None]:
    """-> (dx bf16 [rows, cols], dw fp32 [cols]).

    ``dres``: gradient of the residual connection around the norm, added to dx inside the kernel (fp32 add, one
    rounding, no separate pass).  ``weight``: the norm's parameter; when the sharding engine lets its gradient be
    written in place, dw goes (or is added) straight into ``weight.grad``, the engine is notified and the returned dw
    is ``None`` — no fp32 temporary, no dtype conversion, no AccumulateGrad launch."""
    rows, cols = x2.shape
    L = _lib.lib()
    dx = torch.empty_like(x2)
    nparts = L.llmt_rmsnorm_bwd_num_partials(_lib.c_int(rows))
    partial = torch.empty((nparts, cols), dtype=torch.float32, device=x2.device)
    target = _direct_grad_target(weight)
    if target is not None:
        fresh = getattr(weight, '_llmt_grad_fresh', False)       # first gradient of the step: overwrite, else add
        dw, is_fp32, accumulate = target, target.dtype == torch.float32, not fresh
    else:
        dw, is_fp32, accumulate = torch.empty(cols, dtype=torch.float32, device=x2.device), True, False
    _lib.check(L.llmt_rmsnorm_bwd_res(_lib.ptr(dy2), _lib.ptr(x2), _lib.ptr(w), _lib.ptr(rstd), _lib.ptr(dres),
                                      _lib.ptr(dx), _lib.ptr(dw), _lib.ptr(partial), _lib.c_int(rows),
                                      _lib.c_int(cols), _lib.c_int(int(is_fp32)), _lib.c_int(int(accumulate)),
                                      _lib.stream()), 'rmsnorm_bwd')
    if target is not None:
        weight._llmt_grad_fresh = False
        weight._llmt_grad_ready(weight)
        return dx, None
    return dx, dw


class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
        shape = x.shape
        cols = shape[-1]
        x2 = x.reshape(-1, cols)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        w = weight if weight.dtype == torch.bfloat16 else weight.to(torch.bfloat16)
        y = torch.empty_like(x2)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        _lib.check(_lib.lib().llmt_rmsnorm_fwd(_lib.ptr(x2), _lib.ptr(w), _lib.ptr(y), _lib.ptr(rstd),
                                               _lib.c_int(rows), _lib.c_int(cols), _lib.c_f32(eps),
                                               _lib.stream()), 'rmsnorm_fwd')
        ctx.save_for_backward(x2, w, rstd)
        ctx.shape = shape
        ctx.weight_dtype = weight.dtype
        ctx.weight = weight
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2, w, rstd = ctx.saved_tensors
        rows, cols = x2.shape
        dy2 = dy.reshape(rows, cols)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx, dw = rmsnorm_bwd_raw(dy2, x2, w, rstd, weight=ctx.weight if ctx.needs_input_grad[1] else None)
        return dx.view(ctx.shape), (None if dw is None else dw.to(ctx.weight_dtype)), None


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    if _lib.use_kernels(x):
        return _RMSNormFn.apply(x, weight, eps)
    return rms_norm_ref(x, weight, eps)
