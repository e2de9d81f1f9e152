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


def rmsnorm_bwd_raw(dy2: torch.Tensor, x2: torch.Tensor, w: torch.Tensor, rstd: torch.Tensor
                    ) -> tuple[torch.Tensor, torch.Tensor]:
    """-> (dx bf16 [rows, cols], dw fp32 [cols])."""
    rows, cols = x2.shape
    L = _lib.lib()
    dx = torch.empty_like(x2)
    nparts = L.llmt_rmsnorm_bwd_num_partials(_lib.c_int(rows))
    partial = torch.empty((nparts, cols), dtype=torch.float32, device=x2.device)
    dw = torch.empty(cols, dtype=torch.float32, device=x2.device)
    _lib.check(L.llmt_rmsnorm_bwd(_lib.ptr(dy2), _lib.ptr(x2), _lib.ptr(w), _lib.ptr(rstd), _lib.ptr(dx), _lib.ptr(dw),
                                  _lib.ptr(partial), _lib.c_int(rows), _lib.c_int(cols), _lib.c_int(1), _lib.c_int(0),
                                  _lib.stream()), 'rmsnorm_bwd')
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
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2, w, rstd = ctx.saved_tensors
        rows, cols = x2.shape
        dy2 = dy.reshape(rows, cols)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        L = _lib.lib()
        dx = torch.empty_like(x2)
        nparts = L.llmt_rmsnorm_bwd_num_partials(_lib.c_int(rows))
        partial = torch.empty((nparts, cols), dtype=torch.float32, device=x2.device)
        dw = torch.empty(cols, dtype=torch.float32, device=x2.device)
        _lib.check(L.llmt_rmsnorm_bwd(_lib.ptr(dy2), _lib.ptr(x2), _lib.ptr(w), _lib.ptr(rstd), _lib.ptr(dx),
                                      _lib.ptr(dw), _lib.ptr(partial), _lib.c_int(rows), _lib.c_int(cols),
                                      _lib.c_int(1), _lib.c_int(0), _lib.stream()), 'rmsnorm_bwd')
        return dx.view(ctx.shape), dw.to(ctx.weight_dtype), None


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    if _lib.use_kernels(x):
        return _RMSNormFn.apply(x, weight, eps)
    return rms_norm_ref(x, weight, eps)
