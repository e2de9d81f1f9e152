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
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x2, w = ctx.saved_tensors
        dy2 = _as_2d(dy)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, b_mn_major=True).view(ctx.x_shape)        # [T,out] @ [out,in]
        if ctx.needs_input_grad[1]:
            dw = gemm(dy2, x2, a_mn_major=True, b_mn_major=True)        # [T,out]^T @ [T,in]
        dres = dy if ctx.has_residual else None
        return dx, dw, dres


def linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None,
           residual: torch.Tensor | None = None) -> torch.Tensor:
    if _lib.use_kernels(x) and weight.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and weight.shape[0] % 8 == 0:
        y = _LinearFn.apply(x, weight, residual)
        if bias is not None:
            y = y + bias
        return y
    y = F.linear(x, weight, bias)
    return y if residual is None else y + residual
