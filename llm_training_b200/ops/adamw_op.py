"""Fused AdamW step + squared-norm reduction on flat buffers (``csrc/kernels/optimizer.cu``)."""
import torch

from . import _lib
from .reference import adamw_ref


def sumsq(x: torch.Tensor, out: torch.Tensor, accumulate: bool = False) -> torch.Tensor:
    """out[0] (+)= sum(x^2) on device, no host sync. ``out`` is an fp32 tensor with >= 1 element."""
    if x.is_cuda and x.dtype in (torch.bfloat16, torch.float32) and x.is_contiguous():
        L = _lib.lib()
        partial = torch.empty(L.llmt_sumsq_num_partials(), dtype=torch.float32, device=x.device)
        _lib.check(L.llmt_sumsq(_lib.ptr(x), _lib.c_i64(x.numel()), _lib.c_int(x.dtype == torch.float32),
                                _lib.ptr(partial), _lib.ptr(out), _lib.c_int(accumulate), _lib.stream()), 'sumsq')
        return out
    s = x.float().pow(2).sum()
    if accumulate:
        out[0] += s
    else:
        out[0] = s
    return out


def adamw_step(master: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, grad: torch.Tensor,
               param: torch.Tensor | None, *, lr: float, beta1: float, beta2: float, eps: float,
               weight_decay: float, step: int, grad_scale: float = 1.0, max_norm: float = 0.0,
               norm_sq: torch.Tensor | None = None) -> None:
    """One AdamW update of a flat fp32 ``master`` shard; ``param`` (bf16, optional) receives the
    rounded copy.  With ``max_norm > 0`` the gradient is scaled by min(1, max_norm / (sqrt(norm_sq)+1e-6))
    where ``norm_sq`` is a device scalar — clipping without a host round trip."""
    if master.is_cuda:
        _lib.check(_lib.lib().llmt_adamw(
            _lib.ptr(master), _lib.ptr(exp_avg), _lib.ptr(exp_avg_sq), _lib.ptr(grad), _lib.ptr(param),
            _lib.c_i64(master.numel()), _lib.c_int(grad.dtype == torch.float32), _lib.c_f32(lr), _lib.c_f32(beta1),
            _lib.c_f32(beta2), _lib.c_f32(eps), _lib.c_f32(weight_decay), _lib.c_int(step), _lib.c_f32(grad_scale),
            _lib.c_f32(max_norm), _lib.ptr(norm_sq), _lib.stream()), 'adamw')
        return
    scale = grad_scale
    if max_norm > 0 and norm_sq is not None:
        norm = float(norm_sq.sqrt()) * abs(grad_scale)
        scale *= min(1.0, max_norm / (norm + 1e-6))
    adamw_ref(master, grad, exp_avg, exp_avg_sq, lr=lr, beta1=beta1, beta2=beta2, eps=eps,
              weight_decay=weight_decay, step=step, grad_scale=scale)
    if param is not None:
        param.copy_(master)
