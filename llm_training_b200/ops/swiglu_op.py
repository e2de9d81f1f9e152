"""SiLU(gate) * up, and the three-matrix SwiGLU MLP
(reference: src/llm_training/ops/swiglu_op.py:5-29, ops/liger_kernel/swiglu_op.py:6-39)."""
import torch

from . import _lib
from .reference import silu_mul_ref


def silu_mul_fwd_raw(gu: torch.Tensor) -> torch.Tensor:
    """Kernel launch only: gu contiguous bf16 [rows, 2F] (gate | up) -> [rows, F]."""
    rows, f2 = gu.shape
    f = f2 // 2
    out = torch.empty((rows, f), dtype=gu.dtype, device=gu.device)
    _lib.check(_lib.lib().llmt_silu_mul_fwd(_lib.ptr(gu), _lib.ctypes.c_void_p(gu.data_ptr() + 2 * f), _lib.ptr(out),
                                            _lib.c_int(rows), _lib.c_int(f), _lib.c_i64(f2), _lib.c_i64(f),
                                            _lib.stream()), 'silu_mul_fwd')
    return out


def silu_mul_bwd_raw(d: torch.Tensor, gu: torch.Tensor) -> torch.Tensor:
    """d contiguous bf16 [rows, F], gu [rows, 2F] -> d(gate | up) [rows, 2F]."""
    rows, f2 = gu.shape
    f = f2 // 2
    dgu = torch.empty_like(gu)
    c = _lib.ctypes.c_void_p
    _lib.check(_lib.lib().llmt_silu_mul_bwd(_lib.ptr(d), _lib.ptr(gu), c(gu.data_ptr() + 2 * f), _lib.ptr(dgu),
                                            c(dgu.data_ptr() + 2 * f), _lib.c_int(rows), _lib.c_int(f),
                                            _lib.c_i64(f2), _lib.c_i64(f), _lib.c_i64(f2), _lib.stream()),
               'silu_mul_bwd')
    return dgu


def silu_mul_bwd_act_raw(d: torch.Tensor, gu: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """-> (d(gate | up) [rows, 2F], silu(gate) * up [rows, F]) from ONE pass over ``gu``."""
    rows, f2 = gu.shape
    f = f2 // 2
    dgu = torch.empty_like(gu)
    act = torch.empty((rows, f), dtype=gu.dtype, device=gu.device)
    c = _lib.ctypes.c_void_p
    _lib.check(_lib.lib().llmt_silu_mul_bwd_act(_lib.ptr(d), _lib.ptr(gu), c(gu.data_ptr() + 2 * f), _lib.ptr(dgu),
                                                c(dgu.data_ptr() + 2 * f), _lib.ptr(act), _lib.c_int(rows),
                                                _lib.c_int(f), _lib.c_i64(f2), _lib.c_i64(f), _lib.c_i64(f2),
                                                _lib.stream()), 'silu_mul_bwd_act')
    return dgu, act


class _SiluMulFn(torch.autograd.Function):
    """Operates on a fused ``[rows, 2F]`` tensor (gate | up) so the GEMM before it is one launch and
    the gradient comes back as one ``[rows, 2F]`` tensor for one dgrad GEMM."""

    @staticmethod
    def forward(ctx, gate_up: torch.Tensor) -> torch.Tensor:
        shape = gate_up.shape
        f2 = shape[-1]
        f = f2 // 2
        gu = gate_up.reshape(-1, f2)
        if not gu.is_contiguous():
            gu = gu.contiguous()
        rows = gu.shape[0]
        out = torch.empty((rows, f), dtype=gu.dtype, device=gu.device)
        _lib.check(_lib.lib().llmt_silu_mul_fwd(_lib.ptr(gu), _lib.ctypes.c_void_p(gu.data_ptr() + 2 * f),
                                                _lib.ptr(out),
                                                _lib.c_int(rows), _lib.c_int(f), _lib.c_i64(f2), _lib.c_i64(f),
                                                _lib.stream()), 'silu_mul_fwd')
        ctx.save_for_backward(gu)
        ctx.shape = shape
        return out.view(*shape[:-1], f)

    @staticmethod
    def backward(ctx, dout: torch.Tensor):
        (gu,) = ctx.saved_tensors
        rows, f2 = gu.shape
        f = f2 // 2
        d = dout.reshape(rows, f)
        if not d.is_contiguous():
            d = d.contiguous()
        dgu = torch.empty_like(gu)
        c = _lib.ctypes.c_void_p
        _lib.check(_lib.lib().llmt_silu_mul_bwd(_lib.ptr(d), _lib.ptr(gu), c(gu.data_ptr() + 2 * f), _lib.ptr(dgu),
                                                c(dgu.data_ptr() + 2 * f), _lib.c_int(rows), _lib.c_int(f),
                                                _lib.c_i64(f2), _lib.c_i64(f), _lib.c_i64(f2), _lib.stream()),
                   'silu_mul_bwd')
        return dgu.view(ctx.shape)


def silu_mul_fused(gate_up: torch.Tensor) -> torch.Tensor:
    """``gate_up[..., :F]`` is the gate, ``gate_up[..., F:]`` the up projection."""
    if _lib.use_kernels(gate_up) and (gate_up.shape[-1] // 2) % 8 == 0:
        return _SiluMulFn.apply(gate_up)
    gate, up = gate_up.chunk(2, dim=-1)
    return silu_mul_ref(gate, up)


def silu_mul(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    if _lib.use_kernels(gate) and gate.shape[-1] % 8 == 0:
        return _SiluMulFn.apply(torch.cat((gate, up), dim=-1))
    return silu_mul_ref(gate, up)


def swiglu(x: torch.Tensor, *, w3: torch.Tensor, w1w2: torch.Tensor | None = None,
           w1: torch.Tensor | None = None, w2: torch.Tensor | None = None,
           b1: torch.Tensor | None = None, b2: torch.Tensor | None = None,
           b3: torch.Tensor | None = None, b1b2: torch.Tensor | None = None) -> torch.Tensor:
    """``w3(silu(w1 x) * (w2 x))`` with either a fused ``w1w2`` or separate ``w1``/``w2``."""
    from .linear_op import linear
    fused = w1w2 is not None
    assert fused != (w1 is not None and w2 is not None), 'pass either w1w2 or both w1 and w2'
    if fused:
        h = silu_mul_fused(linear(x, w1w2, b1b2))
    else:
        h = silu_mul(linear(x, w1, b1), linear(x, w2, b2))
    return linear(h, w3, b3)
