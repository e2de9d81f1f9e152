"""Python face of the tcgen05 GEMM (``csrc/kernels/gemm.cu``)."""
import os

import torch

from . import _lib

# CTA pairs (cta_group::2, 256x256 tiles) are the fast configuration; LLMT_GEMM_CTA_GROUP=1 selects the
# single-CTA form (debugging / A-B measurements).
_CTA_GROUP = int(os.environ.get('LLMT_GEMM_CTA_GROUP', '2'))


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, 'row-major 2-D view required'
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
         out: torch.Tensor | None = None, residual: torch.Tensor | None = None,
         out_dtype: torch.dtype = torch.bfloat16, cta_group: int | None = None,
         max_ctas: int = 0) -> torch.Tensor:
    """``C[M,N] = op(A) @ op(B) (+ residual)`` on bf16 operands with fp32 accumulation.

    ``a`` is ``[M, K]`` (or ``[K, M]`` if ``a_mn_major``); ``b`` is ``[N, K]`` (or ``[K, N]`` if
    ``b_mn_major``).  ``residual is out`` accumulates in place (``C += A @ B``).
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    m, k = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    n, kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    assert k == kb, f'K mismatch: {a.shape} vs {b.shape}'
    if out is None:
        out = torch.empty((m, n), dtype=out_dtype, device=a.device)
    else:
        assert out.shape == (m, n) and out.dtype in (torch.bfloat16, torch.float32)
    if residual is not None:
        assert residual.shape == (m, n) and residual.dtype == out.dtype
    cg = _CTA_GROUP if cta_group is None else cta_group
    rc = _lib.lib().llmt_gemm_bf16(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _lib.ptr(residual), _lib.c_int(m), _lib.c_int(n), _lib.c_int(k),
        _lib.c_i64(_ld(a)), _lib.c_i64(_ld(b)), _lib.c_i64(_ld(out)),
        _lib.c_i64(_ld(residual) if residual is not None else 0),
        _lib.c_int(a_mn_major), _lib.c_int(b_mn_major), _lib.c_int(out.dtype == torch.float32),
        _lib.c_int(cg), _lib.c_int(max_ctas), _lib.stream())
    _lib.check(rc, 'gemm_bf16')
    return out
