"""The allocation-free RMSNorm backward entry point, called through the C ABI.

Who needs it: hosts that cannot hand the library a scratch buffer or let anything allocate on the device —
processes whose device memory is owned in full by a fixed external plan or another allocator, loops that cannot
take the implicit device synchronisation of cudaMalloc / cudaFree, and plain C callers without a tensor library.
The test drives ``llmt_rmsnorm_bwd_noalloc`` with ctypes
exactly as such a caller would (raw pointers + stream) and checks dx / dw against ``llmt_rmsnorm_bwd`` (the
entry point with a caller-provided scratch buffer), twice in a row because the built-in scratch keeps the
previous call's contents.  This file sorts last so a ``-x`` run has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _run_pair(lib, rows, cols, seed):
    torch.manual_seed(seed)
    dev = 'cuda'
    x = torch.randn(rows, cols, device=dev).bfloat16()
    w = (1 + 0.1 * torch.randn(cols, device=dev)).bfloat16()
    dy = torch.randn(rows, cols, device=dev).bfloat16()
    y = torch.empty_like(x)
    rstd = torch.empty(rows, dtype=torch.float32, device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.llmt_rmsnorm_fwd(_ptr(x), _ptr(w), _ptr(y), _ptr(rstd), ctypes.c_int(rows), ctypes.c_int(cols),
                                ctypes.c_float(1e-5), stream) == 0
    # existing entry point: scratch comes from the caller
    nparts = lib.llmt_rmsnorm_bwd_num_partials(ctypes.c_int(rows))
    scratch = torch.empty((nparts, cols), dtype=torch.float32, device=dev)
    dx_ref, dw_ref = torch.empty_like(x), torch.empty(cols, dtype=torch.float32, device=dev)
    assert lib.llmt_rmsnorm_bwd(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dx_ref), _ptr(dw_ref), _ptr(scratch),
                                ctypes.c_int(rows), ctypes.c_int(cols), ctypes.c_int(1), ctypes.c_int(0), stream) == 0
    # new entry point: inputs and outputs only
    dx, dw = torch.empty_like(x), torch.empty(cols, dtype=torch.float32, device=dev)
    rc = lib.llmt_rmsnorm_bwd_noalloc(_ptr(dy), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dx), _ptr(dw), ctypes.c_int(rows),
                                      ctypes.c_int(cols), ctypes.c_int(1), ctypes.c_int(0), stream)
    assert rc == 0, lib.llmt_last_error().decode()
    torch.cuda.synchronize()
    assert torch.equal(dx, dx_ref)                     # same kernel, same per-row math
    # both results are fp32 sums of the same per-row fp32 terms; only the grouping into partial rows differs
    # (148 vs up to 592 partials), so they agree to fp32 summation-order error — far inside the bf16-level tolerance
    # test_kernels_gpu.py::test_rmsnorm uses against the PyTorch oracle
    err = (dw - dw_ref).abs().max().item()
    bound = 1e-3 * max(1.0, dw_ref.abs().max().item())
    assert err <= bound, f'dw max abs err {err} > {bound}'


def test_rmsnorm_bwd_noalloc_matches_scratch_version_twice():
    from llm_training_b200.ops import _lib
    lib = _lib.lib()
    _run_pair(lib, rows=777, cols=4096, seed=0)
    _run_pair(lib, rows=90, cols=1024, seed=1)         # second call reuses (dirty) built-in scratch, other shape


def test_rmsnorm_bwd_noalloc_rejects_oversized_rows():
    from llm_training_b200.ops import _lib
    lib = _lib.lib()
    max_cols = lib.llmt_rmsnorm_bwd_noalloc_max_cols()
    cols = max_cols + 8
    x = torch.zeros(2, cols, device='cuda', dtype=torch.bfloat16)
    r = torch.ones(2, device='cuda')
    dw = torch.zeros(cols, device='cuda')
    rc = lib.llmt_rmsnorm_bwd_noalloc(_ptr(x), _ptr(x), _ptr(x[0]), _ptr(r), _ptr(x.clone()), _ptr(dw),
                                      ctypes.c_int(2), ctypes.c_int(cols), ctypes.c_int(1), ctypes.c_int(0),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b'exceeds the built-in scratch' in lib.llmt_last_error()
