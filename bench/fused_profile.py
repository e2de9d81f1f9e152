"""The three fused compute+collective GEMMs at LLaMA-3-8B shapes with TWO ranks emulated on one GPU (the peer's
buffers are ordinary buffers of the same device), for single-GPU ncu captures: ncu cannot profile a multi-rank run.
The pull warps therefore read local HBM instead of NVLink; pipeline behaviour of the kernel is what is captured,
multi-GPU timings are in profiles/tp_fused_bench_*.json."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_training_b200.ops import _lib  # noqa: E402

DEV = 'cuda'
T, d, ffn, world = 8192, 4096, 14336, 2
L = _lib.lib()


def arr(ts):
    return (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


g = torch.Generator(device=DEV).manual_seed(0)
flags = torch.zeros(256, dtype=torch.int32, device=DEV)
# all-gather -> GEMM (qkv, tp=2)
blocks = [torch.randn(T // world, d, device=DEV, generator=g).bfloat16() for _ in range(world)]
w_qkv = (torch.randn(6144 // world, d, device=DEV, generator=g) * 0.05).bfloat16()
landed = torch.empty(T, d, dtype=torch.bfloat16, device=DEV)
out = torch.empty(T, 6144 // world, dtype=torch.bfloat16, device=DEV)
# GEMM -> reduce-scatter (down, tp=2)
x_dn = torch.randn(T, ffn // world, device=DEV, generator=g).bfloat16()
w_dn = (torch.randn(d, ffn // world, device=DEV, generator=g) * 0.05).bfloat16()
targets = [torch.zeros(T // world, d, dtype=torch.bfloat16, device=DEV) for _ in range(world)]
# weight gather -> GEMM (ZeRO-3 qkv, dp=2)
unit = (torch.randn(6144 * d, device=DEV, generator=g) * 0.05).bfloat16()
shards = [unit[r * (unit.numel() // world):(r + 1) * (unit.numel() // world)].clone() for r in range(world)]
full = torch.empty_like(unit)
x = torch.randn(T, d, device=DEV, generator=g).bfloat16()
y = torch.empty(T, 6144, dtype=torch.bfloat16, device=DEV)
null = (ctypes.c_void_p * 1)(0)
ci, c64 = _lib.c_int, _lib.c_i64


def run():
    _lib.check(L.llmt_gemm_bf16_ag(arr(blocks), _lib.ptr(landed), _lib.ptr(w_qkv), _lib.ptr(out), _lib.ptr(flags),
                                   ci(0), ci(world), ci(T // world), ci(T), ci(6144 // world), ci(d), c64(d), c64(d),
                                   c64(d), c64(6144 // world), ci(0), ci(2), ci(0), _lib.stream()), 'ag')
    _lib.check(L.llmt_gemm_bf16_dist(ci(2), _lib.ptr(x_dn), null, _lib.ptr(w_dn), _lib.ptr(None), arr(targets),
                                     ci(world), ci(T // world), ci(T), ci(d), ci(ffn // world), c64(ffn // world),
                                     c64(ffn // world), c64(d), ci(0), ci(2), ci(0), _lib.stream()), 'rs')
    _lib.check(L.llmt_gemm_bf16_wgather(_lib.ptr(x), arr(shards), _lib.ptr(full), c64(0), c64(unit.numel() // world),
                                        _lib.ptr(y), _lib.ptr(flags), ci(0), ci(world), ci(T), ci(6144), ci(d),
                                        c64(d), c64(6144), ci(2), ci(0), _lib.stream()), 'wgather')


for _ in range(3):
    run()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
evs[0].record()
run()
evs[1].record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(f'three fused GEMMs: {evs[0].elapsed_time(evs[1]):.3f} ms')
