"""Weight-gathering GEMM (ZeRO-3, kDistA == 3) vs the plain tcgen05 GEMM at the model's shapes, on ONE GPU with the
`world` slices emulated in local memory: what the pull warps, the arrival counters and the n-major tile walk cost when
the "network" is free.  Prints one JSON line per shape."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_training_b200.ops import _lib  # noqa: E402
from llm_training_b200.ops.gemm_op import gemm  # noqa: E402


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    world, m = 8, 8192
    flags = torch.zeros(4096, dtype=torch.int32, device='cuda')
    for name, n, k in (('qkv', 6144, 4096), ('gate_up', 28672, 4096), ('o_proj', 4096, 4096), ('down', 4096, 14336)):
        g = torch.Generator(device='cuda').manual_seed(1)
        x = (torch.randn(m, k, device='cuda', generator=g) * 0.5).bfloat16()
        w = (torch.randn(n, k, device='cuda', generator=g) * 0.05).bfloat16()
        flat = w.reshape(-1)
        shard_numel = flat.numel() // world
        shards = [flat[r * shard_numel:(r + 1) * shard_numel].clone() for r in range(world)]
        ptrs = (ctypes.c_void_p * world)(*[s.data_ptr() for s in shards])
        full = torch.zeros_like(flat)
        y = torch.empty(m, n, dtype=torch.bfloat16, device='cuda')

        def fused():
            _lib.check(_lib.lib().llmt_gemm_bf16_wgather(
                _lib.ptr(x), ptrs, _lib.ptr(full), _lib.c_i64(0), _lib.c_i64(shard_numel), _lib.ptr(y), _lib.ptr(flags),
                _lib.c_int(0), _lib.c_int(world), _lib.c_int(m), _lib.c_int(n), _lib.c_int(k), _lib.c_i64(k),
                _lib.c_i64(n), _lib.c_int(2), _lib.c_int(0), _lib.stream()), 'gemm_bf16_wgather')

        t_plain = timeit(lambda: gemm(x, w))
        t_fused = timeit(fused)
        ok = bool(torch.equal(full, flat))
        print(json.dumps({'op': name, 'shape': [m, n, k], 'plain_gemm_ms': t_plain, 'weight_gather_gemm_ms': t_fused,
                          'overhead_pct': round(100 * (t_fused / t_plain - 1), 1), 'landed_weights_ok': ok,
                          'plain_tflops': 2 * m * n * k / t_plain / 1e9}))


if __name__ == '__main__':
    main()
