"""Fused collective GEMMs vs NCCL + GEMM back to back, LLaMA-3-8B tensor-parallel shapes (torchrun, N ranks).
Times are CUDA-event medians, max over ranks.  Roofline = max(FLOPs / measured bf16 peak, NVLink bytes / 770 GB/s)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        dist.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], device='cuda')
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
    dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
    os.environ['LLMT_COMM'] = 'nvlink'
    from llm_training_b200.ops.gemm_op import gemm
    from llm_training_b200.parallel import init_mesh
    from llm_training_b200.parallel.tensor_parallel import (TensorParallelContext, _FusedGatherLinear, _gemm_dist,
                                                            _fused_linear_reduce_scatter)
    mesh = init_mesh(1, world)
    tp = TensorParallelContext(mesh.tp)
    assert tp.comm is not None
    T, d, ffn = 8192, 4096, 14336
    peak = 1424.4e12
    try:
        peak = json.load(open('MEASURED_PEAKS.json'))['bf16_tflops_sustained'] * 1e12
    except Exception:
        pass
    link = 770e9
    results = []
    g = torch.Generator('cuda').manual_seed(rank)
    for name, n_local in (('qkv', 6144 // world), ('gate_up', 2 * ffn // world)):
        x = (torch.randn(T // world, d, device='cuda', generator=g) * 0.5).bfloat16()
        w = (torch.randn(n_local, d, device='cuda', generator=g) * 0.05).bfloat16()
        full = torch.empty(T, d, dtype=x.dtype, device='cuda')

        def nccl():
            dist.all_gather_into_tensor(full, x)
            return gemm(full, w)

        def direct():       # earlier variant: A tiles TMA-loaded straight from the owner, once per N tile
            buf, view = tp.ring.next(T // world, d)
            view.copy_(x)
            tp.comm.barrier()
            out = torch.empty((T, n_local), dtype=x.dtype, device='cuda')
            _gemm_dist(1, None, buf.peer_ptrs_at(0), w, out, None, world, T // world, T, n_local, d, d, n_local, False)
            return out

        with torch.no_grad():
            fused = timeit(lambda: _FusedGatherLinear.apply(x, tp, w))
            base = timeit(nccl)
            direct_ms = timeit(direct)
            gemm_only = timeit(lambda: gemm(full, w))
            err = float((_FusedGatherLinear.apply(x, tp, w).float() - nccl().float()).abs().max())
        flops = 2.0 * T * n_local * d
        n_tiles = (n_local + 255) // 256
        bytes_min = (world - 1) / world * T * d * 2                   # every remote row once
        roof = max(flops / peak, bytes_min / link)
        results.append(dict(op=f'AG->GEMM {name}', shape=[T, n_local, d], fused_ms=fused, nccl_plus_gemm_ms=base,
                            roofline_ms=roof * 1e3, frac_of_roofline=roof * 1e3 / fused, max_abs_diff=err,
                            direct_from_peer_ms=direct_ms, local_gemm_only_ms=gemm_only,
                            direct_nvlink_reread_factor=n_tiles))
    for name, k_local in (('o_proj', d // world), ('down', ffn // world)):
        x = (torch.randn(T, k_local, device='cuda', generator=g) * 0.5).bfloat16()
        w = (torch.randn(d, k_local, device='cuda', generator=g) * 0.05).bfloat16()
        out = torch.empty(T // world, d, dtype=x.dtype, device='cuda')

        def nccl():
            y = gemm(x, w)
            dist.reduce_scatter_tensor(out, y)
            return out

        with torch.no_grad():
            fused = timeit(lambda: _fused_linear_reduce_scatter(tp, x, w, None, False))
            base = timeit(nccl)
            err = float((_fused_linear_reduce_scatter(tp, x, w, None, False).float() - nccl().float()).abs().max())
        flops = 2.0 * T * d * k_local
        bytes_min = (world - 1) / world * T * d * 2
        roof = max(flops / peak, bytes_min / link)
        results.append(dict(op=f'GEMM->RS {name}', shape=[T, d, k_local], fused_ms=fused, nccl_plus_gemm_ms=base,
                            roofline_ms=roof * 1e3, frac_of_roofline=roof * 1e3 / fused, max_abs_diff=err))
    if rank == 0:
        for r in results:
            print(json.dumps(r))
        os.makedirs('gpurun_out', exist_ok=True)
        json.dump(dict(world=world, results=results), open(f'gpurun_out/tp_fused_bench_{world}gpu.json', 'w'), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
