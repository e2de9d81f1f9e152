"""Achieved HBM bandwidth of the bandwidth-bound kernels at the model's shapes (CUDA events, L2 flushed between runs)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_training_b200.ops.rms_norm_op import rmsnorm_bwd_raw, rmsnorm_fwd_raw  # noqa: E402
from llm_training_b200.ops.swiglu_op import silu_mul_bwd_raw, silu_mul_fwd_raw  # noqa: E402


def timeit(fn, iters=10):
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.int8, device='cuda')
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    peak = 6544.0
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                           'MEASURED_PEAKS.json')))['hbm_gbs']
    except Exception:
        pass
    T, d, f = 8192, 4096, 14336
    gu = torch.randn(T, 2 * f, device='cuda').bfloat16()
    dact = torch.randn(T, f, device='cuda').bfloat16()
    x = torch.randn(T, d, device='cuda').bfloat16()
    w = torch.ones(d, device='cuda').bfloat16()
    _, rstd = rmsnorm_fwd_raw(x, w, 1e-5)
    rows = [('silu_mul_fwd', lambda: silu_mul_fwd_raw(gu), (2 * f + f) * T * 2),
            ('silu_mul_bwd', lambda: silu_mul_bwd_raw(dact, gu), (2 * f + f + 2 * f) * T * 2),
            ('rmsnorm_fwd', lambda: rmsnorm_fwd_raw(x, w, 1e-5), 2 * T * d * 2),
            ('rmsnorm_bwd', lambda: rmsnorm_bwd_raw(x, x, w, rstd), 3 * T * d * 2)]
    for name, fn, nbytes in rows:
        ms = timeit(fn)
        print(json.dumps({'kernel': name, 'ms': ms, 'algorithmic_GB': nbytes / 1e9, 'GBps': nbytes / ms / 1e6,
                          'of_measured_hbm_peak': nbytes / ms / 1e6 / peak}))


if __name__ == '__main__':
    main()
