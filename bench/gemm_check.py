"""GEMM correctness sweep + throughput vs cuBLAS (torch.matmul), one case per subprocess so that a
faulting configuration cannot take the rest of the sweep down.  Usage (on a B200):
    python bench/gemm_check.py            # all cases
    python bench/gemm_check.py --case ... # internal
"""
import argparse
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run_case(args) -> dict:
    import torch
    from llm_training_b200.ops.gemm_op import gemm
    torch.manual_seed(0)
    M, N, K = args.m, args.n, args.k
    dev = 'cuda'
    a = (torch.randn((K, M) if args.amn else (M, K), device=dev) * 0.5).bfloat16()
    b = (torch.randn((K, N) if args.bmn else (N, K), device=dev) * 0.5).bfloat16()
    out_dtype = torch.float32 if args.fp32 else torch.bfloat16
    res = (torch.randn((M, N), device=dev)).to(out_dtype) if args.res else None
    A = a.float().t() if args.amn else a.float()
    B = b.float() if args.bmn else b.float().t()
    ref = A @ B
    if res is not None:
        ref = ref + res.float()
    c = gemm(a, b, a_mn_major=bool(args.amn), b_mn_major=bool(args.bmn), residual=res, out_dtype=out_dtype,
             cta_group=args.cg)
    torch.cuda.synchronize()
    err = (c.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    out = {'max_abs_err': err, 'ref_max': scale, 'ok': bool(err <= 2e-2 * scale + 1e-3)}
    if args.time:
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.int8, device=dev)

        def timeit(fn, iters=10):
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
        ms = timeit(lambda: gemm(a, b, a_mn_major=bool(args.amn), b_mn_major=bool(args.bmn), residual=res,
                                 out_dtype=out_dtype, cta_group=args.cg))
        Ab = a.t() if args.amn else a
        Bb = b if args.bmn else b.t()
        ms_ref = timeit(lambda: torch.matmul(Ab, Bb))
        fl = 2.0 * M * N * K
        out.update(ms=ms, tflops=fl / ms / 1e9, cublas_ms=ms_ref, cublas_tflops=fl / ms_ref / 1e9)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--case', action='store_true')
    for name in ('m', 'n', 'k', 'amn', 'bmn', 'fp32', 'res', 'cg', 'time'):
        ap.add_argument('--' + name, type=int, default=0)
    ap.add_argument('--quick', action='store_true')
    args = ap.parse_args()
    if args.case:
        print('RESULT ' + json.dumps(run_case(args)))
        return
    cases = []
    # correctness: odd shapes, every variant
    for cg in (1, 2):
        for amn, bmn in ((0, 0), (0, 1), (1, 1)):
            cases.append(dict(m=256, n=256, k=128, amn=amn, bmn=bmn, cg=cg))
            cases.append(dict(m=392, n=712, k=328, amn=amn, bmn=bmn, cg=cg, res=1))
            cases.append(dict(m=1024, n=1536, k=512, amn=amn, bmn=bmn, cg=cg, fp32=1, res=1))
    if not args.quick:
        for cg in (1, 2):
            cases.append(dict(m=8192, n=6144, k=4096, cg=cg, time=1))
            cases.append(dict(m=8192, n=4096, k=4096, cg=cg, time=1, res=1))
            cases.append(dict(m=8192, n=28672, k=4096, cg=cg, time=1))
            cases.append(dict(m=8192, n=4096, k=14336, cg=cg, time=1, res=1))
            cases.append(dict(m=8192, n=4096, k=6144, bmn=1, cg=cg, time=1))           # dgrad qkv
            cases.append(dict(m=28672, n=4096, k=8192, amn=1, bmn=1, cg=cg, time=1))   # wgrad gate_up
            cases.append(dict(m=8192, n=8192, k=8192, cg=cg, time=1))
    results = []
    for c in cases:
        cmd = [sys.executable, __file__, '--case'] + [f'--{k}={v}' for k, v in c.items()]
        try:
            p = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
            line = [l for l in p.stdout.splitlines() if l.startswith('RESULT ')]
            r = json.loads(line[0][7:]) if line else {'ok': False, 'error': (p.stderr or p.stdout)[-400:]}
        except subprocess.TimeoutExpired:
            r = {'ok': False, 'error': 'timeout'}
        r['case'] = c
        results.append(r)
        print(json.dumps(r), flush=True)
    bad = [r for r in results if not r.get('ok')]
    print(f'{len(results) - len(bad)}/{len(results)} ok')
    os.makedirs('gpurun_out', exist_ok=True)
    with open('gpurun_out/gemm_check.json', 'w') as f:
        json.dump(results, f, indent=1)
    sys.exit(1 if bad else 0)


if __name__ == '__main__':
    main()
