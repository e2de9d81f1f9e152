"""Packed attention fwd/bwd throughput vs flash-attn 2.8 (the reference's kernel) on LLaMA-3-8B shapes."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_training_b200.ops.attention_op import packed_attention  # noqa: E402


def timeit(fn, iters=10):
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
    dev = 'cuda'
    results = []
    shapes = [(8192, 32, 8, 128), (16384, 32, 8, 128), (8192, 32, 32, 96)]
    if os.environ.get('ATTN_SHAPES'):       # e.g. ATTN_SHAPES=131072,4,1,128 (one rank's heads of the 128k TP=8 config)
        shapes = [tuple(int(v) for v in s.split(',')) for s in os.environ['ATTN_SHAPES'].split(';')]
    for tokens, hq, hkv, d in shapes:
        torch.manual_seed(0)
        qkv = (torch.randn(tokens, (hq + 2 * hkv) * d, device=dev) * 0.5).bfloat16().requires_grad_()
        start = torch.zeros(tokens, dtype=torch.int32, device=dev)
        end = torch.full((tokens,), tokens, dtype=torch.int32, device=dev)
        dout = torch.randn(tokens, hq * d, device=dev).bfloat16()
        fl_fwd = 4.0 * tokens * tokens * d * hq / 2
        out = packed_attention(qkv, start, end, hq, hkv, d)
        ms_f = timeit(lambda: packed_attention(qkv, start, end, hq, hkv, d))
        ms_b = timeit(lambda: torch.autograd.grad(out, qkv, dout, retain_graph=True))
        out_det = packed_attention(qkv, start, end, hq, hkv, d, deterministic=True)
        ms_bd = timeit(lambda: torch.autograd.grad(out_det, qkv, dout, retain_graph=True))
        r = dict(tokens=tokens, hq=hq, hkv=hkv, d=d, fwd_ms=ms_f, fwd_tflops=fl_fwd / ms_f / 1e9, bwd_ms=ms_b,
                 bwd_tflops=2.5 * fl_fwd / ms_b / 1e9, bwd_deterministic_ms=ms_bd)
        try:
            from flash_attn import flash_attn_func
            q, k, v = qkv.detach().split([hq * d, hkv * d, hkv * d], dim=-1)
            q = q.reshape(1, tokens, hq, d).contiguous().requires_grad_()
            k = k.reshape(1, tokens, hkv, d).contiguous().requires_grad_()
            v = v.reshape(1, tokens, hkv, d).contiguous().requires_grad_()
            o2 = flash_attn_func(q, k, v, causal=True)
            r['fa2_fwd_ms'] = timeit(lambda: flash_attn_func(q, k, v, causal=True))
            do2 = dout.view(1, tokens, hq, d)
            r['fa2_bwd_ms'] = timeit(lambda: torch.autograd.grad(o2, (q, k, v), do2, retain_graph=True))
            o2d = flash_attn_func(q, k, v, causal=True, deterministic=True)
            r['fa2_bwd_deterministic_ms'] = timeit(lambda: torch.autograd.grad(o2d, (q, k, v), do2, retain_graph=True))
            r['max_diff_vs_fa2'] = (o2.view(tokens, -1).float() - out.float()).abs().max().item()
            ours = torch.autograd.grad(out, qkv, dout, retain_graph=True)[0]
            theirs = torch.cat([t.reshape(tokens, -1) for t in torch.autograd.grad(o2, (q, k, v), do2, retain_graph=True)],
                               dim=-1)
            r['max_grad_diff_vs_fa2'] = (ours.float() - theirs.float()).abs().max().item()
            r['grad_abs_max'] = theirs.float().abs().max().item()
        except Exception as ex:  # flash-attn missing or failing on this box
            r['fa2_error'] = repr(ex)[:200]
        results.append(r)
        print(json.dumps(r), flush=True)
    os.makedirs('gpurun_out', exist_ok=True)
    json.dump(results, open('gpurun_out/attn_check.json', 'w'), indent=1)


if __name__ == '__main__':
    main()
