"""One packed-attention forward + backward at LLaMA-3-8B shape (for ncu captures of attn_fwd / attn_bwd)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_training_b200.ops.attention_op import packed_attention  # noqa: E402

T, HQ, HKV, D = int(os.environ.get('TOKENS', '8192')), 32, 8, 128
torch.manual_seed(0)
qkv = (torch.randn(T, (HQ + 2 * HKV) * D, device='cuda') * 0.5).bfloat16().requires_grad_()
start = torch.zeros(T, dtype=torch.int32, device='cuda')
end = torch.full((T,), T, dtype=torch.int32, device='cuda')
dout = torch.randn(T, HQ * D, device='cuda').bfloat16()
for _ in range(3):
    out = packed_attention(qkv, start, end, HQ, HKV, D)
    out.backward(dout)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
out = packed_attention(qkv, start, end, HQ, HKV, D)
s.record()
out.backward(dout)
e.record()
torch.cuda.synchronize()
print(f'attn_bwd_total_ms {s.elapsed_time(e):.3f}')
