"""Reference-EQUIVALENT baseline (NOT the unmodified reference: that cannot start here, see DESIGN.md §Reference arm).

It assembles, on this image's stack, the kernels and collectives the reference's hot loop resolves to
(SURVEY §3.2), through the reference's own ``HFCausalLM`` recipe (src/llm_training/models/hf_causal_lm/
hf_causal_lm.py:25-114): a Hugging Face ``LlamaForCausalLM`` with FlashAttention-2, Liger RMSNorm / SwiGLU /
cross-entropy, ``fully_shard`` per decoder layer + root (FSDP2 over NCCL), activation checkpointing as in the
reference's 8B example config, the reference's own ``MasterWeightsOptimizer`` (imported unmodified from
``baseline/_ref``) around ``torch.optim.AdamW``, and ``clip_grad_norm_``.  cuBLAS does the GEMMs.

Same metric, model, shapes and timing rules as ``bench.py``; prints one JSON line with ``"impl": "ref_equiv"``.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        baseline/ref_equiv_bench.py --gpus N --steps K --warmup W [--no-checkpointing]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--seq-len', type=int, default=8192)
    ap.add_argument('--layers', type=int, default=32)
    ap.add_argument('--no-checkpointing', action='store_true')
    ap.add_argument('--no-liger', action='store_true')
    args = ap.parse_args()

    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29533')
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', local))

    from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard
    from transformers import LlamaConfig, LlamaForCausalLM

    notes = []
    liger = False
    if not args.no_liger:
        try:
            from liger_kernel.transformers import apply_liger_kernel_to_llama
            apply_liger_kernel_to_llama(rope=False, cross_entropy=False, fused_linear_cross_entropy=False,
                                        rms_norm=True, swiglu=True)
            liger = True
        except Exception as e:  # noqa: BLE001
            notes.append(f'liger patch failed: {type(e).__name__}: {e}')

    cfg = LlamaConfig(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=args.layers,
                      num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=args.seq_len,
                      rope_theta=500000.0, rms_norm_eps=1e-5, tie_word_embeddings=False)
    cfg._attn_implementation = 'flash_attention_2'
    torch.manual_seed(0)
    with torch.device('cuda'):
        model = LlamaForCausalLM(cfg).to(torch.bfloat16)
    if not args.no_checkpointing:
        model.gradient_checkpointing_enable(gradient_checkpointing_kwargs={'use_reentrant': False})
    model.train()
    policy = MixedPrecisionPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16)
    for layer in model.model.layers:
        fully_shard(layer, mp_policy=policy)
    fully_shard(model, mp_policy=policy)

    sys.path.insert(0, os.path.join(ROOT, 'baseline', '_ref'))
    try:
        from llm_training.optim.master_weight_wrapper import MasterWeightsOptimizer
        opt = MasterWeightsOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-5))
        notes.append('optimizer: reference MasterWeightsOptimizer(torch.optim.AdamW)')
    except Exception as e:  # noqa: BLE001
        opt = torch.optim.AdamW(model.parameters(), lr=1e-5)
        notes.append(f'reference optimizer wrapper not importable ({type(e).__name__}); plain AdamW')

    ce = None
    if liger:
        try:
            from liger_kernel.transformers import LigerCrossEntropyLoss
            ce = LigerCrossEntropyLoss(ignore_index=-100)
        except Exception as e:  # noqa: BLE001
            notes.append(f'liger CE unavailable: {e}')

    T = args.seq_len
    gen = torch.Generator().manual_seed(1234 + rank)
    pool = [torch.randint(0, cfg.vocab_size, (1, T), generator=gen).pin_memory() for _ in range(4)]

    def step(i: int) -> torch.Tensor:
        ids = pool[i % len(pool)].cuda(non_blocking=True)
        labels = torch.roll(ids, -1, dims=1)
        labels[:, -1] = -100
        logits = model(input_ids=ids, use_cache=False).logits
        flat = logits.float().view(-1, logits.shape[-1])          # reference: clm.py:147 casts the logits to fp32
        loss = ce(flat, labels.view(-1)) if ce is not None else F.cross_entropy(flat, labels.view(-1))
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss.detach()

    for i in range(args.warmup):
        step(i)
    dist.barrier()
    torch.cuda.synchronize()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    last = None
    for i in range(args.steps):
        last = float(step(args.warmup + i))      # device -> host read of the loss every step, like the e2e arm
    end.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([start.elapsed_time(end) / args.steps], device='cuda')
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    peak_mem = torch.cuda.max_memory_allocated() / 2**30
    if rank == 0:
        print(json.dumps({
            'impl': 'ref_equiv', 'metric': 'tokens_per_sec', 'value': world * T / (ms / 1e3), 'unit': 'tokens/s',
            'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True,
            'scaling': 'weak', 'dtype': 'bf16', 'data': 'synthetic',
            'config': {'model': 'llama-3-8b' if args.layers == 32 else f'llama-3-8b-{args.layers}layers',
                       'global_batch': world, 'seq_len': T, 'parallelism': f'fsdp2 dp{world}',
                       'activation_checkpointing': not args.no_checkpointing, 'liger': liger,
                       'attention': 'flash_attention_2 (flash-attn 2.8)', 'gemm': 'cuBLAS'},
            'peak_mem_gib': round(peak_mem, 1), 'last_loss': last, 'notes': notes}))
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
