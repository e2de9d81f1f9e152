"""DPO / ORPO step time on Phi-3-mini (BASELINE config 4): fused linear + log-prob loss, FSDP engine, synthetic pairs.

    python bench/objective_bench.py --objective dpo --steps 3 --warmup 3               (1 GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench/objective_bench.py --objective orpo                                      (N GPUs, data parallel)

One step = policy (and, for DPO, the frozen reference) forward on chosen + rejected sequences of ``--seq-len`` tokens,
backward, clip, AdamW.  Timed with CUDA events between barriers, max over ranks; prints one JSON line.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

PHI3_MINI = dict(vocab_size=32064, hidden_size=3072, intermediate_size=8192, num_hidden_layers=32,
                 num_attention_heads=32, num_key_value_heads=32, max_position_embeddings=4096,
                 original_max_position_embeddings=4096, rms_norm_eps=1e-5, pad_token_id=32000)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--objective', default='dpo', choices=['dpo', 'orpo'])
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--seq-len', type=int, default=4096)
    ap.add_argument('--pairs', type=int, default=2, help='preference pairs per GPU per step')
    ap.add_argument('--layers', type=int, default=0)
    ap.add_argument('--reshard', action='store_true', help='ZeRO-3: policy AND frozen reference model sharded')
    ap.add_argument('--loss', default='fused', choices=['fused', 'materialized'],
                    help="'materialized' computes the sequence log-probs the reference's way (full [B, T, V] logits -> "
                         'fp32 log_softmax -> gather; /root/reference/src/llm_training/lms/dpo/dpo.py:110-114, '
                         'lms/orpo/orpo.py:89-93) on the same model and kernels, so the fused linear + log-prob loss '
                         'has a like-for-like ratio')
    args = ap.parse_args()
    world = int(os.environ.get('WORLD_SIZE', 1))
    rank = int(os.environ.get('RANK', 0))

    from llm_training_b200.lightning import FSDP2Strategy, Trainer
    from llm_training_b200.lms import DPO, DPOConfig, ORPO, ORPOConfig
    cfg = dict(PHI3_MINI)
    if args.layers:
        cfg['num_hidden_layers'] = args.layers
    common = dict(init_weights=True, load_weights=False,
                  model=dict(model_class='llm_training.models.Phi3', model_config=cfg),
                  optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=1e-6)))
    lm = DPO(DPOConfig.model_validate(common)) if args.objective == 'dpo' else ORPO(ORPOConfig.model_validate(common))
    trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=world,
                                             reshard_after_forward=True if args.reshard else 'auto'),
                      precision='bf16-true', logger=False,
                      enable_progress_bar=False, enable_checkpointing=False, gradient_clip_val=1.0)
    lm.trainer = trainer
    trainer.lightning_module = lm
    trainer.strategy.setup_environment(trainer)
    with trainer.strategy.model_init_context():
        lm.configure_model()
    if args.objective == 'dpo' and getattr(lm, 'ref_model', None) is not None:
        # the reference model starts as a copy of the policy (both load the same pre-trained weights in real use)
        lm.ref_model.load_state_dict(lm.model.state_dict())
    if args.loss == 'materialized':
        for m in (lm.model, getattr(lm, 'ref_model', None)):
            if m is not None:
                m.supports_fused_loss = False       # sequence_logps then takes its logits -> log_softmax branch
    trainer.strategy.setup(trainer)
    dev = trainer.strategy.root_device
    g = torch.Generator().manual_seed(7 + rank)
    batch = {}
    for side in ('chosen', 'rejected'):
        ids = torch.randint(0, 32000, (args.pairs, args.seq_len), generator=g)
        labels = ids.clone()
        labels[:, :args.seq_len // 4] = -100                     # prompt tokens carry no loss
        batch[f'{side}_input_ids'] = ids.to(dev)
        batch[f'{side}_labels'] = labels.to(dev)
        batch[f'{side}_attention_mask'] = torch.ones_like(ids).to(dev)
        batch[f'{side}_position_ids'] = torch.arange(args.seq_len).unsqueeze(0).expand(args.pairs, -1).to(dev)
    opt = trainer.optimizers[0]

    def step() -> torch.Tensor:
        loss = trainer.strategy.training_step(batch, 0)
        trainer.strategy.backward(loss)
        trainer.strategy.optimizer_step(opt)
        opt.zero_grad()
        return loss

    for _ in range(args.warmup):
        step()
    trainer.strategy.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(args.steps):
        loss = step()
    trainer.strategy.synchronize_optimizer()
    e.record()
    trainer.strategy.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e) / args.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    tokens = 2 * args.pairs * args.seq_len * world
    if rank == 0:
        print(json.dumps({'metric': f'{args.objective}_tokens_per_sec', 'value': tokens / ms * 1e3, 'unit': 'tokens/s',
                          'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms,
                          'dtype': 'bf16', 'data': 'synthetic',
                          'config': {'model': 'phi-3-mini' + (f'-{args.layers}layers' if args.layers else ''),
                                     'pairs_per_gpu': args.pairs, 'seq_len': args.seq_len,
                                     'parallelism': f'dp{world}', 'objective': args.objective, 'loss': args.loss,
                                     'reshard': args.reshard,
                                     'frozen_engines': len(trainer.strategy.frozen_engines)},
                          'peak_mem_gib': round(torch.cuda.max_memory_allocated() / 2**30, 1),
                          'last_loss': float(loss)}))
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
