#!/usr/bin/env python
"""Headline benchmark: LLaMA-3-8B pre-training throughput (tokens/s, whole job) with FSDP-style sharding,
bf16, 8192-token packed sequences, synthetic data, random-init weights (BASELINE.json config 2).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 5 --warmup 3
    python bench.py --impl reference ...      # the reference arm (unavailable on this image)

Two numbers are measured on the same live model:
  * ``e2e``   : through the public API — ``Trainer.fit`` on ``CLM`` + ``DummyDataModule``; every timed step
                copies its batch from pinned host memory to the device and reads the loss back to the host.
  * ``value`` : the same training step driven directly, inputs already resident on the device, no host read.
Both are timed with CUDA events on the launching stream between barrier + synchronize, max over ranks.
The step's working set (16 GB of bf16 weights alone) is far larger than the 126 MB L2, so no explicit flush
is needed between iterations (stated in ``config.l2``).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MODELS = {
    'llama-3-8b': dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                       num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=8192,
                       rope_theta=500000.0, rms_norm_eps=1e-5, bos_token_id=128000, eos_token_id=128001),
    # BASELINE config 5 (run with --tp 8 --recompute full|selective)
    'llama-3.1-70b': dict(vocab_size=128256, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                          num_attention_heads=64, num_key_value_heads=8, max_position_embeddings=131072,
                          rope_theta=500000.0, rms_norm_eps=1e-5, bos_token_id=128000, eos_token_id=128001),
    # development proxy: at --tp 2 every rank holds the same weight shapes as one rank of llama-3-8b at --tp 8 (4 query
    # heads, 1 kv head, 1792 MLP columns, 16032 vocabulary rows), so host-launch-bound behaviour of the tensor-parallel
    # path can be studied on 2 GPUs
    'llama-3-8b-tp8-rank-slice-x2': dict(vocab_size=32064, hidden_size=4096, intermediate_size=3584,
                                         num_hidden_layers=32, num_attention_heads=8, num_key_value_heads=2,
                                         head_dim=128, max_position_embeddings=8192, rope_theta=500000.0,
                                         rms_norm_eps=1e-5),
    # for smoke runs on small memory / CPU
    'llama-tiny': dict(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                       num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=8192),
}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs (B200_PROFILING.md recipe)."""
    QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index: int) -> None:
        self.gpu_index = gpu_index
        self.proc = None
        self.lines: list[str] = []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.QUERY}', '--format=csv,noheader,nounits',
                                          '-lms', '200', '-i', str(self.gpu_index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self) -> None:
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.lines:
            parts = [p.strip() for p in line.split(',')]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[5:9]):
                if flag.lower().startswith('active'):
                    reasons.add(name)
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(smax) if smax else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def reference_arm(args) -> None:
    """Run the UNMODIFIED reference tree in baseline/_ref on the bench config (see baseline/reference_arm.py for
    exactly which of its modules run).  Nothing from llm_training_b200 is imported on this path."""
    ref = os.path.join(ROOT, 'baseline', '_ref')
    if not os.path.isdir(os.path.join(ref, 'llm_training')):
        if int(os.environ.get('RANK', '0')) == 0:
            print(json.dumps({'impl': 'reference', 'unavailable': 'baseline/_ref not installed (pip install '
                              '--no-index --no-deps --target baseline/_ref /root/reference; see DESIGN.md)'}))
        return
    sys.path.insert(0, os.path.join(ROOT, 'baseline'))
    import reference_arm as arm
    arm.run(args, MODELS, ClockSampler)


def relaunch_under_torchrun(n: int) -> None:
    """``python bench.py --gpus N`` (N > 1) without a launcher: start one rank per GPU ourselves."""
    import random
    port = str(29500 + random.randrange(0, 2000))
    os.execv(sys.executable, [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', f'--nproc-per-node={n}',
                              '--master-addr', '127.0.0.1', '--master-port', port, os.path.abspath(__file__),
                              *sys.argv[1:]])


def main() -> None:
    if os.environ.get('NCCL_DEBUG', '').upper() == 'VERSION':
        os.environ['NCCL_DEBUG'] = 'WARN'      # keep stdout to the ONE JSON line (NCCL prints its version there)
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--model', default='llama-3-8b', choices=sorted(MODELS))
    ap.add_argument('--seq-len', type=int, default=8192)
    ap.add_argument('--layers', type=int, default=0, help='override the layer count (comparison runs only)')
    ap.add_argument('--micro-batch', type=int, default=1)
    ap.add_argument('--recompute', default='auto', choices=['auto', 'none', 'full', 'selective'])
    ap.add_argument('--tp', type=int, default=1)
    ap.add_argument('--reshard', dest='reshard', action='store_true', default=True,
                    help='ZeRO-3 / FSDP reshard_after_forward=True: parameters, gradients and optimizer state sharded '
                         '(the default: it is what BASELINE config 2 names and what the reference arm runs)')
    ap.add_argument('--resident', dest='reshard', action='store_false',
                    help='keep the bf16 parameters resident on every GPU (ZeRO-2 style; ~2.5%% faster at 8 GPUs)')
    ap.add_argument('--docs', type=int, default=1,
                    help='documents per packed sequence (> 1: instruction-tuning shaped batches, BASELINE config 3: '
                         'segment-id attention masks that block cross-document attention, position ids restart per '
                         'document, prompt tokens carry no loss)')
    ap.add_argument('--neftune', type=float, default=0.0, help='NEFTune alpha (config 3 uses 5)')
    ap.add_argument('--phases', action='store_true', help='also report forward / backward / optimizer device time')
    ap.add_argument('--kernel-profile', default='', help='write rank 0\'s per-kernel device time of the timed steps '
                    '(torch.profiler / CUPTI) to this file; the printed numbers of such a run are NOT benchmark values')
    ap.add_argument('--host-profile', default='', help='write rank 0\'s cProfile of the timed steps (top functions by '
                    'own time) to this file; the printed numbers of such a run are NOT benchmark values')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    env_world = int(os.environ.get('WORLD_SIZE', '1'))
    if env_world != args.gpus:
        if 'RANK' not in os.environ and env_world == 1 and args.gpus > 1:
            relaunch_under_torchrun(args.gpus)
        print(f'bench.py: --gpus {args.gpus} but WORLD_SIZE={env_world}; the two must agree', file=sys.stderr)
        raise SystemExit(2)
    if args.impl == 'reference':
        reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    from llm_training_b200.data import DummyDataModule, DummyDataModuleConfig
    from llm_training_b200.lightning import Callback, FSDP2Strategy, Trainer, seed_everything
    from llm_training_b200.lms import CLM, CLMConfig
    from llm_training_b200.ops import _lib

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    cuda = torch.cuda.is_available()
    seed_everything(1234)

    recompute = args.recompute
    if recompute == 'auto':
        # 8B on one GPU: 128 GB of weights + grads + fp32 Adam state, ~24 GB of saved activations (the fused
        # norm+linear / swiglu+down blocks rebuild their elementwise intermediates) -> fits 180 GB without
        # layer-level checkpointing
        recompute = 'none'
    base_cfg = dict(MODELS[args.model])
    if args.layers > 0:
        base_cfg['num_hidden_layers'] = args.layers
    model_cfg = dict(base_cfg, enable_gradient_checkpointing=recompute != 'none',
                     recompute_granularity='selective' if recompute == 'selective' else 'full',
                     attn_implementation='flash_attention_2' if cuda else 'sdpa')
    lm = CLM(CLMConfig.model_validate(dict(
        init_weights=True, load_weights=False, neftune_alpha=args.neftune if args.neftune > 0 else None,
        model=dict(model_class='llm_training.models.Llama', model_config=model_cfg),
        optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=1e-5, weight_decay=0.1),
                   lr_scheduler_class='llm_training.lr_schedulers.CosineAnnealingWarmupLR',
                   lr_scheduler_kwargs=dict(num_warmup_steps=2, min_lr=1e-6)))))
    total_steps = args.warmup + args.steps
    dp = world // args.tp
    dm = DummyDataModule(DummyDataModuleConfig(batch_size=args.micro_batch, vocab_size=model_cfg['vocab_size'],
                                               max_length=args.seq_len,
                                               num_samples=(total_steps + 2) * args.micro_batch * dp, base_seed=11,
                                               pin_memory=True, num_workers=0))

    def pack_documents(sample: dict, seed: int) -> dict:
        """Turn one dummy sample into an instruction-tuning shaped pack: `--docs` documents of uneven length
        (segment ids 1..N in `attention_mask`, like InstructionTuningDataCollator with GROUP_BY_LENGTH packing),
        position ids restarting per document, the first third of every document masked out of the loss."""
        g = torch.Generator().manual_seed(seed)
        n = args.seq_len
        cuts = sorted(torch.randperm(n - 1, generator=g)[:args.docs - 1].add(1).tolist())
        bounds = [0] + cuts + [n]
        seg = torch.empty(n, dtype=torch.long)
        pos = torch.empty(n, dtype=torch.long)
        labels = sample['input_ids'].clone()
        for d, (a, b) in enumerate(zip(bounds[:-1], bounds[1:])):
            seg[a:b] = d + 1
            pos[a:b] = torch.arange(b - a)
            labels[a:a + (b - a) // 3] = -100
        return dict(input_ids=sample['input_ids'], attention_mask=seg, position_ids=pos, labels=labels)

    if args.docs > 1:
        from llm_training_b200.data.dummy.dummy_dataset import DummyDataset
        plain_getitem = DummyDataset.__getitem__
        DummyDataset.__getitem__ = lambda self, i: pack_documents(plain_getitem(self, i), 7000 + i)

    tokens_per_step = args.micro_batch * args.seq_len * dp

    class E2ETimer(Callback):
        """Times optimizer steps [warmup, warmup+steps) of Trainer.fit, reading the loss on the host each step."""

        def __init__(self) -> None:
            self.ms = None
            self.h2d = self.d2h = 0
            self.launch0 = 0

        def on_train_batch_start(self, trainer, module, batch, batch_idx) -> None:
            if batch_idx == 0:
                self.h2d = sum(v.numel() * v.element_size() for v in batch.values() if torch.is_tensor(v))

        def on_train_batch_end(self, trainer, module, outputs, batch, batch_idx) -> None:
            loss_host = float(outputs.detach())          # device -> host read of the step's loss, every step
            self.d2h = outputs.element_size()
            self.last_loss = loss_host
            if batch_idx == args.warmup - 1:    # timed region starts BEFORE the next batch is fetched / copied
                trainer.strategy.barrier()
                if cuda:
                    torch.cuda.synchronize()
                    self.start = torch.cuda.Event(enable_timing=True)
                    self.start.record()
                self.t0 = time.perf_counter()
            if batch_idx == args.warmup + args.steps - 1:
                trainer.strategy.synchronize_optimizer()     # the last step's overlapped update counts
                if cuda:
                    end = torch.cuda.Event(enable_timing=True)
                    end.record()
                trainer.strategy.barrier()
                if cuda:
                    torch.cuda.synchronize()
                    self.ms = self.start.elapsed_time(end)
                else:
                    self.ms = (time.perf_counter() - self.t0) * 1e3
                trainer.should_stop = True

    timer = E2ETimer()
    strategy = FSDP2Strategy(data_parallel_size=dp, tensor_parallel_size=args.tp,
                             reshard_after_forward=True if args.reshard else 'auto')
    # devices = --gpus: on a multi-GPU node a plain `python bench.py --gpus 1` must use exactly one GPU
    trainer = Trainer(strategy=strategy, precision='bf16-true' if cuda else '32-true', logger=False,
                      callbacks=[timer], max_steps=total_steps, gradient_clip_val=1.0, enable_progress_bar=False,
                      enable_checkpointing=False, accelerator='auto', devices=args.gpus)
    trainer.fit(lm, dm)                         # ---- e2e arm: the user's call
    device = trainer.strategy.root_device
    if trainer.strategy.world_size != args.gpus:
        print(f'bench.py: trainer runs on {trainer.strategy.world_size} ranks, --gpus says {args.gpus}', file=sys.stderr)
        raise SystemExit(2)

    def reduce_max(x: float) -> float:
        if dist.is_initialized() and world > 1:
            t = torch.tensor([x], dtype=torch.float64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    e2e_ms = reduce_max(timer.ms) / args.steps

    # ---- device-timed arm: same step, inputs resident on the device, no host read
    lm.train()
    g = torch.Generator().manual_seed(99 + rank)
    ids = torch.randint(0, model_cfg['vocab_size'], (args.micro_batch, args.seq_len), generator=g)
    batch = {'input_ids': ids, 'attention_mask': torch.ones_like(ids),
             'position_ids': torch.arange(args.seq_len).unsqueeze(0).expand(args.micro_batch, -1).contiguous(),
             'labels': ids}
    if args.docs > 1:
        rows = [pack_documents({'input_ids': ids[r]}, 9000 + r) for r in range(args.micro_batch)]
        batch = {k: torch.stack([row[k] for row in rows]) for k in rows[0]}
    batch = {k: v.to(device) for k, v in batch.items()}
    optimizer = trainer.optimizers[0]

    phase_events: list = []

    def step() -> None:
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if (args.phases and cuda) else None
        if marks:
            marks[0].record()
        loss = trainer.strategy.training_step(batch, 0)
        if marks:
            marks[1].record()
        trainer.strategy.backward(loss)
        if marks:
            marks[2].record()
        trainer.strategy.optimizer_step(optimizer)
        optimizer.zero_grad()
        if marks:
            marks[3].record()
            phase_events.append(marks)

    for _ in range(args.warmup):
        step()
    phase_events.clear()
    sampler = ClockSampler(local_rank)
    have_lib = cuda and _lib.kernels_available()
    if have_lib:
        _lib.lib().llmt_launch_count.restype = __import__('ctypes').c_ulonglong
    trainer.strategy.barrier()
    if cuda:
        torch.cuda.synchronize()
    if rank == 0 and cuda:
        sampler.start()
    launches0 = _lib.lib().llmt_launch_count() if have_lib else 0
    if cuda:
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
    if cuda and os.environ.get('LLMT_SYNC_DEBUG'):      # development: warn (with a stack) on every host-blocking op
        torch.cuda.set_sync_debug_mode(1)
    t0 = time.perf_counter()
    prof = None
    if args.host_profile and rank == 0:
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    kprof = None
    if args.kernel_profile and rank == 0 and cuda:
        from torch.profiler import ProfilerActivity, profile
        kprof = profile(activities=[ProfilerActivity.CUDA])
        kprof.__enter__()
    for _ in range(args.steps):
        step()
    if kprof is not None:
        torch.cuda.synchronize()
        kprof.__exit__(None, None, None)
        os.makedirs(os.path.dirname(os.path.abspath(args.kernel_profile)), exist_ok=True)
        open(args.kernel_profile, 'w').write(
            f'{args.steps} steps\n' + kprof.key_averages().table(sort_by='cuda_time_total', row_limit=40,
                                                                  max_name_column_width=90))
    if prof is not None:
        import io
        import pstats
        prof.disable()
        out = io.StringIO()
        pstats.Stats(prof, stream=out).sort_stats('tottime').print_stats(45)
        os.makedirs(os.path.dirname(os.path.abspath(args.host_profile)), exist_ok=True)
        open(args.host_profile, 'w').write(f'host wall {1e3 * (time.perf_counter() - t0) / args.steps:.2f} ms/step '
                                           f'(enqueue only)\n' + out.getvalue())
    if cuda and os.environ.get('LLMT_SYNC_DEBUG'):
        torch.cuda.set_sync_debug_mode(0)
    trainer.strategy.synchronize_optimizer()     # the last step's overlapped optimizer update is inside the timing
    if cuda:
        end.record()
    trainer.strategy.barrier()
    if cuda:
        torch.cuda.synchronize()
        ms_total = start.elapsed_time(end)
    else:
        ms_total = (time.perf_counter() - t0) * 1e3
    launches = (_lib.lib().llmt_launch_count() - launches0) if have_lib else 0
    clocks = sampler.stop() if (rank == 0 and cuda) else {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': []}
    ms = reduce_max(ms_total) / args.steps

    # model FLOPs per token: 6*N_layer_params + 6*d*V + 6*L*d*T (causal attention, fwd+bwd)
    c = model_cfg
    hd = c.get('head_dim') or c['hidden_size'] // c['num_attention_heads']
    layer_params = c['hidden_size'] * (c['num_attention_heads'] + 2 * c['num_key_value_heads']) * hd \
        + c['hidden_size'] * c['hidden_size'] + 3 * c['hidden_size'] * c['intermediate_size']
    # effective attended length: the whole sequence for pre-training packs, sum(len_i^2) / T for packed documents
    attended = args.seq_len
    if args.docs > 1:
        seg = batch['attention_mask'][0]
        attended = float(torch.bincount(seg)[1:].double().pow(2).sum()) / args.seq_len
    flops_per_token = 6 * layer_params * c['num_hidden_layers'] + 6 * c['hidden_size'] * c['vocab_size'] \
        + 6 * c['num_hidden_layers'] * c['hidden_size'] * attended
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['bf16_tflops_sustained']
    except Exception:
        pass
    value = tokens_per_step / (ms / 1e3)
    e2e_value = tokens_per_step / (e2e_ms / 1e3)
    phases = None
    if phase_events:
        fwd = sum(m[0].elapsed_time(m[1]) for m in phase_events) / len(phase_events)
        bwd = sum(m[1].elapsed_time(m[2]) for m in phase_events) / len(phase_events)
        opt = sum(m[2].elapsed_time(m[3]) for m in phase_events) / len(phase_events)
        phases = {'forward_ms': reduce_max(fwd), 'backward_ms': reduce_max(bwd), 'optimizer_ms': reduce_max(opt)}
    engines = trainer.strategy.engines
    zero3 = any(getattr(e, 'reshard', False) for e in engines)
    # which data plane actually ran (read from the live engines, not from the request)
    if world == 1:
        comm_name = 'none (single GPU)'
    elif all(e.nvlink for e in engines) and all(e.nvlink_grads for e in engines):
        comm_name = 'nvlink'
    elif any(e.nvlink for e in engines):
        comm_name = 'nvlink (parameters) + nccl (gradients)'
    else:
        comm_name = 'nccl'
    tp_ctx = getattr(getattr(lm, 'model', None), 'tp_context', None)
    if args.tp > 1:
        tp_plane = 'nvlink fused GEMM+collective kernels' if getattr(tp_ctx, 'comm', None) is not None else 'nccl'
        comm_name = f'tp: {tp_plane}; dp: {comm_name}'
    multicast = any(bool(getattr(getattr(e, a, None), 'multicast_ptr', 0)) and getattr(e.comm, 'use_multicast', False)
                    for e in engines for a in ('param_arena', 'grad_arena', 'shard_arena'))
    if rank == 0:
        mfu = (value * flops_per_token / 1e12 / (peak * world)) if peak else None
        print(json.dumps({
            'metric': 'tokens_per_sec', 'value': value, 'unit': 'tokens/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'bf16' if cuda else 'fp32', 'data': 'synthetic', 'impl': 'ours',
            'config': {'model': args.model if args.layers <= 0 else f'{args.model}-{args.layers}layers', 'global_batch': args.micro_batch * dp, 'seq_len': args.seq_len,
                       'parallelism': f'dp{dp}' + (f'xtp{args.tp}' if args.tp > 1 else ''),
                       'sharding': ('flat-param ZeRO-3 (params, grads and optimizer state sharded)' if zero3 else
                                    'flat-param ZeRO (optimizer+grad sharded, params resident)'),
                       'comm': comm_name, 'reshard': zero3, 'multicast': multicast,
                       'optimizer': 'AdamW fp32 master, grad clip 1.0', 'recompute': recompute,
                       **({'documents_per_sequence': args.docs} if args.docs > 1 else {}),
                       **({'neftune_alpha': args.neftune} if args.neftune > 0 else {}),
                       'l2': 'working set (16 GB bf16 weights) >> 126 MB L2, no flush needed'},
            'mfu_of_measured_sustained_peak': mfu, 'model_tflops_per_gpu': value * flops_per_token / 1e12 / world,
            'clocks': clocks,
            'e2e': {'value': e2e_value, 'unit': 'tokens/s', 'ms_per_step': e2e_ms,
                    'h2d_bytes_per_step': timer.h2d, 'd2h_bytes_per_step': timer.d2h,
                    'api': 'Trainer.fit(CLM, DummyDataModule)'},
            'gpu_launches': int(launches), 'last_loss': getattr(timer, 'last_loss', None),
            'peak_mem_gib': round(torch.cuda.max_memory_allocated() / 2 ** 30, 1) if cuda else None,
            **({'phases': phases} if phases else {}),
        }))
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
