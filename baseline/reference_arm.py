"""Reference arm of bench.py: the UNMODIFIED reference tree installed in ``baseline/_ref`` on the bench config.

What runs is the reference's own code (nothing from ``llm_training_b200``, none of its kernels):

* model     ``llm_training.models.Llama`` (`/root/reference/src/llm_training/models/llama/llama_model.py:32-192`):
            its q/k/v/o + gate/up/down ``nn.Linear`` (cuBLAS), ``flash_attention_forward``
            (`ops/attention_op.py:538-654`, flash-attn 2.8), Liger RMSNorm / SiLU-mul wrappers, torch RoPE, and the
            per-layer ``torch.cuda.synchronize()`` at `llama_model.py:629`;
* sharding  ``Llama.parallelize`` -> ``configure_tensor_parallel`` / ``configure_fully_sharded_data_parallel``
            (`llama_model.py:197-268`) with FSDP2Strategy's defaults (``reshard_after_forward=True``,
            `lightning/strategy/fsdp2/fsdp2_strategy.py:56`) and FSDP2Precision's ``bf16-true`` policy
            (`lightning/strategy/fsdp2/fsdp2_precision.py:72-76`);
* step      ``CLM.training_step``'s sequence (`lms/clm/clm.py:136-168`): ``shift_labels``, forward,
            ``logits.float()``, Liger ``cross_entropy``; ``clip_grad_norm_`` (+ ``full_tensor()``) as
            `fsdp2_precision.py:165-169`; ``MasterWeightsOptimizer(torch.optim.AdamW)``
            (`optim/master_weight_wrapper.py:10-96`, wrapped as `fsdp2_strategy.py:249-250`);
            ``CosineAnnealingWarmupLR`` (`lr_schedulers/cosine.py`).

What does NOT run: the Lightning ``Trainer`` / ``LightningCLI`` / torchmetrics layer (not installable offline), so the
loop below is a bare loop that issues the calls above in the order Lightning would.  Missing third-party symbols are
shimmed in ``baseline/stubs`` (see ``ref_compat.py``); the reference's files are untouched.

Same timing rules as the main arm: W warm-up steps, K timed steps between barrier + synchronize, CUDA events, max
over ranks; ``e2e`` copies every step's batch from pinned host memory and reads the loss back; ``value`` keeps the
batch on the device and reads nothing.
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _isolate_imports() -> None:
    """`llm_training` must resolve to baseline/_ref, never to the repo's alias package of the same name."""
    drop = {ROOT, '', '.', os.getcwd()}
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or '.') not in {os.path.abspath(d or '.') for d in drop}]
    sys.path.insert(0, os.path.join(HERE, '_ref'))
    sys.path.insert(0, os.path.join(HERE, 'stubs'))
    for name in list(sys.modules):
        if name == 'llm_training' or name.startswith('llm_training.'):
            del sys.modules[name]


def run(args, models: dict, clock_sampler_cls) -> None:
    _isolate_imports()
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    cuda = torch.cuda.is_available()

    def bail(why: str) -> None:
        if rank == 0:
            print(json.dumps({'impl': 'reference', 'unavailable': why[:300]}))

    if world != args.gpus:
        raise SystemExit(f'reference arm: WORLD_SIZE={world} but --gpus {args.gpus}; launch with torch.distributed.run')
    try:
        import ref_compat
        notes = ref_compat.install()
        import llm_training
        assert os.path.abspath(llm_training.__file__).startswith(os.path.join(HERE, '_ref')), llm_training.__file__
        from llm_training.lr_schedulers import CosineAnnealingWarmupLR
        from llm_training.models import Llama, LlamaConfig
        from llm_training.ops import shift_labels
        from llm_training.ops.liger_kernel import cross_entropy
        from llm_training.optim.master_weight_wrapper import MasterWeightsOptimizer
    except Exception as e:  # noqa: BLE001 - any import problem means the arm is unavailable, not a crash
        bail(f'reference modules do not import on this image: {type(e).__name__}: {e}')
        return

    from torch.distributed._composable.fsdp import MixedPrecisionPolicy, OffloadPolicy
    from torch.distributed.device_mesh import init_device_mesh
    from torch.distributed.tensor import DTensor
    from torch.distributed.tensor.parallel import loss_parallel

    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29541')
    if cuda:
        torch.cuda.set_device(local_rank)
    device = torch.device('cuda', local_rank) if cuda else torch.device('cpu')
    dist.init_process_group('nccl' if cuda else 'gloo', rank=rank, world_size=world,
                            **({'device_id': device} if cuda else {}))
    tp = args.tp
    dp = world // tp
    mesh = init_device_mesh('cuda' if cuda else 'cpu', (dp, tp), mesh_dim_names=('data_parallel', 'tensor_parallel'))
    dtype = torch.bfloat16 if cuda else torch.float32

    base = dict(models[args.model])
    if args.layers > 0:
        base['num_hidden_layers'] = args.layers
    # activation checkpointing: the reference's 8B example turns it on (llama-3.1-8b_pt_example.yaml:40); the main arm
    # runs without.  'auto' gives the reference whichever is FASTER and fits: none when the per-GPU static state
    # (10 B/param sharded over dp: bf16 param + grad, fp32 master + grad + 2 moments = 20 B/param at peak) leaves room
    # for ~45 GB of un-checkpointed activations of an 8k-token 8B step, full otherwise.
    recompute = args.recompute
    if recompute == 'auto':
        recompute = 'none'
        if cuda:
            hd = base['hidden_size'] // base['num_attention_heads']
            lp = base['hidden_size'] * (base['num_attention_heads'] + 2 * base['num_key_value_heads']) * hd \
                + base['hidden_size'] ** 2 + 3 * base['hidden_size'] * base['intermediate_size']
            n_params = lp * base['num_hidden_layers'] + 2 * base['hidden_size'] * base['vocab_size']
            static_gb = 20 * n_params / (dp * tp) / 1e9
            act_gb = 1.35 * base['num_hidden_layers'] * (args.seq_len * args.micro_batch / 8192) \
                * (base['hidden_size'] / 4096) / tp + 12
            total_gb = torch.cuda.get_device_properties(device).total_memory / 1e9
            if static_gb + act_gb > 0.9 * total_gb:
                recompute = 'full'
    cfg = LlamaConfig(**base, enable_gradient_checkpointing=recompute != 'none',
                      recompute_granularity='selective' if recompute == 'selective' else 'full',
                      attn_implementation='flash_attention_2' if cuda else 'sdpa')

    torch.manual_seed(1234)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dtype)           # FSDP2Precision.module_init_context / convert_module: bf16-true
    try:
        with torch.device(device):
            model = Llama(cfg)               # BaseModel.__init__ runs init_weights() (random init, no checkpoint)
    finally:
        torch.set_default_dtype(prev)
    model.train()
    mp_policy = MixedPrecisionPolicy(param_dtype=dtype, reduce_dtype=dtype, output_dtype=dtype,
                                     cast_forward_inputs=False)
    model.parallelize(dp_mesh=mesh['data_parallel'], tp_mesh=mesh['tensor_parallel'],
                      reshard_after_forward=True, mp_policy=mp_policy, offload_policy=OffloadPolicy())

    # fused=True: the reference's example configs use a fused Adam (deepspeed FusedAdam); torch's default foreach
    # path materialises a 4 B/param temporary and does not fit 8B on one 180 GB GPU
    inner = torch.optim.AdamW(model.parameters(), lr=1e-5, weight_decay=0.1, **({'fused': True} if cuda else {}))
    optimizer = MasterWeightsOptimizer(inner) if dtype != torch.float32 else inner
    total_steps = 2 * (args.warmup + args.steps) + 2
    scheduler = CosineAnnealingWarmupLR(inner, num_warmup_steps=2, num_total_steps=total_steps, min_lr=1e-6)
    params = [p for p in model.parameters() if p.requires_grad]

    V, T, B = base['vocab_size'], args.seq_len, args.micro_batch

    def make_batch(seed: int) -> dict:
        g = torch.Generator().manual_seed(seed)
        ids = torch.randint(0, V, (B, T), generator=g)
        batch = {'input_ids': ids, 'attention_mask': torch.ones_like(ids),
                 'position_ids': torch.arange(T).unsqueeze(0).expand(B, T).contiguous(), 'labels': ids.clone()}
        return {k: (v.pin_memory() if cuda else v) for k, v in batch.items()}

    def train_step(batch: dict) -> torch.Tensor:
        labels = shift_labels(batch['labels'], -100)
        logits = model(input_ids=batch['input_ids'], attention_mask=batch['attention_mask'],
                       position_ids=batch['position_ids']).logits.float()
        if tp > 1:
            with loss_parallel():
                loss = torch.nn.functional.cross_entropy(logits.flatten(end_dim=1), labels.flatten(end_dim=1),
                                                         ignore_index=-100)
                loss.backward()
        else:
            loss = cross_entropy(logits=logits, labels=labels, ignore_index=-100)
            loss.backward()
        norm = torch.nn.utils.clip_grad_norm_(params, 1.0)
        if isinstance(norm, DTensor):
            norm = norm.full_tensor()
        optimizer.step()
        optimizer.zero_grad()
        scheduler.step()
        return loss

    def sync() -> None:
        dist.barrier()
        if cuda:
            torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(n: int, fn) -> float:
        sync()
        if cuda:
            start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        if cuda:
            end.record()
        sync()
        return reduce_max(start.elapsed_time(end) if cuda else (time.perf_counter() - t0) * 1e3) / n

    try:
        # ---- e2e: every step copies its batch from pinned host memory and reads the loss on the host
        host_batches = [make_batch(11 + 1000 * rank + i) for i in range(args.warmup + args.steps)]
        h2d = sum(v.numel() * v.element_size() for v in host_batches[0].values())
        last = {}

        def e2e_step(i: int, offset: int = 0) -> None:
            batch = {k: v.to(device, non_blocking=True) for k, v in host_batches[offset + i].items()}
            last['loss'] = float(train_step(batch).detach())

        for i in range(args.warmup):
            e2e_step(i)
        e2e_ms = timed(args.steps, lambda i: e2e_step(i, args.warmup))

        # ---- device-timed: batch resident, no host read
        resident = {k: v.to(device) for k, v in make_batch(99 + rank).items()}
        for _ in range(args.warmup):
            train_step(resident)
        sampler = clock_sampler_cls(local_rank)
        if rank == 0 and cuda:
            sampler.start()
        ms = timed(args.steps, lambda i: train_step(resident))
        clocks = sampler.stop() if (rank == 0 and cuda) else {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': []}
    except torch.OutOfMemoryError as e:
        bail(f'reference runs out of memory on this config at {world} GPU(s) (recompute={recompute}): {str(e)[:120]}')
        dist.destroy_process_group()
        return

    tokens = B * T * dp
    if rank == 0:
        print(json.dumps({
            'metric': 'tokens_per_sec', 'value': tokens / (ms / 1e3), 'unit': 'tokens/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16' if cuda else 'fp32', 'data': 'synthetic',
            'impl': 'reference', 'reference_class': 'reference modules, bare loop',
            'config': {'model': args.model if args.layers <= 0 else f'{args.model}-{args.layers}layers',
                       'global_batch': B * dp, 'seq_len': T,
                       'parallelism': f'dp{dp}' + (f'xtp{tp}' if tp > 1 else ''),
                       'sharding': ('none (single GPU, Llama.configure_fully_sharded_data_parallel returns at dp=1)'
                                    if dp == 1 else 'FSDP2 fully_shard per layer + embed/norm/lm_head, '
                                                    'reshard_after_forward=True'),
                       'comm': 'nccl' if cuda else 'gloo', 'reshard': dp > 1,
                       'optimizer': 'reference MasterWeightsOptimizer(torch.optim.AdamW fused), grad clip 1.0',
                       'recompute': recompute,
                       'l2': 'working set (16 GB bf16 weights) >> 126 MB L2, no flush needed'},
            'same_config': {'model': True, 'seq_len': True, 'dtype': True, 'optimizer': True,
                            'recompute': recompute == 'none', 'sharding': 'reference default (ZeRO-3 style)'},
            'clocks': clocks,
            'e2e': {'value': tokens / (e2e_ms / 1e3), 'unit': 'tokens/s', 'ms_per_step': e2e_ms,
                    'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 4,
                    'api': 'reference Llama + parallelize + MasterWeightsOptimizer, CLM.training_step sequence'},
            'gpu_launches': None, 'last_loss': last.get('loss'), 'notes': notes,
            'ours_loaded': any(m.startswith('llm_training_b200') for m in sys.modules),
        }))
    dist.barrier()
    dist.destroy_process_group()
