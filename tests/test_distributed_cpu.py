"""Multi-process plumbing on CPU/gloo (BASELINE config 1): sharded-vs-single-process parity of the
flat-parameter engine, gradient accumulation without sync, checkpoint save -> resume."""

import pytest
import torch

from tests.dist_utils import run_distributed

TINY = dict(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
            num_key_value_heads=2, attn_implementation='sdpa')


def _make_model(seed: int = 0):
    from llm_training_b200.models import Llama, LlamaConfig
    torch.manual_seed(seed)
    return Llama(LlamaConfig(**TINY))


def _batch(seed: int, rows: int = 4, seq: int = 24):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, TINY['vocab_size'], (rows, seq), generator=g)
    return ids


def _loss(model, ids):
    logits = model(input_ids=ids).logits.float()
    return torch.nn.functional.cross_entropy(logits[:, :-1].flatten(0, 1), ids[:, 1:].flatten())


def _dp_worker(rank: int, world: int, steps: int, accum: int, reshard: bool = False):
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import apply_fully_sharded, init_mesh
    mesh = init_mesh(world, 1)
    model = _make_model()
    engine = apply_fully_sharded(model, mesh.dp, reshard_after_forward=reshard)
    assert engine.reshard == reshard
    opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.1), [engine],
                           max_grad_norm=1.0)
    losses = []
    for step in range(steps):
        opt.zero_grad()
        for micro in range(accum):
            ids = _batch(100 * step + micro)             # global batch of 4 rows
            local = ids[rank::world]                     # this rank's rows
            engine.set_requires_gradient_sync(micro == accum - 1)
            loss = _loss(model, local) / accum
            loss.backward()
            if micro != accum - 1:
                engine._rearm()
            losses.append(loss.item())
        opt.step()
        if reshard:      # ZeRO-3: between steps no unit keeps its full parameter or gradient buffer
            assert all(u.is_sharded and not u.grad_live for u in engine.units), \
                [(u.name, u.is_sharded, u.grad_live) for u in engine.units]
    with engine.unsharded():
        params = {n: p.detach().clone() for n, p in model.named_parameters()}
    return params, losses, float(opt.last_grad_norm)


@pytest.mark.parametrize('accum,reshard', [(1, False), (2, False), (1, True), (2, True)])
def test_sharded_data_parallel_matches_single_process(accum, reshard):
    """Resident (ZeRO-2) and resharding (ZeRO-3: gather before use, drop after use) engines both reproduce the
    single-process run."""
    steps = 3
    results = run_distributed(_dp_worker, world=2, args=(steps, accum, reshard))
    ref = _make_model()
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.1)
    for step in range(steps):
        ropt.zero_grad()
        for micro in range(accum):
            ids = _batch(100 * step + micro)
            # mean over ranks of per-rank means == what the reduce-scatter(avg) produces
            loss = sum(_loss(ref, ids[r::2]) for r in range(2)) / 2 / accum
            loss.backward()
        norm = torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        ropt.step()
    for rank in range(2):
        params, _, gnorm = results[rank]
        assert abs(gnorm - float(norm)) < 1e-4 * max(1.0, float(norm))
        for n, p in ref.named_parameters():
            assert torch.allclose(params[n], p.detach(), atol=2e-5, rtol=1e-4), n
    # both ranks hold identical full parameters after the in-place all-gather
    for n in results[0][0]:
        assert torch.equal(results[0][0][n], results[1][0][n]), n


def _fit_worker(rank: int, world: int, tmpdir: str, ckpt_path, max_steps: int, reshard: bool = False):
    from llm_training_b200.data import DummyDataModule, DummyDataModuleConfig
    from llm_training_b200.lightning import FSDP2Strategy, ModelCheckpoint, Trainer, seed_everything
    from llm_training_b200.lms import CLM, CLMConfig
    seed_everything(42)
    lm = CLM(CLMConfig.model_validate(dict(
        init_weights=True, load_weights=False,
        model=dict(model_class='llm_training.models.Llama', model_config=TINY),
        optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=5e-3),
                   lr_scheduler_class='llm_training.lr_schedulers.CosineAnnealingWarmupLR',
                   lr_scheduler_kwargs=dict(num_warmup_steps=2, min_lr=1e-4)))))
    dm = DummyDataModule(DummyDataModuleConfig(batch_size=2, vocab_size=128, max_length=16, num_samples=48,
                                               base_seed=5, pin_memory=False))
    trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=world, reshard_after_forward=reshard),
                      precision='32-true', logger=False,
                      enable_progress_bar=False, max_steps=max_steps, gradient_clip_val=1.0, accelerator='cpu',
                      default_root_dir=tmpdir, max_epochs=1,
                      callbacks=[ModelCheckpoint(dirpath=tmpdir + '/ckpt', every_n_train_steps=4, save_top_k=-1)])
    trainer.fit(lm, dm, ckpt_path=ckpt_path)
    assert all(e.reshard == reshard for e in trainer.strategy.engines)
    import contextlib
    with contextlib.ExitStack() as stack:
        for e in trainer.strategy.engines:
            stack.enter_context(e.unsharded())               # ZeRO-3: parameters only exist whole inside
        params = {n: p.detach().clone() for n, p in lm.named_parameters()}
    return params, trainer.global_step, int(lm.consumed_samples.compute())


def test_fit_checkpoint_resume_is_bit_identical(tmp_path):
    full = run_distributed(_fit_worker, world=2, args=(str(tmp_path / 'a'), None, 8))
    assert full[0][1] == 8
    ckpt = str(tmp_path / 'a' / 'ckpt' / 'epoch=0-step=4.ckpt')
    import os
    assert os.path.isdir(ckpt) and os.path.exists(os.path.join(ckpt, 'meta.pt')) \
        and os.path.exists(os.path.join(ckpt, '.metadata'))
    resumed = run_distributed(_fit_worker, world=2, args=(str(tmp_path / 'b'), ckpt, 8))
    assert resumed[0][1] == 8
    assert resumed[0][2] == full[0][2]                        # consumed-samples metric restored and continued
    for n, p in full[0][0].items():
        assert torch.allclose(resumed[0][0][n], p, atol=1e-6), n


def test_zero3_fit_checkpoint_resume_matches_resident_run(tmp_path):
    """The same Trainer.fit in ZeRO-3 mode (reshard_after_forward=True): identical weights to the resident run, and a
    sharded checkpoint written from slices-only state resumes to the same result."""
    resident = run_distributed(_fit_worker, world=2, args=(str(tmp_path / 'r'), None, 8, False))
    zero3 = run_distributed(_fit_worker, world=2, args=(str(tmp_path / 'z'), None, 8, True))
    for n, p in resident[0][0].items():
        assert torch.allclose(zero3[0][0][n], p, atol=1e-6), n
    ckpt = str(tmp_path / 'z' / 'ckpt' / 'epoch=0-step=4.ckpt')
    resumed = run_distributed(_fit_worker, world=2, args=(str(tmp_path / 'z2'), ckpt, 8, True))
    assert resumed[0][1] == 8
    for n, p in zero3[0][0].items():
        assert torch.allclose(resumed[0][0][n], p, atol=1e-6), n
    # and a ZeRO-3 checkpoint loads into a resident engine (same file format)
    crossed = run_distributed(_fit_worker, world=2, args=(str(tmp_path / 'z3'), ckpt, 8, False))
    for n, p in zero3[0][0].items():
        assert torch.allclose(crossed[0][0][n], p, atol=1e-6), n


def _tp_worker(rank: int, world: int, model_kind: str, dp: int, tp: int, reshard: bool = False):
    from llm_training_b200.models import Llama, LlamaConfig, Phi3, Phi3Config
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import apply_fully_sharded, init_mesh
    from llm_training_b200.parallel.comm import get_comm
    from llm_training_b200.ops import shift_labels
    mesh = init_mesh(dp, tp)
    torch.manual_seed(0)
    if model_kind == 'llama':
        model = Llama(LlamaConfig(**TINY))
    else:
        model = Phi3(Phi3Config(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                                num_attention_heads=4, num_key_value_heads=2, pad_token_id=0,
                                attn_implementation='sdpa'))
    model.parallelize(dp_mesh=mesh.dp, tp_mesh=mesh.tp, reshard_after_forward=reshard)
    engine = model._flat_engine if getattr(model, '_flat_engine', None) is not None \
        else apply_fully_sharded(model, mesh.dp)
    opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.1), [engine],
                           max_grad_norm=1.0, norm_extra_axes=[get_comm(mesh.tp, 'torch')])
    losses = []
    for step in range(3):
        ids = _batch(200 + step)                      # [4, 24]
        local = ids[mesh.dp.rank::dp]                 # dp split; tp peers see the same rows
        opt.zero_grad()
        out = model(input_ids=local, skip_logits=True)
        labels = shift_labels(local).flatten()
        loss = model.tp_context.vocab_parallel_cross_entropy(out.normed_hidden_states, model.lm_head.weight, labels)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    full = {}
    with engine.unsharded():                          # no-op for the resident engine, gathers under ZeRO-3
        for n, p in model.named_parameters():
            info = getattr(p, '_tp_info', None)
            full[n] = info.gather(p.detach()) if info is not None else p.detach().clone()
    return full, losses, float(opt.last_grad_norm)


@pytest.mark.parametrize('model_kind,dp,tp,reshard', [('llama', 1, 2, False), ('phi3', 1, 2, False),
                                                      ('llama', 2, 2, False), ('llama', 2, 2, True)])
def test_tensor_sequence_parallel_matches_single_process(model_kind, dp, tp, reshard):
    from llm_training_b200.models import Llama, LlamaConfig, Phi3, Phi3Config
    results = run_distributed(_tp_worker, world=dp * tp, args=(model_kind, dp, tp, reshard), timeout=600)
    torch.manual_seed(0)
    if model_kind == 'llama':
        ref = Llama(LlamaConfig(**TINY))
    else:
        ref = Phi3(Phi3Config(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                              num_attention_heads=4, num_key_value_heads=2, pad_token_id=0,
                              attn_implementation='sdpa'))
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.1)
    ref_losses = []
    for step in range(3):
        ids = _batch(200 + step)
        ropt.zero_grad()
        loss = sum(_loss(ref, ids[r::dp]) for r in range(dp)) / dp
        loss.backward()
        norm = torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        ropt.step()
        ref_losses.append(loss.item())
    full, losses, gnorm = results[0]
    if dp == 1:
        for a, b in zip(losses, ref_losses):
            assert abs(a - b) < 1e-4, (losses, ref_losses)
    assert abs(gnorm - float(norm)) < 1e-3 * max(1.0, float(norm)), (gnorm, float(norm))
    for n, p in ref.named_parameters():
        diff = (full[n] - p.detach()).abs()
        # Adam turns round-off in near-zero gradients into O(lr) differences for a handful of elements
        assert diff.max().item() < 2e-3 and (diff > 5e-5).float().mean().item() < 1e-3, (n, diff.max().item())


# ---------------------------------------------------------------------------------------------------
# BASELINE config 1: tiny LLaMA (2 layers, d=256) pre-training with best-fit bin packing, CPU/gloo, world_size=2
# ---------------------------------------------------------------------------------------------------
def _config1_worker(rank: int, world: int, data_path: str, tmpdir: str):
    import datasets
    from llm_training_b200.data import PreTrainingDataModule, PreTrainingDataModuleConfig
    from llm_training_b200.lightning import FSDP2Strategy, Trainer, seed_everything
    from llm_training_b200.lms import CLM, CLMConfig
    from tests.helpers import tiny_tokenizer
    seed_everything(7)

    class DM(PreTrainingDataModule):
        def load_data(self):
            return datasets.DatasetDict({'train': datasets.load_from_disk(data_path)})

    dm = DM(PreTrainingDataModuleConfig(tokenizer=tiny_tokenizer(), max_length=48,
                                        packing_method='best_fit_bin_packing', batch_size=2, pin_memory=False,
                                        enable_cache=False, num_workers=0))
    lm = CLM(CLMConfig.model_validate(dict(
        init_weights=True, load_weights=False,
        model=dict(model_class='llm_training.models.Llama',
                   model_config=dict(vocab_size=256, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                                     num_attention_heads=4, num_key_value_heads=2, attn_implementation='sdpa')),
        optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=3e-3)))))
    losses = []

    class Record:
        def on_train_batch_end(self, trainer, module, outputs, batch, batch_idx):
            losses.append(float(outputs.detach()))

    from llm_training_b200.lightning import Callback
    rec = type('Rec', (Callback,), {'on_train_batch_end': Record.on_train_batch_end})()
    trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=world), precision='32-true', logger=False,
                      enable_progress_bar=False, enable_checkpointing=False, max_steps=6, gradient_clip_val=1.0,
                      accelerator='cpu', default_root_dir=tmpdir, callbacks=[rec])
    trainer.fit(lm, dm)
    packed = dm.dataset_dict['train']
    return losses, [len(x) for x in packed['input_ids']], {n: p.detach().clone() for n, p in lm.named_parameters()}


def test_baseline_config1_best_fit_packing_pretraining_two_ranks(tmp_path):
    import datasets
    texts = [' '.join(f'w{(i * 13 + j) % 200}' for j in range(4 + (i * 5) % 23)) for i in range(96)]
    data_path = str(tmp_path / 'raw')
    datasets.Dataset.from_dict({'text': texts}).save_to_disk(data_path)
    two = run_distributed(_config1_worker, world=2, args=(data_path, str(tmp_path / 'w2')), timeout=600)
    losses0, rows, params0 = two[0]
    losses1, _, params1 = two[1]
    assert max(rows) <= 48 and sum(r > 40 for r in rows) >= len(rows) // 2      # best-fit bins are well filled
    assert len(losses0) == 6 and losses0[-1] < losses0[0] and losses1[-1] < losses1[0]
    assert losses0 != losses1                                                    # the ranks saw different rows
    for n, p in params0.items():                                                 # and hold identical weights
        assert torch.equal(p, params1[n]), n


def _global_tokens_worker(rank: int, world: int, mode: str):
    from llm_training_b200.lightning import FSDP2Strategy, Trainer, seed_everything
    from llm_training_b200.lms import CLM, CLMConfig
    seed_everything(3)
    lm = CLM(CLMConfig.model_validate(dict(
        init_weights=True, load_weights=False, loss_normalization=mode,
        model=dict(model_class='llm_training.models.Llama', model_config=TINY),
        optim=dict(optimizer_class='torch.optim.SGD', optimizer_kwargs=dict(lr=0.1)))))
    trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=world), precision='32-true', logger=False,
                      enable_progress_bar=False, enable_checkpointing=False, accelerator='cpu')
    lm.trainer = trainer
    trainer.lightning_module = lm
    trainer.strategy.setup_environment(trainer)
    with trainer.strategy.model_init_context():
        lm.configure_model()
    trainer.strategy.setup(trainer)
    ids = _batch(77, rows=2, seq=24)[rank:rank + 1]
    labels = ids.clone()
    if rank == 1:
        labels[:, :20] = -100                       # rank 1 has 4 valid tokens, rank 0 has 24
    batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids), position_ids=torch.arange(24)[None], labels=labels)
    loss = trainer.strategy.training_step(batch, 0)
    trainer.strategy.backward(loss)
    trainer.strategy.optimizer_step(trainer.optimizers[0])
    with trainer.strategy.engines[0].unsharded():
        return {n: p.detach().clone() for n, p in lm.model.named_parameters()}


@pytest.mark.parametrize('mode', ['micro_batch', 'global_tokens'])
def test_loss_normalization_over_data_parallel_ranks(mode):
    """`micro_batch` reproduces the reference (mean of the ranks' means); `global_tokens` equals the single-process
    mean over all valid tokens of the step."""
    from llm_training_b200.lightning import seed_everything
    from llm_training_b200.ops import shift_labels
    results = run_distributed(_global_tokens_worker, world=2, args=(mode,))
    seed_everything(3)
    ref = _make_model_seeded()
    ids = _batch(77, rows=2, seq=24)
    labels = ids.clone()
    labels[1, :20] = -100
    logits = ref(input_ids=ids).logits.float()
    shifted = shift_labels(labels)
    per_row = [torch.nn.functional.cross_entropy(logits[r], shifted[r], ignore_index=-100, reduction='sum')
               for r in range(2)]
    counts = [(shifted[r] != -100).sum() for r in range(2)]
    if mode == 'micro_batch':
        loss = (per_row[0] / counts[0] + per_row[1] / counts[1]) / 2
    else:
        loss = (per_row[0] + per_row[1]) / (counts[0] + counts[1])
    loss.backward()
    with torch.no_grad():
        for p in ref.parameters():
            p -= 0.1 * p.grad
    for n, p in ref.named_parameters():
        assert torch.allclose(results[0][n], p.detach(), atol=2e-5, rtol=1e-4), (mode, n)


def _make_model_seeded():
    from llm_training_b200.models import Llama, LlamaConfig
    return Llama(LlamaConfig(**TINY))


def _hf_causal_lm_worker(rank: int, world: int, reshard: bool):
    from transformers import LlamaConfig as HFLlamaConfig
    from llm_training_b200.models import HFCausalLM, HFCausalLMConfig
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import init_mesh
    mesh = init_mesh(world, 1)
    torch.manual_seed(0)
    hf_cfg = HFLlamaConfig(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                           num_attention_heads=4, num_key_value_heads=2, attn_implementation='sdpa')
    model = HFCausalLM(HFCausalLMConfig(hf_config=hf_cfg))
    names = [n for n, _ in model.sharding_units()]
    model.parallelize(dp_mesh=mesh.dp, tp_mesh=None, reshard_after_forward=reshard)
    engine = model._flat_engine
    assert engine.reshard == reshard
    opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2), [engine], max_grad_norm=1.0)
    losses = []
    for step in range(3):
        ids = _batch(300 + step)[rank::world]
        opt.zero_grad()
        logits = model(input_ids=ids, attention_mask=torch.ones_like(ids)).logits.float()
        loss = torch.nn.functional.cross_entropy(logits[:, :-1].flatten(0, 1), ids[:, 1:].flatten())
        loss.backward()
        opt.step()
        losses.append(loss.item())
    with engine.unsharded():
        params = {n: p.detach().clone() for n, p in model.named_parameters()}
    return params, losses, names


@pytest.mark.parametrize('reshard', [False, True])
def test_hf_causal_lm_wrapper_trains_sharded(reshard):
    """Any Hugging Face causal LM through the same engine (reference: models/hf_causal_lm/hf_causal_lm.py:88-114):
    one sharding unit per `_no_split_modules` block plus the leftovers; ZeRO-2 and ZeRO-3 agree with a plain
    single-process AdamW run on the HF module."""
    from transformers import LlamaConfig as HFLlamaConfig, LlamaForCausalLM
    results = run_distributed(_hf_causal_lm_worker, world=2, args=(reshard,), timeout=600)
    params, losses, names = results[0]
    assert sum('layers.' in n for n in names) == 2 and len(names) >= 4        # decoder blocks + embed / norm / head
    torch.manual_seed(0)
    ref = LlamaForCausalLM(HFLlamaConfig(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                                         num_attention_heads=4, num_key_value_heads=2, attn_implementation='sdpa'))
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2)
    for step in range(3):
        ids = _batch(300 + step)
        ropt.zero_grad()
        loss = 0
        for r in range(2):
            part = ids[r::2]
            logits = ref(input_ids=part, attention_mask=torch.ones_like(part)).logits.float()
            loss = loss + torch.nn.functional.cross_entropy(logits[:, :-1].flatten(0, 1), part[:, 1:].flatten()) / 2
        loss.backward()
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        ropt.step()
    for n, p in ref.named_parameters():
        assert torch.allclose(params['hf_model.' + n], p.detach(), atol=3e-5, rtol=1e-4), n


def _offload_worker(rank: int, world: int, offload: bool):
    from llm_training_b200.lightning import FSDP2Strategy, Trainer, seed_everything
    from llm_training_b200.lightning.strategy.precision import CPUOffloadPolicy
    from llm_training_b200.lms import CLM, CLMConfig
    seed_everything(11)
    lm = CLM(CLMConfig.model_validate(dict(
        init_weights=True, load_weights=False,
        model=dict(model_class='llm_training.models.Llama', model_config=TINY),
        optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=5e-3, weight_decay=0.1)))))
    strategy = FSDP2Strategy(data_parallel_size=world, offload_policy=CPUOffloadPolicy() if offload else None)
    trainer = Trainer(strategy=strategy, precision='bf16-true', logger=False, enable_progress_bar=False,
                      enable_checkpointing=False, accelerator='cpu', gradient_clip_val=1.0)
    lm.trainer = trainer
    trainer.lightning_module = lm
    trainer.strategy.setup_environment(trainer)
    with trainer.strategy.model_init_context():
        lm.configure_model()
    trainer.strategy.setup(trainer)
    opt = trainer.optimizers[0]
    assert opt.offload == offload and all(st['master'].device.type == 'cpu' for st in opt.unit_state)
    losses = []
    for step in range(4):
        ids = _batch(500 + step)[rank::world]
        batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids),
                     position_ids=torch.arange(ids.shape[1])[None].expand(ids.shape[0], -1), labels=ids)
        loss = trainer.strategy.training_step(batch, 0)
        trainer.strategy.backward(loss)
        trainer.strategy.optimizer_step(opt)
        opt.zero_grad()
        losses.append(float(loss.detach()))
    return {n: p.detach().float().clone() for n, p in lm.model.named_parameters()}, losses


def test_cpu_offload_policy_moves_optimizer_state_without_changing_the_update():
    """`CPUOffloadPolicy` (reference: fsdp2_strategy.py:58) selects the host-resident optimizer state; the update is
    the same AdamW (bf16 weights, fp32 masters)."""
    plain = run_distributed(_offload_worker, world=2, args=(False,))
    off = run_distributed(_offload_worker, world=2, args=(True,))
    assert plain[0][1] == off[0][1]
    for n, p in plain[0][0].items():
        assert torch.equal(off[0][0][n], p), n


def _groups(model, opt_cls, **kw):
    """Two param groups the usual way: matrices decay, norm weights / embeddings do not and get another lr."""
    decay = [p for p in model.parameters() if p.dim() == 2 and p.shape[0] != TINY['vocab_size']]
    rest = [p for p in model.parameters() if not (p.dim() == 2 and p.shape[0] != TINY['vocab_size'])]
    if opt_cls is torch.optim.SGD:
        return opt_cls([{'params': decay, 'weight_decay': 0.1}, {'params': rest, 'weight_decay': 0.0, 'lr': 3e-2}],
                       lr=1e-2, momentum=0.9, **kw)
    return opt_cls([{'params': decay, 'weight_decay': 0.1}, {'params': rest, 'weight_decay': 0.0, 'lr': 3e-3}],
                   lr=1e-2, **kw)


def _param_group_worker(rank: int, world: int, opt_name: str, steps: int):
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import apply_fully_sharded, init_mesh
    mesh = init_mesh(world, 1)
    model = _make_model()
    engine = apply_fully_sharded(model, mesh.dp, reshard_after_forward=False)
    opt = ShardedOptimizer(_groups(model, getattr(torch.optim, opt_name)), [engine], max_grad_norm=1.0)
    assert opt.fused == (opt_name == 'AdamW')
    assert any(len(opt._segments(u)) > 1 for u in engine.units)      # a unit really spans both groups
    for step in range(steps):
        opt.zero_grad()
        _loss(model, _batch(100 * step)[rank::world]).backward()
        opt.step()
    return {n: p.detach().clone() for n, p in model.named_parameters()}


@pytest.mark.parametrize('opt_name', ['AdamW', 'SGD'])
def test_param_groups_keep_their_own_hyperparameters(opt_name):
    """Per-group lr / weight_decay (no-decay groups) survive the flat-shard optimizer, on the fused AdamW path and
    on the generic path (round-1 advisor finding: every slice silently used group 0's values)."""
    steps = 3
    results = run_distributed(_param_group_worker, world=2, args=(opt_name, steps))
    ref = _make_model()
    ropt = _groups(ref, getattr(torch.optim, opt_name))
    for step in range(steps):
        ropt.zero_grad()
        ids = _batch(100 * step)
        (sum(_loss(ref, ids[r::2]) for r in range(2)) / 2).backward()
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        ropt.step()
    for rank in range(2):
        for n, p in ref.named_parameters():
            assert torch.allclose(results[rank][n], p.detach(), atol=2e-5, rtol=1e-4), n


def _dpo_worker(rank: int, world: int, tmpdir: str, reshard: bool):
    """DPO with both models on the same mesh: the frozen reference model gets a parameter-only ZeRO-3 engine."""
    from llm_training_b200.lightning import FSDP2Strategy, Trainer, seed_everything
    from llm_training_b200.lms import DPO, DPOConfig
    seed_everything(7)
    lm = DPO(DPOConfig.model_validate(dict(
        init_weights=True, load_weights=False,
        model=dict(model_class='llm_training.models.Llama', model_config=TINY),
        optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=1e-2)))))
    g = torch.Generator().manual_seed(3)
    rows = []
    for i in range(8):
        c, r = torch.randint(0, 128, (2, 12), generator=g), torch.randint(0, 128, (2, 12), generator=g)
        rows.append({'chosen_input_ids': c[0], 'chosen_labels': c[0].clone(), 'chosen_attention_mask': torch.ones(12, dtype=torch.long),
                     'chosen_position_ids': torch.arange(12), 'rejected_input_ids': r[0], 'rejected_labels': r[0].clone(),
                     'rejected_attention_mask': torch.ones(12, dtype=torch.long), 'rejected_position_ids': torch.arange(12)})
    loader = torch.utils.data.DataLoader(rows[rank::world], batch_size=2)
    trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=world, reshard_after_forward=reshard),
                      precision='32-true', logger=False, enable_progress_bar=False, max_steps=2, accelerator='cpu',
                      default_root_dir=tmpdir, enable_checkpointing=False, use_distributed_sampler=False)
    trainer.fit(lm, train_dataloaders=loader)
    frozen = trainer.strategy.frozen_engines
    info = {'n_frozen': len(frozen), 'n_trainable': len(trainer.strategy.engines)}
    if frozen:
        e = frozen[0]
        info['frozen_units_sharded'] = all(u.is_sharded and not u.with_grad and u.flat_grad.numel() == 0
                                           for u in e.units)
        info['ref_shard_numel'] = sum(u.shard_numel for u in e.units)
        info['ref_numel'] = sum(u.padded_numel for u in e.units)
    from llm_training_b200.lightning.checkpoint_io import module_state_dict
    sd = module_state_dict(trainer.strategy, balanced=False)
    return info, {k: v for k, v in sd.items() if k.startswith(('ref_model.', 'model.'))}


def test_dpo_reference_model_is_sharded_with_the_policy(tmp_path):
    """ZeRO-3: the frozen reference model is sharded too (parameter-only engine, no gradients / optimizer state) and
    the run is bit-identical to the resident run, where the frozen model stays whole (reference: both models are
    parallelised with the same mesh, lms/dpo/dpo.py:69-71)."""
    z3 = run_distributed(_dpo_worker, world=2, args=(str(tmp_path / 'z3'), True))
    res = run_distributed(_dpo_worker, world=2, args=(str(tmp_path / 'res'), False))
    info = z3[0][0]
    assert info['n_frozen'] == 1 and info['n_trainable'] == 1 and info['frozen_units_sharded']
    assert info['ref_shard_numel'] * 2 == info['ref_numel']
    assert res[0][0]['n_frozen'] == 0
    for k, v in res[0][1].items():
        assert torch.equal(z3[0][1][k], v), k
    ref_keys = [k for k in z3[0][1] if k.startswith('ref_model.')]
    assert ref_keys and all(z3[0][1][k].numel() > 0 for k in ref_keys)


def _uneven_worker(rank: int, world: int, reshard: bool):
    """Rank 1 never uses decoder layer 0 (its parameters get no gradient there); rank 0 uses everything."""
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import apply_fully_sharded, init_mesh
    mesh = init_mesh(world, 1)
    model = _make_model()
    if rank == 1:
        model.layers[0].forward = lambda hidden_states, ctx: hidden_states
    engine = apply_fully_sharded(model, mesh.dp, reshard_after_forward=reshard)
    opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.0), [engine],
                           max_grad_norm=1.0)
    for step in range(2):
        opt.zero_grad()
        _loss(model, _batch(step)[rank::world]).backward()
        opt.step()
    with engine.unsharded():
        params = {n: p.detach().clone() for n, p in model.named_parameters()}
    return params, float(opt.last_grad_norm)


@pytest.mark.parametrize('reshard', [False, True])
def test_ranks_with_unused_parameters_still_reduce_in_one_order(reshard):
    """The engine issues gradient reductions in a fixed order on every rank.  Before, a rank whose unit received no
    gradient reduced it last while its peers reduced it in backward order: the collectives paired up different units
    (mismatched sizes on gloo / NCCL, silently wrong data on the epoch-paired NVLink kernels)."""
    results = run_distributed(_uneven_worker, world=2, args=(reshard,))
    (p0, n0), (p1, n1) = results
    assert abs(n0 - n1) < 1e-6                      # one global gradient norm
    for name in p0:
        assert torch.equal(p0[name], p1[name]), name      # every rank holds the same updated parameters
