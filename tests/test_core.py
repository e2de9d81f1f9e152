"""Config loader / CLI, LR schedules, RoPE initialisers, op references, objectives, checkpoint + HF export."""
import math
import os

import pytest
import torch

from tests.helpers import TINY_LLAMA


# ---------------------------------------------------------------------------------------------------- config / CLI
def test_alias_package_shares_module_objects():
    import llm_training
    import llm_training.lms
    import llm_training_b200.lms
    from llm_training.models import Llama
    from llm_training_b200.models import Llama as Llama2
    assert llm_training.lms is llm_training_b200.lms and Llama is Llama2


def test_yaml_loader_dotted_keys_overrides_and_instantiation(tmp_path):
    from llm_training_b200.lightning import FSDP2Strategy, LearningRateMonitor
    from llm_training_b200.lightning.cli.config import (instantiate, instantiate_nested, load_config, parse_overrides,
                                                        set_by_path)
    cfg_file = tmp_path / 'c.yaml'
    cfg_file.write_text('''
trainer:
  strategy:
    class_path: llm_training.lightning.FSDP2Strategy
    init_args:
      tensor_parallel_size: 2
  callbacks:
    - class_path: LearningRateMonitor
model:
  class_path: llm_training.lms.CLM
  init_args.config:
    model:
      model_class: llm_training.models.Llama
      model_config: {vocab_size: 64, hidden_size: 32, intermediate_size: 64, num_hidden_layers: 1,
                     num_attention_heads: 2, num_key_value_heads: 1}
    optim:
      optimizer_class: deepspeed.ops.adam.FusedAdam
      optimizer_kwargs: {lr: 1e-5}
''')
    configs, overrides, positional = parse_overrides(['fit', '--config', str(cfg_file), '--trainer.num_nodes', '2',
                                                      '--model.init_args.config.optim.optimizer_kwargs.lr=3e-4'])
    assert positional == ['fit'] and overrides['trainer.num_nodes'] == 2
    cfg = load_config(configs)
    for k, v in overrides.items():
        set_by_path(cfg, k, v)
    assert cfg['model']['init_args']['config']['optim']['optimizer_kwargs']['lr'] == 3e-4      # '1e-5' style floats
    assert cfg['trainer']['num_nodes'] == 2
    trainer_kwargs = instantiate_nested(cfg['trainer'])
    assert isinstance(trainer_kwargs['strategy'], FSDP2Strategy)
    assert isinstance(trainer_kwargs['callbacks'][0], LearningRateMonitor)          # short name resolution
    lm = instantiate(cfg['model'])
    from llm_training_b200.lms import CLM, ModelProvider
    from llm_training_b200.optim import FusedAdamW
    assert isinstance(lm, CLM) and isinstance(lm.config.model, ModelProvider)
    assert lm.config.optim.optimizer_class is FusedAdamW                            # DeepSpeed class aliased
    assert lm.config.model.model_config.hidden_size == 32


def test_cli_fit_runs_and_checkpoint_embeds_config(tmp_path):
    from llm_training_b200.lightning.cli import LightningCLI
    cfg_file = tmp_path / 'fit.yaml'
    cfg_file.write_text(f'''
seed_everything: 1
trainer:
  precision: 32-true
  accelerator: cpu
  logger:
    class_path: llm_training.lightning.CSVLogger
    init_args: {{save_dir: {tmp_path}/logs}}
  max_epochs: 1
  enable_progress_bar: false
  callbacks:
    - class_path: llm_training.lightning.ModelCheckpoint
      init_args: {{save_on_train_epoch_end: true, save_top_k: -1}}
model:
  class_path: llm_training.lms.CLM
  init_args.config:
    init_weights: true
    load_weights: false
    model:
      model_class: llm_training.models.Llama
      model_config: {dict(TINY_LLAMA)}
    optim:
      optimizer_class: torch.optim.AdamW
      optimizer_kwargs: {{lr: 1e-3}}
data:
  class_path: llm_training.data.DummyDataModule
  init_args.config: {{batch_size: 2, vocab_size: 256, max_length: 16, num_samples: 8, base_seed: 1, pin_memory: false}}
''')
    cli = LightningCLI(['fit', '--config', str(cfg_file)])
    assert cli.trainer.global_step == 4
    ckpt = os.path.join(cli.trainer.log_dir, 'checkpoints', 'epoch=0-step=4.ckpt')
    assert os.path.exists(ckpt)
    from llm_training_b200.lightning.checkpoint_io import read_checkpoint_state_dict
    state, meta = read_checkpoint_state_dict(ckpt)
    assert meta['config']['model']['class_path'] == 'llm_training.lms.CLM' and meta['config']['world_size'] == 1
    assert 'model.layers.0.self_attn.q_proj.weight' in state and 'consumed_samples.n' in state
    assert int(state['consumed_samples.n']) == 8


# ---------------------------------------------------------------------------------------------------- schedules
def _lrs(cls, steps=12, **kw):
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    sched = cls(opt, **kw)
    out = []
    for _ in range(steps):
        out.append(opt.param_groups[0]['lr'])
        opt.step()
        sched.step()
    return out, sched, opt


def test_lr_schedules_values_and_resume():
    from llm_training_b200.lr_schedulers import ConstantWarmupLR, CosineAnnealingWarmupLR, LinearWarmupLR
    cos, _, _ = _lrs(CosineAnnealingWarmupLR, num_warmup_steps=3, num_total_steps=10, min_lr=0.1)
    assert cos[:3] == pytest.approx([1 / 3, 2 / 3, 1.0]) and cos[3] == pytest.approx(1.0)
    assert cos[10] == pytest.approx(0.1) and all(a >= b - 1e-12 for a, b in zip(cos[3:10], cos[4:11]))
    assert cos[6] == pytest.approx(0.1 + 0.9 * (1 + math.cos(math.pi * 3 / 7)) / 2)
    const, _, _ = _lrs(ConstantWarmupLR, num_warmup_steps=4)
    assert const[:4] == pytest.approx([0.25, 0.5, 0.75, 1.0]) and set(const[4:]) == {1.0}
    lin, _, _ = _lrs(LinearWarmupLR, num_warmup_steps=3, num_total_steps=10, min_lr=0.1)
    assert lin[:4] == pytest.approx([0.25, 0.5, 0.75, 1.0]) and lin[10] == pytest.approx(0.1)
    # state_dict round trip continues the same curve
    first, sched, opt = _lrs(CosineAnnealingWarmupLR, steps=5, num_warmup_steps=3, num_total_steps=10, min_lr=0.1)
    state = sched.state_dict()
    p = torch.nn.Parameter(torch.zeros(1))
    opt2 = torch.optim.SGD([p], lr=1.0)
    sched2 = CosineAnnealingWarmupLR(opt2, num_warmup_steps=3, num_total_steps=10, min_lr=0.1)
    sched2.load_state_dict(state)
    opt2.param_groups[0]['lr'] = opt.param_groups[0]['lr']
    rest = []
    for _ in range(5):
        rest.append(opt2.param_groups[0]['lr'])
        opt2.step()
        sched2.step()
    assert first + rest == pytest.approx(cos[:10])


# ---------------------------------------------------------------------------------------------------- rope
@pytest.mark.parametrize('kind,scaling', [
    ('default', None), ('linear', {'factor': 4.0}), ('dynamic', {'factor': 2.0}),
    ('yarn', {'factor': 8.0}), ('longrope', {'factor': 4.0, 'short_factor': [1.0] * 16, 'long_factor': [2.0] * 16}),
    ('llama3', {'factor': 8.0, 'low_freq_factor': 1.0, 'high_freq_factor': 4.0,
                'original_max_position_embeddings': 64})])
def test_rope_initialisers(kind, scaling):
    from llm_training_b200.ops import ROPE_INIT_FUNCTIONS, RoPEConfig
    cfg = RoPEConfig(type=kind, base=10000.0, dim=32, max_position_embeddings=256, scaling_config=scaling)
    inv, factor = ROPE_INIT_FUNCTIONS[kind](cfg, None, 1024)
    base = 1.0 / (10000.0 ** (torch.arange(0, 32, 2).float() / 32))
    assert inv.shape == (16,) and torch.isfinite(inv).all() and factor > 0
    if kind == 'default':
        assert torch.allclose(inv, base)
    if kind == 'linear':
        assert torch.allclose(inv, base / 4)
    if kind == 'longrope':
        assert torch.allclose(inv, base / 2) and factor == pytest.approx(math.sqrt(1 + math.log(4) / math.log(256)))
    if kind == 'llama3':
        assert torch.allclose(inv[0], base[0]) and torch.allclose(inv[-1], base[-1] / 8)    # high freq kept, low / 8
    if kind == 'yarn':
        assert factor == pytest.approx(0.1 * math.log(8) + 1) and (inv <= base + 1e-9).all()
    if kind == 'dynamic':
        assert (inv <= base + 1e-9).all() and inv[0] == pytest.approx(1.0)
    with pytest.raises(KeyError):
        RoPEConfig(type='linear', base=1e4, dim=32, max_position_embeddings=8, scaling_config={})


# ---------------------------------------------------------------------------------------------------- ops (CPU)
def test_shift_labels_and_segment_bounds_and_packed_attention_reference():
    from llm_training_b200.ops import prepare_4d_causal_attention_mask, segment_bounds, shift_labels
    from llm_training_b200.ops.reference import attention_ref
    labels = torch.tensor([[1, 2, 3], [4, 5, 6]])
    assert shift_labels(labels).tolist() == [[2, 3, -100], [5, 6, -100]]
    mask = torch.tensor([[1, 1, 2, 2, 2, 0], [1, 1, 1, 1, 1, 1]])
    start, end = segment_bounds(mask, 2, 6, torch.device('cpu'))
    assert start.tolist() == [0, 0, 2, 2, 2, 5, 6, 6, 6, 6, 6, 6]
    assert end.tolist() == [2, 2, 5, 5, 5, 6, 12, 12, 12, 12, 12, 12]
    # packed reference == eager attention with the 4-D block-diagonal mask, per batch row
    torch.manual_seed(0)
    hq, hkv, d = 4, 2, 8
    q, k, v = torch.randn(12, hq, d), torch.randn(12, hkv, d), torch.randn(12, hkv, d)
    packed = attention_ref(q, k, v, start)
    add = prepare_4d_causal_attention_mask(mask, (2, 6), torch.zeros(1), 0)
    for b in range(2):
        qb = q[b * 6:(b + 1) * 6].transpose(0, 1)
        kb = k[b * 6:(b + 1) * 6].transpose(0, 1).repeat_interleave(2, 0)
        vb = v[b * 6:(b + 1) * 6].transpose(0, 1).repeat_interleave(2, 0)
        w = torch.softmax(qb @ kb.transpose(1, 2) / math.sqrt(d) + add[b], dim=-1)
        eager = (w @ vb).transpose(0, 1)
        valid = mask[b] != 0
        assert torch.allclose(packed[b * 6:(b + 1) * 6][valid], eager[valid], atol=1e-5)


def test_fused_loss_functions_cpu_paths_match_reference_math():
    from llm_training_b200.ops import cross_entropy, fused_linear_cross_entropy, fused_linear_logprob
    torch.manual_seed(0)
    h, w = torch.randn(10, 16), torch.randn(32, 16)
    labels = torch.randint(0, 32, (10,))
    labels[3] = -100
    logits = h @ w.t()
    ref = torch.nn.functional.cross_entropy(logits, labels, ignore_index=-100)
    assert fused_linear_cross_entropy(h, w, labels) == pytest.approx(ref.item(), rel=1e-5)
    assert cross_entropy(logits.view(2, 5, 32), labels.view(2, 5)) == pytest.approx(ref.item(), rel=1e-5)
    lp = fused_linear_logprob(h, w, labels)
    manual = logits.log_softmax(-1).gather(1, labels.clamp_min(0)[:, None])[:, 0] * (labels != -100)
    assert torch.allclose(lp, manual, atol=1e-5)


# ---------------------------------------------------------------------------------------------------- objectives
def _preference_batch(vocab=256, rows=2):
    g = torch.Generator().manual_seed(0)
    batch = {}
    for side, n in (('chosen', 12), ('rejected', 9)):
        ids = torch.randint(1, vocab, (rows, n), generator=g)
        labels = ids.clone()
        labels[:, :4] = -100
        batch |= {f'{side}_input_ids': ids, f'{side}_attention_mask': torch.ones_like(ids),
                  f'{side}_position_ids': torch.arange(n)[None], f'{side}_labels': labels}
    return batch


def _lm(kind, **extra):
    from llm_training_b200.lms import DPO, ORPO, DPOConfig, ORPOConfig
    cfg = dict(init_weights=True, load_weights=False,
               model=dict(model_class='llm_training.models.Llama', model_config=dict(TINY_LLAMA)), **extra)
    lm = DPO(DPOConfig.model_validate(cfg)) if kind == 'dpo' else ORPO(ORPOConfig.model_validate(cfg))
    torch.manual_seed(0)
    lm.configure_model()
    return lm


def test_dpo_loss_matches_manual_computation():
    lm = _lm('dpo', beta=0.2, label_smoothing=0.1)
    batch = _preference_batch()
    with torch.no_grad():                       # make policy != reference
        for p in lm.model.parameters():
            p.add_(0.01 * torch.randn_like(p))
    loss = lm.training_step(batch, 0)

    def logps(model, side):
        ids, labels = batch[f'{side}_input_ids'], batch[f'{side}_labels']
        logp = model(input_ids=ids).logits.float().log_softmax(-1)[:, :-1]
        tgt = labels[:, 1:]
        m = tgt != -100
        return (logp.gather(2, tgt.clamp_min(0)[..., None])[..., 0] * m).sum(-1)

    pi = logps(lm.model, 'chosen') - logps(lm.model, 'rejected')
    ref = logps(lm.ref_model, 'chosen') - logps(lm.ref_model, 'rejected')
    z = 0.2 * (pi - ref)
    manual = (-torch.nn.functional.logsigmoid(z) * 0.9 - torch.nn.functional.logsigmoid(-z) * 0.1).mean()
    assert loss.item() == pytest.approx(manual.item(), rel=1e-4, abs=1e-5)
    assert not any(p.requires_grad for p in lm.ref_model.parameters())
    loss.backward()
    assert all(p.grad is not None for p in lm.model.parameters())


def test_orpo_loss_matches_manual_computation():
    lm = _lm('orpo', beta=0.25)
    batch = _preference_batch()
    loss = lm.training_step(batch, 0)

    def mean_logps(side):
        ids, labels = batch[f'{side}_input_ids'], batch[f'{side}_labels']
        logits = lm.model(input_ids=ids).logits.float()
        logp = logits.log_softmax(-1)[:, :-1]
        tgt = labels[:, 1:]
        m = tgt != -100
        per = (logp.gather(2, tgt.clamp_min(0)[..., None])[..., 0] * m).sum(-1) / m.sum(-1)
        return per, logits, tgt

    c, c_logits, c_tgt = mean_logps('chosen')
    r, _, _ = mean_logps('rejected')
    log_odds = (c - r) - (torch.log1p(-torch.exp(c)) - torch.log1p(-torch.exp(r)))
    or_loss = -(0.25 * torch.nn.functional.logsigmoid(log_odds)).mean()
    ce = torch.nn.functional.cross_entropy(c_logits[:, :-1].flatten(0, 1), c_tgt.flatten(), ignore_index=-100)
    assert loss.item() == pytest.approx((or_loss + ce).item(), rel=1e-4)


def test_clm_neftune_and_frozen_modules():
    from llm_training_b200.lms import CLM, CLMConfig
    cfg = CLMConfig.model_validate(dict(init_weights=True, load_weights=False, neftune_alpha=5.0,
                                        frozen_modules=[r'layers\.0\.'],
                                        model=dict(model_class='llm_training.models.Llama',
                                                   model_config=dict(TINY_LLAMA))))
    lm = CLM(cfg)
    lm.configure_model()
    assert lm.model.neftune_alpha == 5.0
    assert not any(p.requires_grad for p in lm.model.layers[0].parameters())
    assert all(p.requires_grad for p in lm.model.layers[1].parameters())
    ids = torch.randint(0, 256, (2, 10))
    batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids), labels=ids)
    lm.train()
    a = lm.training_step(batch, 0).item()
    b = lm.training_step(batch, 0).item()
    assert a != b                                   # fresh noise every step
    lm.eval()
    assert lm._forward_loss(batch)[0].item() == lm._forward_loss(batch)[0].item()
    assert 'model.lm_head.weight' in lm.required_keys


# ---------------------------------------------------------------------------------------------------- HF interop
def test_hf_state_dict_round_trip_and_logit_parity_with_transformers():
    transformers = pytest.importorskip('transformers')
    from llm_training_b200.models import Llama, LlamaConfig
    torch.manual_seed(0)
    model = Llama(LlamaConfig(**TINY_LLAMA, rope_theta=10000.0, rms_norm_eps=1e-5)).eval()
    sd = model.state_dict()
    hf_sd = model.convert_state_dict_to_hf(sd)
    assert 'model.layers.0.self_attn.q_proj.weight' in hf_sd and 'lm_head.weight' in hf_sd
    assert set(model.convert_state_dict_from_hf(hf_sd)) == set(sd)
    hf = model.get_hf_model().eval()
    assert isinstance(hf, transformers.LlamaForCausalLM)
    ids = torch.randint(0, 256, (2, 12))
    with torch.no_grad():
        ours = model(input_ids=ids).logits
        theirs = hf(input_ids=ids).logits
    assert torch.allclose(ours, theirs, atol=2e-4), (ours - theirs).abs().max()


def test_convert_to_hf_from_dcp_checkpoint(tmp_path):
    pytest.importorskip('transformers')
    from llm_training_b200.lightning.cli import LightningCLI
    cfg_file = tmp_path / 'fit.yaml'
    cfg_file.write_text(f'''
trainer:
  strategy:
    class_path: llm_training.lightning.FSDP2Strategy
    init_args: {{data_parallel_size: 1}}
  precision: 32-true
  accelerator: cpu
  logger:
    class_path: llm_training.lightning.CSVLogger
    init_args: {{save_dir: {tmp_path}/logs}}
  max_epochs: 1
  enable_progress_bar: false
  callbacks:
    - class_path: llm_training.lightning.ModelCheckpoint
      init_args: {{save_on_train_epoch_end: true}}
model:
  class_path: llm_training.lms.CLM
  init_args.config:
    init_weights: true
    load_weights: false
    model:
      model_class: llm_training.models.Llama
      model_config: {dict(TINY_LLAMA)}
    optim:
      optimizer_class: torch.optim.AdamW
      optimizer_kwargs: {{lr: 1e-3}}
data:
  class_path: llm_training.data.DummyDataModule
  init_args.config: {{batch_size: 2, vocab_size: 256, max_length: 16, num_samples: 4, base_seed: 1, pin_memory: false}}
''')
    cli = LightningCLI(['fit', '--config', str(cfg_file)])
    ckpt = os.path.join(cli.trainer.log_dir, 'checkpoints', 'epoch=0-step=2.ckpt')
    assert os.path.isdir(ckpt)                                   # sharded layout: DCP directory + meta.pt
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        'convert_to_hf', os.path.join(os.path.dirname(os.path.dirname(__file__)), 'scripts', 'convert_to_hf.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = mod.convert(ckpt, str(tmp_path / 'hf'))
    from transformers import AutoModelForCausalLM
    hf = AutoModelForCausalLM.from_pretrained(out)
    trained = cli.model.model
    assert torch.allclose(hf.model.layers[1].mlp.down_proj.weight, trained.layers[1].mlp.down_proj.weight.detach())
    assert torch.allclose(hf.lm_head.weight, trained.lm_head.weight.detach())


def test_fp16_mixed_dynamic_loss_scaling_skips_overflow_steps():
    """`16-mixed`: the loss is scaled, the fused optimizer unscales inside its update, a non-finite gradient norm skips
    the step and halves the scale (reference: GradScaler path of FSDP2Precision, fsdp2_precision.py:130-163; its
    scaler is never created because of SURVEY B7)."""
    from llm_training_b200.lightning import FSDP2Strategy, Trainer
    from llm_training_b200.lms import CLM, CLMConfig
    torch.manual_seed(0)
    lm = CLM(CLMConfig.model_validate(dict(init_weights=True, load_weights=False,
        model=dict(model_class='llm_training.models.Llama', model_config=dict(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2, attn_implementation='sdpa')),
        optim=dict(optimizer_class='torch.optim.AdamW', optimizer_kwargs=dict(lr=1e-2)))))
    trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=1), precision='16-mixed', logger=False, enable_progress_bar=False, enable_checkpointing=False, gradient_clip_val=1.0, accelerator='cpu')
    lm.trainer = trainer
    trainer.lightning_module = lm
    trainer.strategy.setup_environment(trainer)
    with trainer.strategy.model_init_context():
        lm.configure_model()
    trainer.strategy.setup(trainer)
    pp = trainer.strategy.precision_plugin
    assert pp.scaler is not None and pp.scaler.get_scale() == 65536.0
    ids = torch.randint(0, 128, (2, 32))
    batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids), position_ids=torch.arange(32)[None].expand(2, -1), labels=ids)
    opt = trainer.optimizers[0]
    losses = []
    for i in range(8):
        loss = trainer.strategy.training_step(batch, 0)
        trainer.strategy.backward(loss)
        if i == 3:   # inject an overflow
            for e in trainer.strategy.engines:
                e.units[0].flat_grad[0] = float('inf')
        trainer.strategy.optimizer_step(opt)
        opt.zero_grad()
        losses.append(round(float(loss.detach()), 4))
        if i == 3:
            assert opt.last_step_skipped and pp.scaler.get_scale() == 32768.0 and pp.skipped_steps == 1
        else:
            assert not opt.last_step_skipped and float(opt.last_grad_norm) < 100      # norm is reported UNscaled
    assert losses[4] == losses[3]          # the skipped step left the weights untouched
    assert losses[-1] < losses[0]


def test_reads_reference_deepspeed_zero2_checkpoint_directory(tmp_path):
    """The directory layout the reference writes under its DeepSpeed strategy (stage <= 2) is readable without
    DeepSpeed; a ZeRO-3 directory (placeholder tensors in `module`) is rejected with a pointer to zero_to_fp32.py."""
    import os
    from llm_training_b200.lightning.checkpoint_io import read_checkpoint_state_dict
    ckpt = tmp_path / 'epoch=0-step=10.ckpt'
    os.makedirs(ckpt / 'checkpoint')
    (ckpt / 'latest').write_text('checkpoint')
    weights = {'_forward_module.model.embed_tokens.weight': torch.randn(8, 4),
               '_forward_module.model.norm.weight': torch.ones(4)}
    torch.save({'module': weights, 'config': {'trainer': {'precision': 'bf16-true'}}, 'global_step': 10},
               ckpt / 'checkpoint' / 'mp_rank_00_model_states.pt')
    state, rest = read_checkpoint_state_dict(str(ckpt))
    assert set(state) == {'model.embed_tokens.weight', 'model.norm.weight'}
    assert torch.equal(state['model.norm.weight'], torch.ones(4)) and rest['global_step'] == 10 and 'config' in rest
    z3 = tmp_path / 'z3.ckpt'
    os.makedirs(z3 / 'checkpoint')
    (z3 / 'latest').write_text('checkpoint')
    torch.save({'module': {k: torch.empty(0) for k in weights}}, z3 / 'checkpoint' / 'mp_rank_00_model_states.pt')
    with pytest.raises(ValueError, match='zero_to_fp32'):
        read_checkpoint_state_dict(str(z3))


def test_model_checkpoint_top_k_monitor_and_save_last(tmp_path):
    """`save_top_k` with and without a monitored metric, `save_last`, Lightning's `epoch=E-step=S.ckpt` naming."""
    import os
    from llm_training_b200.lightning import ModelCheckpoint

    class FakeStrategy:
        def broadcast(self, obj, src=0):
            return obj

        def barrier(self, name=None):
            pass

        def remove_checkpoint(self, path):
            if os.path.exists(path):
                os.remove(path)

    class FakeTrainer:
        def __init__(self):
            self.strategy, self.is_global_zero, self.log_dir = FakeStrategy(), True, str(tmp_path / 'logs')
            self.current_epoch, self.global_step, self.callback_metrics, self.training = 0, 0, {}, True

        def _resolve(self, v):
            return float(v)

        def save_checkpoint(self, path):
            open(path, 'w').write(str(self.global_step))

    trainer = FakeTrainer()
    plain = ModelCheckpoint(every_n_train_steps=2, save_top_k=2, save_last=True)
    plain.setup(trainer, None, 'fit')
    assert plain.dirpath == os.path.join(trainer.log_dir, 'checkpoints')
    for step in range(1, 9):
        trainer.global_step = step
        plain.on_train_batch_end(trainer, None, None, None, step - 1)
    names = sorted(os.listdir(plain.dirpath))
    assert names == ['epoch=0-step=6.ckpt', 'epoch=0-step=8.ckpt', 'last.ckpt'], names     # two newest + last
    assert open(os.path.join(plain.dirpath, 'last.ckpt')).read() == '8'

    best = ModelCheckpoint(dirpath=str(tmp_path / 'best'), monitor='val_loss', mode='min', save_top_k=2,
                           filename='{step}-{val_loss:.2f}')
    best.setup(trainer, None, 'fit')
    trainer.training = False
    for step, loss in ((10, 3.0), (20, 2.0), (30, 2.5), (40, 4.0), (50, 1.0)):
        trainer.global_step, trainer.callback_metrics = step, {'val_loss': loss}
        best.on_validation_end(trainer, None)
    assert sorted(os.listdir(best.dirpath)) == ['step=20-val_loss=2.00.ckpt', 'step=50-val_loss=1.00.ckpt']
    assert best.best_model_path.endswith('step=50-val_loss=1.00.ckpt')
    restored = ModelCheckpoint(dirpath=str(tmp_path / 'best'), monitor='val_loss', save_top_k=2)
    restored.load_state_dict(best.state_dict())
    assert restored.best_k == best.best_k and restored.best_model_path == best.best_model_path


def test_wandb_logger_directory_convention_and_local_mirror(tmp_path):
    """``save_dir == log_dir == <save_dir>/<project>/<name>`` (reference: lightning/loggers/wandb.py:58-72), created
    lazily; metrics are mirrored to ``metrics.jsonl`` so a run without network still leaves a record."""
    import json
    import os
    from llm_training_b200.lightning.loggers.wandb import WandbLogger, _NullRun
    lg = WandbLogger(name='run-a', project='proj', save_dir=str(tmp_path), experiment=_NullRun(), prefix='train/')
    assert not os.path.exists(tmp_path / 'proj')                       # nothing created before first use
    assert lg.save_dir == lg.log_dir == str(tmp_path / 'proj' / 'run-a') and os.path.isdir(lg.log_dir)
    lg.log_hyperparams({'lr': 1e-3})
    lg.log_metrics({'Loss': 2.5}, step=3)
    lg.finalize('success')
    rows = [json.loads(l) for l in open(os.path.join(lg.log_dir, 'metrics.jsonl'))]
    assert rows[0] == {'hparams': {'lr': 1e-3}} and rows[1] == {'step': 3, 'train/Loss': 2.5}


def test_hf_causal_lm_keeps_packed_documents_isolated():
    """Segment-id masks through the HF wrapper: a packed row gives every document the logits it gets alone."""
    from transformers import LlamaConfig as HFLlamaConfig
    from llm_training_b200.models import HFCausalLM, HFCausalLMConfig
    torch.manual_seed(0)
    model = HFCausalLM(HFCausalLMConfig(hf_config=HFLlamaConfig(
        vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
        num_key_value_heads=2, attn_implementation='sdpa'))).eval()
    ids = torch.randint(0, 128, (1, 12))
    seg = torch.tensor([[1] * 5 + [2] * 4 + [0] * 3])
    pos = torch.tensor([[0, 1, 2, 3, 4, 0, 1, 2, 3, 0, 0, 0]])
    with torch.no_grad():
        packed = model(input_ids=ids, attention_mask=seg, position_ids=pos).logits
        first = model(input_ids=ids[:, :5], attention_mask=torch.ones(1, 5, dtype=torch.long)).logits
        second = model(input_ids=ids[:, 5:9], attention_mask=torch.ones(1, 4, dtype=torch.long)).logits
    assert torch.allclose(packed[:, :5], first, atol=1e-5) and torch.allclose(packed[:, 5:9], second, atol=1e-5)


def test_hf_causal_lm_in_tree_kernel_patching_preserves_outputs():
    """`enable_liger_kernel: true` (reference: Liger patching of HF modules) swaps RMSNorm / SwiGLU MLP forwards for
    the in-tree ops; logits and gradients are unchanged."""
    from transformers import LlamaConfig as HFLlamaConfig
    from llm_training_b200.models import HFCausalLM, HFCausalLMConfig
    hf_cfg = HFLlamaConfig(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                           num_attention_heads=4, num_key_value_heads=2, attn_implementation='sdpa')
    torch.manual_seed(0)
    plain = HFCausalLM(HFCausalLMConfig(hf_config=hf_cfg))
    torch.manual_seed(0)
    patched = HFCausalLM(HFCausalLMConfig(hf_config=hf_cfg, enable_liger_kernel=True))
    assert patched.patched_modules == 2 * 2 + 1 + 2          # two norms + one MLP per layer, the final norm
    patched.load_state_dict(plain.state_dict())
    ids = torch.randint(0, 128, (2, 10))
    a, b = plain(input_ids=ids).logits, patched(input_ids=ids).logits
    assert torch.allclose(a, b, atol=1e-5)
    a.float().pow(2).mean().backward()
    b.float().pow(2).mean().backward()
    for (n, p), (_, q) in zip(plain.named_parameters(), patched.named_parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5), n


def test_precision_modes_route_to_kernels_or_refuse_on_gpu(monkeypatch):
    """bf16-mixed shares the bf16 kernel path (bf16 compute copy + fp32 masters); modes without native kernels raise
    on a GPU instead of silently running PyTorch ops, unless explicitly allowed.  (Device type is passed explicitly,
    so this runs without a GPU.)"""
    import pytest
    from llm_training_b200.lightning.strategy.precision import FSDP2Precision
    for mode in ('bf16-true', 'bf16-mixed'):
        plugin = FSDP2Precision(mode, device='cuda')
        assert plugin.native and plugin.param_dtype == torch.bfloat16 and plugin.autocast_dtype is None
        assert plugin.mp_policy.reduce_dtype == torch.bfloat16
    for mode in ('32-true', '16-true'):
        monkeypatch.delenv('LLMT_ALLOW_TORCH_PRECISION', raising=False)
        with pytest.raises(NotImplementedError, match='no sm_100a kernel path'):
            FSDP2Precision(mode, device='cuda').check_supported()
        monkeypatch.setenv('LLMT_ALLOW_TORCH_PRECISION', '1')
        allowed = FSDP2Precision(mode, device='cuda')
        allowed.check_supported()
        assert allowed.native is False
    # CPU keeps the reference's semantics for every mode (fp32 parameters + autocast for the mixed ones)
    cpu = FSDP2Precision('bf16-mixed', device='cpu')
    assert cpu.param_dtype == torch.float32 and cpu.autocast_dtype == torch.bfloat16
    assert cpu.mp_policy.param_dtype == torch.bfloat16 and cpu.mp_policy.reduce_dtype == torch.bfloat16


def test_strategy_arguments_are_honoured_or_rejected_never_dropped(monkeypatch):
    import pytest
    from llm_training_b200.lightning import DeepSpeedStrategy, FSDP2Strategy
    from llm_training_b200.lightning.strategy.precision import FSDP2Precision, MixedPrecisionPolicy
    ok = FSDP2Strategy(mp_policy=MixedPrecisionPolicy(param_dtype=torch.float32, reduce_dtype=torch.float32))
    ok.precision_plugin = FSDP2Precision('32-true', device='cpu')
    assert ok.mp_policy.reduce_dtype == torch.float32
    bad = FSDP2Strategy(mp_policy=MixedPrecisionPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.float32))
    bad.precision_plugin = FSDP2Precision('bf16-true', device='cpu')
    with pytest.raises(ValueError, match='reduce_dtype'):
        bad.mp_policy
    monkeypatch.delenv('LLMT_IGNORE_UNSUPPORTED', raising=False)
    with pytest.raises(NotImplementedError, match='offload_parameters'):
        DeepSpeedStrategy(stage=3, offload_parameters=True)
    monkeypatch.setenv('LLMT_IGNORE_UNSUPPORTED', '1')
    assert DeepSpeedStrategy(stage=3, offload_parameters=True).stage == 3
    from llm_training_b200.parallel.fsdp import _resolve_reshard
    from llm_training_b200.parallel.mesh import MeshAxis
    two = MeshAxis('data_parallel', [0, 1], 0, None)
    cpu = torch.device('cpu')
    assert _resolve_reshard(2, [], two, cpu) is True and _resolve_reshard(1, [], two, cpu) is False


def test_optimizer_without_master_weights_steps_the_low_precision_slices():
    """FSDP2Strategy(use_master_weights=False): no fp32 copy; the user's optimizer class updates the bf16 slices."""
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import apply_fully_sharded
    torch.manual_seed(0)
    model = torch.nn.Linear(16, 16, bias=False).to(torch.bfloat16)
    ref = torch.nn.Linear(16, 16, bias=False).to(torch.bfloat16)
    ref.load_state_dict(model.state_dict())
    engine = apply_fully_sharded(model, None)
    opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-2), [engine], master_weights=False)
    assert not opt.fused and all('master' not in st for st in opt.unit_state)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2)
    x = torch.randn(4, 16, dtype=torch.bfloat16)
    for m, o in ((model, opt), (ref, ropt)):
        o.zero_grad()
        m(x).float().pow(2).mean().backward()
        o.step()
    assert torch.equal(model.weight, ref.weight)


def test_wandb_logger_uploads_the_source_tree(tmp_path):
    """SaveConfigCallback hands the code to the logger like the reference's `experiment.log_code`
    (save_config_callback.py:39-49): python / shell / template / kernel sources only."""
    import os
    from llm_training_b200.lightning.loggers.wandb import WandbLogger, _NullRun
    lg = WandbLogger(name='run-b', project='proj', save_dir=str(tmp_path), experiment=_NullRun())
    files = lg.log_code()
    assert 'llm_training_b200/lms/clm/clm.py' in files and 'csrc/kernels/gemm.cu' in files
    assert 'scripts/train.sh' in files and not any(f.endswith(('.so', '.md', '.yaml')) for f in files)
    assert os.path.exists(os.path.join(lg.log_dir, 'code_files.txt'))


def test_hf_causal_lm_packed_attention_hook_matches_dense_mask_path():
    """`attn_implementation: flash_attention_2` routes HF's attention layers to the in-tree packed attention through
    HF's AttentionInterface (per-token document bounds, no [B,1,T,T] mask; reference: the `_get_unpad_data` patch,
    hf_causal_lm.py:14-20).  Same logits and gradients as the sdpa path with the block-diagonal 4-D mask."""
    from transformers import LlamaConfig as HFLlamaConfig
    from llm_training_b200.models import HFCausalLM, HFCausalLMConfig
    hf_cfg = dict(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                  num_key_value_heads=2)
    torch.manual_seed(0)
    dense = HFCausalLM(HFCausalLMConfig(hf_config=HFLlamaConfig(**hf_cfg), attn_implementation='sdpa'))
    hooked = HFCausalLM(HFCausalLMConfig(hf_config=HFLlamaConfig(**hf_cfg), attn_implementation='flash_attention_2'))
    assert hooked.uses_packed_attention and not dense.uses_packed_attention
    assert hooked.hf_model.config._attn_implementation == 'llmt_packed'
    hooked.load_state_dict(dense.state_dict())
    ids = torch.randint(0, 128, (2, 12))
    seg = torch.tensor([[1] * 5 + [2] * 4 + [0] * 3, [1] * 12])
    pos = torch.tensor([[0, 1, 2, 3, 4, 0, 1, 2, 3, 0, 0, 0], list(range(12))])
    a = dense(input_ids=ids, attention_mask=seg, position_ids=pos).logits
    b = hooked(input_ids=ids, attention_mask=seg, position_ids=pos).logits
    keep = seg > 0                                            # pad rows are ignored by every loss
    assert torch.allclose(a[keep], b[keep], atol=1e-5)
    (a[keep].float().pow(2).mean()).backward()
    (b[keep].float().pow(2).mean()).backward()
    for (n, p), (_, q) in zip(dense.named_parameters(), hooked.named_parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5), n


def test_deterministic_switches_reach_the_attention_op(monkeypatch):
    """`Trainer(deterministic=True)` sets torch's global switch (Lightning semantics); the attention op reads it, and
    the reference's own `FLASH_ATTENTION_DETERMINISTIC=1` (ops/attention_op.py:590-591), when the caller passes
    `deterministic=None`."""
    from llm_training_b200.lightning import Trainer
    from llm_training_b200.ops import attention_op
    monkeypatch.delenv('FLASH_ATTENTION_DETERMINISTIC', raising=False)
    assert attention_op.deterministic_default() is False
    monkeypatch.setenv('FLASH_ATTENTION_DETERMINISTIC', '1')
    assert attention_op.deterministic_default() is True
    monkeypatch.delenv('FLASH_ATTENTION_DETERMINISTIC')
    try:
        Trainer(deterministic=True, logger=False, enable_progress_bar=False, enable_checkpointing=False)
        assert torch.are_deterministic_algorithms_enabled()
        assert attention_op.deterministic_default() is True
    finally:
        torch.use_deterministic_algorithms(False)
    try:
        Trainer(deterministic='warn', logger=False, enable_progress_bar=False, enable_checkpointing=False)
        assert torch.are_deterministic_algorithms_enabled() and torch.is_deterministic_algorithms_warn_only_enabled()
    finally:
        torch.use_deterministic_algorithms(False)
    # the flag is part of the reference-compatible dispatcher's signature
    q = torch.randn(1, 8, 2, 16)
    out = attention_op.flash_attention_forward(q, q, q, None, 8, deterministic=True)
    assert out.shape == q.shape


def test_norm_multi_linear_residual_passthrough_keeps_values_and_gradients():
    """`return_residual=True` hands `x` back for the residual connection; on the PyTorch path it is `x` itself, and the
    result and every gradient equal the plain formulation."""
    from llm_training_b200.ops import norm_multi_linear
    torch.manual_seed(0)
    x = torch.randn(6, 32, requires_grad=True)
    nw = torch.rand(32, requires_grad=True)
    w1, w2 = torch.randn(16, 32, requires_grad=True), torch.randn(8, 32, requires_grad=True)
    y, x_res = norm_multi_linear(x, nw, 1e-5, [w1, w2], return_residual=True)
    (y.sum() + (x_res * 2).sum()).backward()
    got = [t.grad.clone() for t in (x, nw, w1, w2)]
    for t in (x, nw, w1, w2):
        t.grad = None
    y2 = norm_multi_linear(x, nw, 1e-5, [w1, w2])
    (y2.sum() + (x * 2).sum()).backward()
    assert torch.equal(y, y2)
    for a, t in zip(got, (x, nw, w1, w2)):
        assert torch.allclose(a, t.grad)
