"""Shared, import-neutral case generator for tests/test_reference_oracle.py.

Executed twice with the SAME code: once in a subprocess whose ``llm_training`` is the unmodified reference
(``baseline/_ref`` first on PYTHONPATH), once in-process against this repository (``llm_training`` alias).  Every
case is a function of the imported ``llm_training`` package only, so the two runs are comparable value by value."""
import torch


def build_cases() -> dict:
    from llm_training.lr_schedulers import ConstantWarmupLR, LinearWarmupLR
    from llm_training.ops import apply_rope, rms_norm, shift_labels, swiglu
    from llm_training.ops.attention_op import (get_max_seqlen_in_batch, get_sequence_indices,
                                               prepare_4d_causal_attention_mask,
                                               update_4d_causal_mask_for_packed_sequence)
    from llm_training.ops.rope_utils import ROPE_INIT_FUNCTIONS, RoPEConfig

    out: dict = {}
    g = torch.Generator().manual_seed(0)
    x = torch.randn(3, 17, 64, generator=g)
    w = torch.randn(64, generator=g)
    out['rms_norm_fp32'] = rms_norm(x, w, 1e-5)
    out['rms_norm_bf16'] = rms_norm(x.bfloat16(), w.bfloat16(), 1e-6).float()

    q, k = torch.randn(2, 4, 9, 32, generator=g), torch.randn(2, 2, 9, 32, generator=g)
    ang = torch.randn(2, 9, 32, generator=g)
    qo, ko = apply_rope(q, k, ang.cos(), ang.sin())
    out['rope_q'], out['rope_k'] = qo, ko

    labels = torch.randint(0, 50, (3, 11), generator=g)
    out['shift_labels'] = shift_labels(labels)
    out['shift_labels_ignore'] = shift_labels(labels, ignore_index=-1)

    h = torch.randn(5, 32, generator=g)
    w1, w2, w3 = (torch.randn(48, 32, generator=g), torch.randn(48, 32, generator=g), torch.randn(32, 48, generator=g))
    out['swiglu_separate'] = swiglu(h, w3=w3, w1=w1, w2=w2)
    out['swiglu_fused'] = swiglu(h, w3=w3, w1w2=torch.cat([w1, w2]))

    rope_cfgs = {
        'default': dict(type='default', base=10000.0, dim=64, max_position_embeddings=2048),
        'linear': dict(type='linear', base=10000.0, dim=64, max_position_embeddings=2048,
                       scaling_config=dict(rope_type='linear', factor=4.0)),
        'dynamic': dict(type='dynamic', base=10000.0, dim=64, max_position_embeddings=2048,
                        scaling_config=dict(rope_type='dynamic', factor=2.0)),
        'yarn': dict(type='yarn', base=10000.0, dim=64, max_position_embeddings=8192,
                     scaling_config=dict(rope_type='yarn', factor=4.0, original_max_position_embeddings=2048)),
        'longrope': dict(type='longrope', base=10000.0, dim=64, max_position_embeddings=8192,
                         scaling_config=dict(rope_type='longrope', factor=4.0, short_factor=[1.0 + 0.01 * i for i in range(32)],
                                             long_factor=[2.0 + 0.05 * i for i in range(32)],
                                             original_max_position_embeddings=2048)),
        'llama3': dict(type='llama3', base=500000.0, dim=128, max_position_embeddings=131072,
                       scaling_config=dict(rope_type='llama3', factor=8.0, low_freq_factor=1.0, high_freq_factor=4.0,
                                           original_max_position_embeddings=8192)),
    }
    for name, kw in rope_cfgs.items():
        for seq_len in (None, 6000):
            try:
                inv, scale = ROPE_INIT_FUNCTIONS[name](RoPEConfig(**kw), device=torch.device('cpu'), seq_len=seq_len)
                out[f'rope_init_{name}_{seq_len}'] = inv.float()
                out[f'rope_init_{name}_{seq_len}_scale'] = torch.tensor(float(scale))
            except Exception as e:  # noqa: BLE001 -- both sides must fail the same way
                out[f'rope_init_{name}_{seq_len}_error'] = type(e).__name__

    # packed (segment-id) and padded masks -> 4-D additive masks
    seg = torch.tensor([[1, 1, 1, 2, 2, 3, 0, 0], [1, 1, 1, 1, 1, 1, 1, 1]])
    emb = torch.zeros(2, 8, 4)
    # rows of padding queries are compared nowhere: the reference leaves them fully masked (uniform softmax over
    # finfo.min), this repository lets a padding token see itself; their outputs never reach a loss
    m = prepare_4d_causal_attention_mask(seg, (2, 8), emb, 0)
    out['mask4d_packed_allowed'] = (m > -1.0).float() * (seg != 0)[:, None, :, None]
    seg3 = torch.tensor([[1, 1, 1, 2, 2, 3, 0, 0], [1, 1, 1, 1, 1, 1, 1, 1], [2, 2, 1, 1, 0, 0, 0, 3]])
    idx = get_sequence_indices(seg3)
    out['sequence_indices'] = torch.tensor(idx)
    out['doc_lengths'] = get_max_seqlen_in_batch(seg3)
    base = torch.zeros(3, 1, 8, 8)
    out['packed_mask_update'] = (update_4d_causal_mask_for_packed_sequence(base, idx) < -1.0).float()
    pad = torch.tensor([[1, 1, 1, 1, 1, 0, 0, 0], [1, 1, 1, 1, 1, 1, 1, 1]])
    m = prepare_4d_causal_attention_mask(pad, (2, 8), emb, 0)
    out['mask4d_padded_allowed'] = (m > -1.0).float() * (pad != 0)[:, None, :, None]
    m = prepare_4d_causal_attention_mask(pad, (2, 8), emb, 0, sliding_window=3)
    out['mask4d_window_allowed'] = (m > -1.0).float() * (pad != 0)[:, None, :, None]

    for cls, kw in ((ConstantWarmupLR, dict(num_warmup_steps=4)), (LinearWarmupLR, dict(num_warmup_steps=3,
                                                                                        num_total_steps=12,
                                                                                        min_lr=1e-5))):
        p = torch.nn.Parameter(torch.zeros(1))
        opt = torch.optim.SGD([p], lr=1e-3)
        sched = cls(opt, **kw)
        lrs = []
        for _ in range(12):
            lrs.append(opt.param_groups[0]['lr'])
            opt.step()
            sched.step()
        out[f'lr_{cls.__name__}'] = torch.tensor(lrs)
    return out


if __name__ == '__main__':
    import sys
    torch.save(build_cases(), sys.argv[1])
