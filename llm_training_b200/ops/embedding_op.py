"""Token embedding gather with NEFTune noise fused in, and its scatter-add backward.

Reference: ``nn.Embedding`` (models/llama/llama_model.py:48,160) and the NEFTune forward hook
(lms/clm/clm.py:45-78): noise ~ U(-1,1) * alpha / sqrt(L * d) on non-pad tokens, L = number of
non-pad tokens of the batch row.
"""
import torch
import torch.nn.functional as F

from . import _lib


def neftune_noise_scale(attention_mask: torch.Tensor | None, batch_size: int, seq_len: int, dim: int, alpha: float,
                        device: torch.device, global_seq_len_factor: int = 1) -> torch.Tensor:
    """fp32 [B*S]: alpha / sqrt(len_row * dim) for real tokens, 0 for padding."""
    if attention_mask is None:
        mask = torch.ones((batch_size, seq_len), dtype=torch.float32, device=device)
    else:
        mask = attention_mask.to(device).bool().float()
    lengths = mask.sum(1, keepdim=True) * global_seq_len_factor
    scale = alpha / torch.sqrt(lengths.clamp_min(1.0) * dim)
    return (mask * scale).reshape(-1).contiguous()


class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids: torch.Tensor, weight: torch.Tensor, noise_scale: torch.Tensor | None, seed: int,
                padding_idx: int | None) -> torch.Tensor:
        flat = ids.reshape(-1).contiguous()
        tokens, (vocab, dim) = flat.numel(), weight.shape
        out = torch.empty((tokens, dim), dtype=weight.dtype, device=weight.device)
        _lib.check(_lib.lib().llmt_embedding_fwd(_lib.ptr(flat), _lib.ptr(weight), _lib.ptr(out),
                                                 _lib.ptr(noise_scale), _lib.c_int(tokens), _lib.c_int(dim),
                                                 _lib.c_int(vocab), _lib.ctypes.c_uint64(seed), _lib.stream()),
                   'embedding_fwd')
        ctx.save_for_backward(flat)
        ctx.meta = (vocab, dim, -1 if padding_idx is None else padding_idx, weight.dtype)
        return out.view(*ids.shape, dim)

    @staticmethod
    def backward(ctx, dout: torch.Tensor):
        (flat,) = ctx.saved_tensors
        vocab, dim, padding_idx, wdtype = ctx.meta
        d2 = dout.reshape(-1, dim).contiguous()
        # Tokens sorted by id: every distinct id's rows are summed in fp32 by one CTA, in a fixed order, and written
        # once as bf16 — no atomics (reproducible), no fp32 [vocab, dim] scratch and no conversion pass.
        sorted_ids, perm = torch.sort(flat, stable=True)
        dw = torch.empty((vocab, dim), dtype=wdtype, device=dout.device)
        _lib.check(_lib.lib().llmt_embedding_bwd_sorted(_lib.ptr(sorted_ids), _lib.ptr(perm), _lib.ptr(d2),
                                                        _lib.ptr(dw), _lib.c_int(flat.numel()), _lib.c_int(dim),
                                                        _lib.c_int(vocab), _lib.c_i64(padding_idx), _lib.stream()),
                   'embedding_bwd')
        return None, dw, None, None, None


def embedding(ids: torch.Tensor, weight: torch.Tensor, padding_idx: int | None = None,
              noise_scale: torch.Tensor | None = None, seed: int = 0) -> torch.Tensor:
    if weight.is_cuda and weight.dtype == torch.bfloat16 and weight.shape[1] % 8 == 0:
        return _EmbeddingFn.apply(ids, weight, noise_scale, seed, padding_idx)
    out = F.embedding(ids, weight, padding_idx)
    if noise_scale is not None:
        gen = torch.Generator(device=out.device).manual_seed(seed)
        noise = torch.rand(out.shape, generator=gen, device=out.device, dtype=torch.float32) * 2 - 1
        out = out + (noise * noise_scale.view(*ids.shape, 1)).to(out.dtype)
    return out
