"""Segment-id mask → per-token attention bounds, computed ONCE per batch, on device, with no
host synchronisation.

The reference re-derives ``cu_seqlens`` from the segment-id mask in every layer with a Python
loop and an ``.item()`` sync (src/llm_training/ops/attention_op.py:375-412),
and drops the mask entirely when the batch contains no pad token (llama_model.py:162-164; B2).

Here a batch ``attention_mask [B, S]`` (0 = pad, k>0 = k-th document of the pack) becomes two
int32 vectors over the flattened ``T = B*S`` tokens:

    seg_start[t] = flat index of the first token of t's document
    seg_end[t]   = flat index one past the last token of t's document

A query at flat position t may attend keys j with ``seg_start[t] <= j <= t``.  Fixed shapes, so
the attention kernels tile the token axis uniformly and need no ragged scheduling; pad runs form
their own "documents" whose outputs are ignored by the loss.
"""
import torch


@torch.no_grad()
def segment_bounds(attention_mask: torch.Tensor | None, batch_size: int, seq_len: int,
                   device: torch.device) -> tuple[torch.Tensor, torch.Tensor]:
    t = batch_size * seq_len
    idx = torch.arange(t, device=device, dtype=torch.int32)
    if attention_mask is None:
        row = idx // seq_len
        start = row * seq_len
        return start, start + seq_len
    seg = attention_mask.reshape(t).to(device=device, dtype=torch.int32)
    prev = torch.roll(seg, 1)
    # position 0 is a row start, so is_start[0] is already True and the roll below wraps it into the last slot: no
    # element assignment (a scalar written into a CUDA tensor is a pageable host-to-device copy that BLOCKS the host
    # until the stream drains — one lost step of launch run-ahead)
    is_start = (seg != prev) | (idx % seq_len == 0)
    start = torch.where(is_start, idx, torch.zeros_like(idx)).cummax(0).values.to(torch.int32)
    nxt_is_start = torch.roll(is_start, -1)
    end_marks = torch.where(nxt_is_start, idx + 1, torch.full_like(idx, t))
    end = end_marks.flip(0).cummin(0).values.flip(0).to(torch.int32)
    return start, end
