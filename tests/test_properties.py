"""Property tests (hypothesis) for the host-side building blocks every training run depends on: the C++ packers, the
segment-bound computation that replaces ``cu_seqlens``, label shifting, and the vectorised mask helpers."""
import torch
from hypothesis import given, settings, strategies as st

from llm_training_b200.data.packing import best_fit_bin_packing, group_indices_by_length
from llm_training_b200.ops import prepare_4d_causal_attention_mask, segment_bounds, shift_labels
from llm_training_b200.ops.attention_op import get_max_seqlen_in_batch, get_sequence_indices, get_unpad_data

lengths_st = st.lists(st.integers(min_value=1, max_value=64), min_size=0, max_size=200)


@settings(max_examples=60, deadline=None)
@given(lengths=lengths_st, capacity=st.integers(min_value=64, max_value=256))
def test_best_fit_packing_invariants(lengths, capacity):
    lengths = sorted(lengths, reverse=True)
    bins = best_fit_bin_packing(capacity, lengths)
    flat = sorted(i for b in bins for i in b)
    assert flat == list(range(len(lengths)))                            # every document exactly once
    assert all(sum(lengths[i] for i in b) <= capacity for b in bins)    # no bin overflows
    assert all(b for b in bins)                                         # no empty bin
    # best-fit-decreasing never needs more than first-fit's classic bound of 2 x optimum (+1)
    assert len(bins) <= 2 * (-(-sum(lengths) // capacity)) + 1


@settings(max_examples=60, deadline=None)
@given(lengths=st.lists(st.integers(min_value=1, max_value=40), min_size=1, max_size=100),
       max_length=st.integers(min_value=40, max_value=128))
def test_group_by_length_invariants(lengths, max_length):
    groups = group_indices_by_length(lengths, max_length)
    assert sorted(i for g in groups for i in g) == list(range(len(lengths)))
    for g in groups:
        # a group holds its documents plus one reserved slot per document after the first (reference semantics)
        assert len(g) == 1 or sum(lengths[i] for i in g) + len(g) - 1 <= max_length


def _segment_mask(draw_lengths, seq):
    """Row of segment ids 1,2,3.. for the given document lengths, zero padded to ``seq``."""
    row, seg = [], 1
    for n in draw_lengths:
        if len(row) + n > seq:
            break
        row += [seg] * n
        seg += 1
    return row + [0] * (seq - len(row))


@settings(max_examples=40, deadline=None)
@given(data=st.data(), rows=st.integers(min_value=1, max_value=3), seq=st.integers(min_value=4, max_value=48))
def test_segment_bounds_and_masks_agree_with_a_per_token_scan(data, rows, seq):
    mask = torch.tensor([_segment_mask(data.draw(st.lists(st.integers(1, 12), min_size=1, max_size=8)), seq)
                         for _ in range(rows)])
    start, end = segment_bounds(mask, rows, seq, torch.device('cpu'))
    start, end = start.view(rows, seq), end.view(rows, seq)
    allowed = prepare_4d_causal_attention_mask(mask, (rows, seq), torch.zeros(rows, seq, 2), 0) > -1.0
    for r in range(rows):
        for t in range(seq):
            if mask[r, t] == 0:
                continue
            first = t
            while first > 0 and mask[r, first - 1] == mask[r, t]:
                first -= 1
            last = t
            while last + 1 < seq and mask[r, last + 1] == mask[r, t]:
                last += 1
            assert int(start[r, t]) == r * seq + first and int(end[r, t]) == r * seq + last + 1
            want = torch.zeros(seq, dtype=torch.bool)
            want[first:t + 1] = True
            assert torch.equal(allowed[r, 0, t], want)
    # document lengths: three independent derivations agree
    runs = [(i, s, e) for i, s, e in get_sequence_indices(mask) if mask[i, s] != 0]
    lens_from_runs = sorted(e - s + 1 for _, s, e in runs)
    assert sorted(get_max_seqlen_in_batch(mask).tolist()) == lens_from_runs
    indices, cu, longest = get_unpad_data(mask)
    assert indices.numel() == int((mask != 0).sum()) and (cu[1:] - cu[:-1]).tolist() == [e - s + 1 for _, s, e in runs]
    assert longest == (max(lens_from_runs) if lens_from_runs else 0)


@settings(max_examples=40, deadline=None)
@given(rows=st.integers(1, 4), seq=st.integers(1, 32), ignore=st.sampled_from([-100, -1]))
def test_shift_labels_is_a_left_roll_with_ignore_at_the_end(rows, seq, ignore):
    labels = torch.arange(rows * seq).view(rows, seq)
    out = shift_labels(labels, ignore)
    assert out.shape == labels.shape and torch.equal(out[:, :-1], labels[:, 1:]) and bool((out[:, -1] == ignore).all())
