"""Sequence packers.  Thin Python over the native implementations in ``csrc/host/packing.cpp``
(pure-Python fallbacks with identical results are used when the host library has not been built)."""
import ctypes

import numpy as np

from ..ops._lib import host_lib


def best_fit_bin_packing(capacity: int, lengths: list[int]) -> list[list[int]]:
    """Bins of item indices; item i goes to the fullest bin that still fits it (ties: lowest index).
    Reference: src/llm_training/data/pre_training/pre_training_datamodule.py:144-179."""
    n = len(lengths)
    if n == 0:
        return []
    lib = host_lib()
    if lib is not None:
        arr = np.ascontiguousarray(lengths, dtype=np.int64)
        bin_of = np.empty(n, dtype=np.int64)
        lib.llmt_best_fit_pack.restype = ctypes.c_int64
        nbins = lib.llmt_best_fit_pack(arr.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(n),
                                       ctypes.c_int64(capacity), bin_of.ctypes.data_as(ctypes.c_void_p))
        bins: list[list[int]] = [[] for _ in range(nbins)]
        for i, b in enumerate(bin_of.tolist()):
            bins[b].append(i)
        return bins
    remaining: list[int] = []
    bins = []
    for i, length in enumerate(lengths):
        best, best_left = -1, None
        for j, space in enumerate(remaining):
            if space >= length and (best_left is None or space - length < best_left):
                best, best_left = j, space - length
        if best >= 0:
            remaining[best] -= length
            bins[best].append(i)
        else:
            remaining.append(capacity - length)
            bins.append([i])
    return bins


def group_indices_by_length(lengths: list[int], max_length: int) -> list[list[int]]:
    """Greedy groups over items sorted by length; every document reserves one extra slot
    (reference: instruction_tuning_datamodule.py:102-121)."""
    n = len(lengths)
    if n == 0:
        return []
    lib = host_lib()
    if lib is not None:
        arr = np.ascontiguousarray(lengths, dtype=np.int64)
        order = np.empty(n, dtype=np.int64)
        group_of = np.empty(n, dtype=np.int64)
        lib.llmt_group_by_length.restype = ctypes.c_int64
        ngroups = lib.llmt_group_by_length(arr.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(n),
                                           ctypes.c_int64(max_length), order.ctypes.data_as(ctypes.c_void_p),
                                           group_of.ctypes.data_as(ctypes.c_void_p))
        groups: list[list[int]] = [[] for _ in range(ngroups)]
        for i in order.tolist():
            groups[group_of[i]].append(i)
        return groups
    groups, current, total = [], [], 0
    for i, length in sorted(enumerate(lengths), key=lambda x: x[1]):
        if current and total + length + len(current) > max_length:
            groups.append(current)
            current, total = [], 0
        current.append(i)
        total += length
    if current:
        groups.append(current)
    return groups
