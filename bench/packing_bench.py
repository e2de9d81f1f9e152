"""CPU benchmark of the pre-training packers: this repository's native best-fit packer (csrc/host/packing.cpp through
``llm_training_b200.data.packing``) against the reference's own function, run UNMODIFIED.

The reference's data module cannot be imported here (it needs `lightning`), so the source of
`PreTrainingDataModule._best_fit_bin_packing` / `_best_fit_decreasing`'s packing loop is read from the installed
reference tree (`baseline/_ref`, or `/root/reference/src`) with `ast` at run time and executed as is — nothing of it
is copied into this repository.

    python bench/packing_bench.py [--docs 1000 10000 100000] [--capacity 8192]

Prints one JSON line per batch size: seconds per call (best of 5), documents / s, identical bins yes / no.
"""
import argparse
import ast
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def load_reference_packer():
    rel = os.path.join('llm_training', 'data', 'pre_training', 'pre_training_datamodule.py')
    for base in (os.path.join(ROOT, 'baseline', '_ref'), '/root/reference/src'):
        path = os.path.join(base, rel)
        if os.path.exists(path):
            tree = ast.parse(open(path).read())
            for node in ast.walk(tree):
                if isinstance(node, ast.FunctionDef) and node.name == '_best_fit_bin_packing':
                    node.decorator_list = []          # a plain function taking (cls, capacity, lengths)
                    module = ast.Module(body=[node], type_ignores=[])
                    scope: dict = {}
                    exec(compile(ast.fix_missing_locations(module), path, 'exec'), scope)
                    return lambda capacity, lengths: scope['_best_fit_bin_packing'](None, capacity, lengths), path
    return None, None


def best_of(fn, repeat=5):
    best = float('inf')
    out = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    return best, out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--docs', type=int, nargs='+', default=[1000, 10000, 100000])
    ap.add_argument('--capacity', type=int, default=8192)
    args = ap.parse_args()
    from llm_training_b200.data.packing import best_fit_bin_packing
    from llm_training_b200.ops._lib import host_lib
    ref, ref_path = load_reference_packer()
    rng = np.random.default_rng(0)
    for n in args.docs:
        # document lengths: log-normal around ~600 tokens, clipped to the capacity, sorted decreasing (best-fit DEcreasing
        # is what both pipelines run: pre_training_datamodule.py:196-198)
        lengths = np.clip(rng.lognormal(6.4, 1.0, n).astype(np.int64), 8, args.capacity)
        lengths = sorted(lengths.tolist(), reverse=True)
        ours_s, ours = best_of(lambda: best_fit_bin_packing(args.capacity, lengths))
        row = {'docs': n, 'capacity': args.capacity, 'bins': len(ours), 'native_library': host_lib() is not None,
               'ours_s': ours_s, 'ours_docs_per_s': n / ours_s}
        if ref is not None:
            ref_s, theirs = best_of(lambda: ref(args.capacity, lengths), repeat=1 if n > 20000 else 3)
            row.update(reference_s=ref_s, reference_docs_per_s=n / ref_s, speedup=ref_s / ours_s,
                       identical_bins=theirs == ours, reference_source=ref_path)
        print(json.dumps(row), flush=True)


if __name__ == '__main__':
    main()
