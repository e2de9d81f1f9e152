#!/usr/bin/env python
"""Build the in-tree native libraries.

  llm_training_b200/_C/libllmt_kernels.so   csrc/kernels/*.cu   nvcc, sm_100a only
  llm_training_b200/_C/libllmt_host.so      csrc/host/*.cpp     g++ (CPU helpers: packing, batching)

Incremental (mtime based), parallel over translation units.  ``--force`` rebuilds everything,
``--verbose`` adds ``-Xptxas -v`` so register / spill / smem numbers are printed.
"""
import argparse
import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(ROOT, 'llm_training_b200', '_C')
OBJ_DIR = os.path.join(ROOT, 'build', 'obj')

NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a',
    '-O3', '-std=c++17', '-lineinfo',
    '--expt-relaxed-constexpr', '--expt-extended-lambda',
    '-Xcompiler', '-fPIC',
    '-Xcompiler', '-fvisibility=default',
    '-I', os.path.join(ROOT, 'include'),
    '-I', os.path.join(ROOT, 'csrc', 'kernels'),
]
CXX_FLAGS = ['-O3', '-std=c++17', '-fPIC', '-Wall', '-pthread', '-I', os.path.join(ROOT, 'include')]


def _nvcc() -> str:
    for cand in (shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found')


def _newer(src_deps: list[str], target: str) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in src_deps)


def _run(cmd: list[str], verbose: bool) -> None:
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0 or verbose:
        sys.stderr.write(' '.join(cmd) + '\n' + proc.stdout + proc.stderr)
    if proc.returncode != 0:
        raise RuntimeError(f'build failed: {cmd[0]} {cmd[-1]}')


def build_kernels(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    sources = sorted(glob.glob(os.path.join(ROOT, 'csrc', 'kernels', '*.cu')))
    headers = glob.glob(os.path.join(ROOT, 'csrc', 'kernels', '*.cuh')) + glob.glob(os.path.join(ROOT, 'include', '*.h'))
    nvcc = _nvcc()
    jobs, objects = [], []
    for src in sources:
        obj = os.path.join(OBJ_DIR, os.path.basename(src) + '.o')
        objects.append(obj)
        if force or _newer([src] + headers, obj):
            cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-c', src, '-o', obj]
            jobs.append(cmd)
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as pool:
        list(pool.map(lambda c: _run(c, verbose), jobs))
    target = os.path.join(OUT_DIR, 'libllmt_kernels.so')
    if jobs or force or not os.path.exists(target):
        _run([nvcc, '-shared', '-o', target] + objects, verbose)
    return target


def build_host(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    sources = sorted(glob.glob(os.path.join(ROOT, 'csrc', 'host', '*.cpp')))
    target = os.path.join(OUT_DIR, 'libllmt_host.so')
    if sources and (force or _newer(sources, target)):
        _run(['g++'] + CXX_FLAGS + ['-shared', '-o', target] + sources, verbose)
    return target


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--force', action='store_true')
    ap.add_argument('--verbose', action='store_true')
    ap.add_argument('--host-only', action='store_true')
    args = ap.parse_args()
    print(build_host(args.force, args.verbose))
    if not args.host_only:
        print(build_kernels(args.force, args.verbose))


if __name__ == '__main__':
    main()
