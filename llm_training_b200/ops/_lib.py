"""ctypes binding to the in-tree sm_100a kernel library (``llm_training_b200/_C/libllmt_kernels.so``).

The library is plain C-ABI (``include/llmt_kernels.h``): raw device pointers, sizes and a
``cudaStream_t``.  Every scratch buffer is allocated by the caller (torch's caching allocator) and
passed in, so the kernels themselves never allocate.  The library is built by
``python build_kernels.py`` / ``__graft_entry__.build()`` with
``nvcc -gencode arch=compute_100a,code=sm_100a``.

On a CUDA device the library is mandatory: ``lib()`` raises if it is missing rather than silently
running a PyTorch fallback.
"""
import ctypes
import os
from functools import lru_cache

import torch

_C_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), '_C')
KERNEL_LIB_PATH = os.path.join(_C_DIR, 'libllmt_kernels.so')
HOST_LIB_PATH = os.path.join(_C_DIR, 'libllmt_host.so')


class KernelLibraryMissing(RuntimeError):
    pass


@lru_cache(maxsize=None)
def lib() -> ctypes.CDLL:
    if not os.path.exists(KERNEL_LIB_PATH):
        raise KernelLibraryMissing(
            f'{KERNEL_LIB_PATH} not found: run `python build_kernels.py` (needs nvcc) — CUDA tensors '
            'must run on the in-tree sm_100a kernels, there is no PyTorch fallback on GPU.')
    handle = ctypes.CDLL(KERNEL_LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    handle.llmt_last_error.restype = ctypes.c_char_p
    return handle


@lru_cache(maxsize=None)
def host_lib() -> ctypes.CDLL | None:
    """CPU-side native helpers (bin packing, batch assembly). Optional: pure-Python fallback exists."""
    if not os.path.exists(HOST_LIB_PATH):
        return None
    return ctypes.CDLL(HOST_LIB_PATH)


def kernels_available() -> bool:
    return os.path.exists(KERNEL_LIB_PATH)


def use_kernels(*tensors: torch.Tensor) -> bool:
    """True when the op must run on the sm_100a kernels: CUDA + bf16 activations."""
    t = tensors[0]
    return t.is_cuda and t.dtype == torch.bfloat16


def ptr(t: torch.Tensor | None) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)
_raw_device = getattr(torch._C, '_cuda_getDevice', None)


def stream() -> ctypes.c_void_p:
    """The current stream's handle.  Called once per kernel launch (~1200 times per step under tensor parallelism):
    the two raw accessors cost ~0.3 us, `torch.cuda.current_stream().cuda_stream` ~15 us."""
    if _raw_stream is not None and _raw_device is not None:
        return ctypes.c_void_p(_raw_stream(_raw_device()))
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def check(code: int, what: str) -> None:
    if code != 0:
        msg = lib().llmt_last_error()
        raise RuntimeError(f'{what} failed (code {code}): {msg.decode() if msg else "?"}')


def c_int(v) -> ctypes.c_int:
    return ctypes.c_int(int(v))


def c_i64(v) -> ctypes.c_int64:
    return ctypes.c_int64(int(v))


def c_f32(v) -> ctypes.c_float:
    return ctypes.c_float(float(v))
