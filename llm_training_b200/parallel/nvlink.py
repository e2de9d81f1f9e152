"""NVLink 5 / NVSwitch communication backend built on symmetric memory + the kernels in ``csrc/kernels/comm.cu``.

``torch.distributed._symmetric_memory`` is used ONLY as the allocator / handle exchange (cuMem + fabric or fd
handles mapped into every peer, plus a signal pad per rank); every byte that crosses NVLink is moved by the
in-tree kernels with plain loads / stores on the mapped peer addresses.

Protocol notes
  * barrier / one-shot all-reduce use uint32 slots in the signal pads with monotonically increasing epochs
    (``st.release.sys`` by the writer, ``ld.acquire.sys`` spin by the reader);
  * all ranks issue the same sequence of collectives, so epochs agree without any host-side negotiation: parameter
    gathers follow the module order, and the sharding engine issues gradient reductions in one fixed order (last unit
    first) even when a rank's gradients arrive in another order or not at all (``FlatParamEngine._launch_ready_reductions``);
  * kernels run on the caller's current stream; the sharding engine puts them on its communication stream.
"""
import ctypes
from typing import Any

import torch
import torch.distributed as dist

from ..ops import _lib
from .comm import Comm
from .mesh import MeshAxis

_BARRIER_SLOT = 256      # uint32 index inside the signal pad (torch's own channels use the low slots)
_ALLREDUCE_SLOT = 320


def _ptr_array(ptrs: list[int]):
    arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
    return arr


class SymmetricBuffer:
    """A tensor that exists at the same offset on every rank of the axis, with the peers' addresses."""

    def __init__(self, tensor: torch.Tensor, handle: Any) -> None:
        self.tensor = tensor
        self.handle = handle
        self.peer_ptrs: list[int] = [int(p) for p in handle.buffer_ptrs]
        self._ptr_cache: dict[int, Any] = {}      # byte offset -> ctypes array of the peers' addresses
        # NVSwitch multicast mapping of the same buffer (0 when the fabric / driver does not offer one)
        try:
            self.multicast_ptr = int(getattr(handle, 'multicast_ptr', 0) or 0)
        except Exception:  # noqa: BLE001
            self.multicast_ptr = 0

    def peer_ptrs_at(self, byte_offset: int):
        arr = self._ptr_cache.get(byte_offset)
        if arr is None:
            if len(self._ptr_cache) > 4096:
                self._ptr_cache.clear()
            arr = self._ptr_cache[byte_offset] = _ptr_array([p + byte_offset for p in self.peer_ptrs])
        return arr


class NVLinkComm(Comm):
    is_nvlink = True

    def __init__(self, axis: MeshAxis) -> None:
        super().__init__(axis)
        import torch.distributed._symmetric_memory as symm_mem
        self._symm = symm_mem
        assert axis.group is not None and axis.size() > 1
        self.group = axis.group
        self.device = torch.device('cuda', torch.cuda.current_device())
        symm_mem.enable_symm_mem_for_group(self.group.group_name) if hasattr(symm_mem, 'enable_symm_mem_for_group') \
            else None
        self._scalars = self.alloc(2 * 32 * max(axis.size(), 1), torch.float32)  # [2][world][32], half = epoch parity
        self._scalars.tensor.zero_()
        pad_bytes = int(self._scalars.handle.signal_pad_size)
        assert pad_bytes >= 4 * (_ALLREDUCE_SLOT + axis.size()), f'signal pad too small ({pad_bytes} B)'
        self._pads = _ptr_array([int(p) for p in self._scalars.handle.signal_pad_ptrs])
        self._barrier_epoch = 0
        self._allreduce_epoch = 0
        import os
        self.use_multicast = os.environ.get('LLMT_MULTICAST', '1') != '0'   # NVSwitch multicast stores when mapped
        # gradient reduce-scatter by in-switch reduction (multimem.ld_reduce) when the buffer has a multicast mapping
        self.use_multimem_rs = os.environ.get('LLMT_MULTIMEM_RS', '1') != '0'
        self.last_rs_kind = None          # 'multimem' | 'pull': what the last reduce-scatter actually ran
        self._lib = _lib.lib()
        self._rs_partial = torch.empty(self._lib.llmt_comm_rs_num_partials(), dtype=torch.float32, device=self.device)
        dist.barrier(group=self.group, device_ids=[self.device.index])

    # ---------------------------------------------------------------- allocation
    def alloc(self, numel: int, dtype: torch.dtype) -> SymmetricBuffer:
        t = self._symm.empty(numel, dtype=dtype, device=self.device)
        handle = self._symm.rendezvous(t, self.group)
        return SymmetricBuffer(t, handle)

    # ---------------------------------------------------------------- primitives
    def barrier(self) -> None:
        self._barrier_epoch += 1
        _lib.check(self._lib.llmt_comm_barrier(self._pads, _lib.c_int(self.rank), _lib.c_int(self.size),
                                               _lib.c_int(_BARRIER_SLOT), ctypes.c_uint32(self._barrier_epoch),
                                               _lib.stream()), 'comm_barrier')

    def all_reduce(self, t: torch.Tensor, op: str = 'sum', async_op: bool = False) -> None:
        if op == 'sum' and t.dtype == torch.float32 and t.numel() <= 32 and t.is_contiguous():
            self._allreduce_epoch += 1
            _lib.check(self._lib.llmt_comm_allreduce_f32(
                self._scalars.peer_ptrs_at(0), self._pads, _lib.c_int(self.rank), _lib.c_int(self.size),
                _lib.c_int(_ALLREDUCE_SLOT), ctypes.c_uint32(self._allreduce_epoch), _lib.ptr(t),
                _lib.c_int(t.numel()), _lib.stream()), 'comm_allreduce_f32')
            return None
        return dist.all_reduce(t, op={'sum': dist.ReduceOp.SUM, 'max': dist.ReduceOp.MAX}[op], group=self.group)

    def broadcast(self, t: torch.Tensor, src_index: int = 0) -> None:
        dist.broadcast(t, src=self.axis.ranks[src_index], group=self.group)

    def all_gather_into(self, full: torch.Tensor, shard: torch.Tensor, async_op: bool = False) -> None:
        """Generic tensors (not in the symmetric heap) fall back to NCCL; the engine uses ``allgather_pull``."""
        return dist.all_gather_into_tensor(full, shard, group=self.group)

    def reduce_scatter_into(self, shard, full, average: bool = True, async_op: bool = False) -> None:
        return dist.reduce_scatter_tensor(shard, full, op=dist.ReduceOp.AVG if average else dist.ReduceOp.SUM,
                                          group=self.group)

    # ---------------------------------------------------------------- engine collectives on symmetric arenas
    def allgather_pull(self, arena: SymmetricBuffer, byte_offset: int, shard_bytes: int) -> None:
        self.barrier()                                           # every owner's slice is final
        _lib.check(self._lib.llmt_comm_allgather_pull(arena.peer_ptrs_at(byte_offset), _lib.c_int(self.rank),
                                                      _lib.c_int(self.size), _lib.c_i64(shard_bytes), _lib.stream()),
                   'comm_allgather_pull')

    def gather_shards(self, arena: SymmetricBuffer, byte_offset: int, shard_bytes: int, full: torch.Tensor,
                      barrier: bool = True) -> None:
        """ZeRO-3 parameter gather: ``full`` (local, not symmetric) <- every rank's slice at ``byte_offset`` of the
        symmetric shard arena.  ``barrier``: the slices may have been written since the last cross-rank barrier."""
        if barrier:
            self.barrier()                                       # every owner's slice is final
        _lib.check(self._lib.llmt_comm_gather_shards(arena.peer_ptrs_at(byte_offset), _lib.ptr(full),
                                                     _lib.c_int(self.rank), _lib.c_int(self.size),
                                                     _lib.c_i64(shard_bytes), _lib.stream()), 'comm_gather_shards')

    def gather_range(self, arena: SymmetricBuffer, byte_offset: int, shard_bytes: int, full: torch.Tensor,
                     lo: int, hi: int, skip_lo: int = 0, skip_hi: int = 0, barrier: bool = True) -> None:
        """Like ``gather_shards`` for the byte range [lo, hi) of the unit minus [skip_lo, skip_hi)."""
        if barrier:
            self.barrier()
        _lib.check(self._lib.llmt_comm_gather_range(arena.peer_ptrs_at(byte_offset), _lib.ptr(full),
                                                    _lib.c_int(self.size), _lib.c_i64(shard_bytes), _lib.c_i64(lo),
                                                    _lib.c_i64(hi), _lib.c_i64(skip_lo), _lib.c_i64(skip_hi),
                                                    _lib.stream()), 'comm_gather_range')

    def gemm_weight_gather(self, x: torch.Tensor, arena: SymmetricBuffer, byte_offset: int, full: torch.Tensor,
                           w_off: int, shard_numel: int, n: int, flags: torch.Tensor) -> torch.Tensor:
        """``x @ W^T`` where W = elements [w_off, w_off + n*k) of the unit: the GEMM's pull warps land W in ``full``
        while the tensor cores work (csrc/kernels/gemm.cu, kDistA == 3).  No barrier here: the caller has already
        waited for one that follows the last write of every slice."""
        m, k = x.shape
        y = torch.empty((m, n), dtype=x.dtype, device=x.device)
        _lib.check(self._lib.llmt_gemm_bf16_wgather(
            _lib.ptr(x), arena.peer_ptrs_at(byte_offset), _lib.ptr(full), _lib.c_i64(w_off), _lib.c_i64(shard_numel),
            _lib.ptr(y), _lib.ptr(flags), _lib.c_int(self.rank), _lib.c_int(self.size), _lib.c_int(m), _lib.c_int(n),
            _lib.c_int(k), _lib.c_i64(x.stride(0)), _lib.c_i64(n), _lib.c_int(2), _lib.c_int(0), _lib.stream()),
            'gemm_bf16_wgather')
        return y

    def reduce_scatter_sumsq(self, arena: SymmetricBuffer, byte_offset: int, shard_numel: int, out_shard: torch.Tensor,
                             norm_sq: torch.Tensor, accumulate_norm: bool) -> None:
        self.barrier()                                           # every rank finished writing this unit's grads
        # the switch's summation order is not specified: deterministic mode sums the ranks' copies in rank order
        if (self.use_multimem_rs and arena.multicast_ptr and out_shard.dtype == torch.bfloat16
                and not torch.are_deterministic_algorithms_enabled()):
            # one multimem.ld_reduce per 16 bytes: the NVSwitch sums every rank's copy, this GPU receives 1/world of
            # what the pull version reads
            self.last_rs_kind = 'multimem'
            _lib.check(self._lib.llmt_comm_rs_sumsq_mc(
                ctypes.c_void_p(arena.multicast_ptr + byte_offset), _lib.c_int(self.rank), _lib.c_int(self.size),
                _lib.c_i64(shard_numel), _lib.ptr(out_shard), _lib.ptr(self._rs_partial), _lib.ptr(norm_sq),
                _lib.c_int(accumulate_norm), _lib.stream()), 'comm_rs_sumsq_mc')
            return
        self.last_rs_kind = 'pull'
        _lib.check(self._lib.llmt_comm_rs_sumsq(
            arena.peer_ptrs_at(byte_offset), _lib.c_int(self.rank), _lib.c_int(self.size), _lib.c_i64(shard_numel),
            _lib.c_int(out_shard.dtype == torch.float32), _lib.ptr(out_shard), _lib.ptr(self._rs_partial),
            _lib.ptr(norm_sq), _lib.c_int(accumulate_norm), _lib.stream()), 'comm_rs_sumsq')

    def adamw_push(self, master: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor,
                   grad_shard: torch.Tensor, arena: SymmetricBuffer, byte_offset: int, shard_numel: int, *,
                   lr: float, beta1: float, beta2: float, eps: float, weight_decay: float, step: int,
                   max_norm: float, norm_sq: torch.Tensor | None, grad_scale: float = 1.0) -> None:
        mc = arena.multicast_ptr + byte_offset if (arena.multicast_ptr and self.use_multicast) else 0
        _lib.check(self._lib.llmt_comm_adamw_push_mc(
            _lib.ptr(master), _lib.ptr(exp_avg), _lib.ptr(exp_avg_sq), _lib.ptr(grad_shard),
            arena.peer_ptrs_at(byte_offset), ctypes.c_void_p(mc), _lib.c_int(self.rank), _lib.c_int(self.size), _lib.c_i64(shard_numel),
            _lib.c_int(grad_shard.dtype == torch.float32), _lib.c_f32(lr), _lib.c_f32(beta1), _lib.c_f32(beta2),
            _lib.c_f32(eps), _lib.c_f32(weight_decay), _lib.c_int(step), _lib.c_f32(grad_scale), _lib.c_f32(max_norm),
            _lib.ptr(norm_sq), _lib.stream()), 'comm_adamw_push')
