"""Collective-communication interface used by the sharding engine and the TP plans.

Two implementations share the interface ("the comm layer needs a gloo-backed implementation
of the same interface as the NVLink kernels"):

* :class:`TorchDistComm` — ``torch.distributed`` (gloo on CPU for tests, NCCL on GPUs).
* :class:`llm_training_b200.parallel.nvlink.NVLinkComm` — symmetric-memory buffers mapped into every peer
  and hand-written sm_100a kernels that move data with peer loads/stores over NVLink 5 / NVSwitch.
"""
from typing import Any

import torch
import torch.distributed as dist

from .mesh import MeshAxis


class Comm:
    """Collectives over ONE mesh axis."""

    def __init__(self, axis: MeshAxis) -> None:
        self.axis = axis

    @property
    def size(self) -> int:
        return self.axis.size()

    @property
    def rank(self) -> int:
        return self.axis.rank

    # every method is a no-op / identity when the axis has a single member
    def all_gather_into(self, full: torch.Tensor, shard: torch.Tensor, async_op: bool = False) -> Any: ...
    def reduce_scatter_into(self, shard: torch.Tensor, full: torch.Tensor, average: bool = True,
                            async_op: bool = False) -> Any: ...
    def all_reduce(self, t: torch.Tensor, op: str = 'sum', async_op: bool = False) -> Any: ...
    def broadcast(self, t: torch.Tensor, src_index: int = 0) -> None: ...
    def barrier(self) -> None: ...


class TorchDistComm(Comm):
    _OPS = {'sum': dist.ReduceOp.SUM, 'max': dist.ReduceOp.MAX, 'min': dist.ReduceOp.MIN}

    def all_gather_into(self, full, shard, async_op=False):
        if self.size == 1:
            if full.data_ptr() != shard.data_ptr():
                full.view(-1)[:shard.numel()].copy_(shard.view(-1))
            return None
        return dist.all_gather_into_tensor(full, shard, group=self.axis.group, async_op=async_op)

    def reduce_scatter_into(self, shard, full, average=True, async_op=False):
        if self.size == 1:
            if shard.data_ptr() != full.data_ptr():
                shard.view(-1).copy_(full.view(-1)[:shard.numel()])
            return None
        backend = dist.get_backend(self.axis.group)
        if backend == 'gloo':
            # gloo has no reduce_scatter_tensor: all-reduce then slice (CPU test path only)
            dist.all_reduce(full, group=self.axis.group)
            n = shard.numel()
            shard.view(-1).copy_(full.view(-1)[self.rank * n:(self.rank + 1) * n])
            if average:
                shard.div_(self.size)
            return None
        op = dist.ReduceOp.AVG if average else dist.ReduceOp.SUM
        return dist.reduce_scatter_tensor(shard, full, op=op, group=self.axis.group, async_op=async_op)

    def all_reduce(self, t, op='sum', async_op=False):
        if self.size == 1:
            return None
        return dist.all_reduce(t, op=self._OPS[op], group=self.axis.group, async_op=async_op)

    def broadcast(self, t, src_index=0):
        if self.size > 1:
            dist.broadcast(t, src=self.axis.ranks[src_index], group=self.axis.group)

    def barrier(self):
        if self.size > 1:
            dist.barrier(group=self.axis.group)


_COMM_CACHE: dict[tuple, Comm] = {}


def get_comm(axis: MeshAxis | None, kind: str = 'auto') -> Comm:
    """Comm for a mesh axis. ``kind``: 'torch' | 'nvlink' | 'auto' (nvlink when the axis is a multi-GPU
    CUDA group on one node and symmetric memory is available, else torch)."""
    import os
    if axis is None:
        axis = MeshAxis('single', [0], 0, None)
    if kind == 'auto':
        kind = os.environ.get('LLMT_COMM', 'auto')      # torch | nvlink | auto
    key = (axis.name, tuple(axis.ranks), kind)
    if key in _COMM_CACHE:
        return _COMM_CACHE[key]
    comm: Comm | None = None
    if kind in ('nvlink', 'auto') and axis.size() > 1 and torch.cuda.is_available() \
            and dist.is_initialized() and dist.get_backend(axis.group) == 'nccl':
        try:
            from .nvlink import NVLinkComm
            comm = NVLinkComm(axis)
        except Exception as e:  # symmetric memory unavailable: fall back, loudly when explicitly requested
            if kind == 'nvlink':
                raise
            import logging
            logging.getLogger(__name__).warning(f'NVLink comm unavailable ({e!r}); using torch.distributed')
    if comm is None:
        comm = TorchDistComm(axis)
    _COMM_CACHE[key] = comm
    return comm
