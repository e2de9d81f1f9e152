"""2-D process mesh ``(data_parallel, tensor_parallel)`` with TP as the inner (intra-node, NVLink) axis
(reference: ``init_device_mesh((dp, tp), ...)``, src/llm_training/lightning/strategy/fsdp2/fsdp2_strategy.py:181-199)."""
from dataclasses import dataclass, field
from typing import Any

import torch.distributed as dist


@dataclass
class MeshAxis:
    """One axis of the mesh as seen from this rank: its peers, in axis order."""
    name: str
    ranks: list[int]                    # global ranks of the peers along this axis (includes self)
    rank: int                           # index of this process inside ``ranks``
    group: Any | None = None            # torch.distributed process group (None when size == 1 / no dist)

    def size(self) -> int:
        return len(self.ranks)

    def get_group(self) -> Any | None:
        return self.group

    def get_local_rank(self) -> int:
        return self.rank


@dataclass
class Mesh:
    world_size: int
    global_rank: int
    dp: MeshAxis
    tp: MeshAxis
    axes: dict[str, MeshAxis] = field(default_factory=dict)

    def __getitem__(self, name: str) -> MeshAxis:
        return self.axes[name]


def init_mesh(dp_size: int, tp_size: int, world_size: int | None = None, rank: int | None = None) -> Mesh:
    """Create every sub-group collectively (all ranks must call this with the same arguments)."""
    initialised = dist.is_available() and dist.is_initialized()
    world_size = world_size if world_size is not None else (dist.get_world_size() if initialised else 1)
    rank = rank if rank is not None else (dist.get_rank() if initialised else 0)
    assert dp_size * tp_size == world_size, f'dp({dp_size}) x tp({tp_size}) != world size ({world_size})'
    dp_index, tp_index = divmod(rank, tp_size)

    my_tp: MeshAxis | None = None
    my_dp: MeshAxis | None = None
    for d in range(dp_size):                       # tp groups: consecutive ranks
        ranks = [d * tp_size + t for t in range(tp_size)]
        group = dist.new_group(ranks) if (initialised and tp_size > 1) else None
        if d == dp_index:
            my_tp = MeshAxis('tensor_parallel', ranks, tp_index, group)
    for t in range(tp_size):                       # dp groups: stride tp_size
        ranks = [d * tp_size + t for d in range(dp_size)]
        group = dist.new_group(ranks) if (initialised and dp_size > 1) else None
        if t == tp_index:
            my_dp = MeshAxis('data_parallel', ranks, dp_index, group)
    assert my_tp is not None and my_dp is not None
    return Mesh(world_size, rank, my_dp, my_tp, {'data_parallel': my_dp, 'tensor_parallel': my_tp})
