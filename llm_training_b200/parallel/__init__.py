"""Parallel runtime: device mesh (dp x tp), communication backends (torch.distributed gloo/nccl and
NVLink symmetric-memory kernels), the flat-parameter sharding engine (FSDP / ZeRO) and the
tensor/sequence-parallel plans.  Replaces torch FSDP2 + DTensor + DeepSpeed in the reference
(SURVEY §2.4)."""
from .comm import Comm, TorchDistComm, get_comm
from .fsdp import FlatParamEngine, ShardedUnit, apply_fully_sharded
from .mesh import Mesh, MeshAxis, init_mesh

__all__ = ['Mesh', 'MeshAxis', 'init_mesh', 'Comm', 'TorchDistComm', 'get_comm', 'FlatParamEngine', 'ShardedUnit',
           'apply_fully_sharded']
