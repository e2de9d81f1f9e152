"""``FSDP2Strategy``: 2-D mesh (data parallel x tensor parallel) over the flat-parameter engine.

Same constructor arguments as the reference (src/llm_training/lightning/strategy/fsdp2/
fsdp2_strategy.py:49-60): ``data_parallel_size``, ``tensor_parallel_size`` (ints or ``'auto'``),
``save_distributed_checkpoint``, ``process_group_backend``, ``timeout``, ``reshard_after_forward``,
``mp_policy``, ``offload_policy``, ``use_master_weights``.  ``'auto'/'auto'`` resolves to dp = number of
nodes, tp = GPUs per node (reference :181-191); TP is the inner, intra-node axis.
"""
import datetime
from typing import Any, Literal

from ..base import Strategy
from ..precision import MixedPrecisionPolicy, OffloadPolicy


class FSDP2Strategy(Strategy):
    def __init__(self, data_parallel_size: int | Literal['auto'] = 'auto',
                 tensor_parallel_size: int | Literal['auto'] = 1, save_distributed_checkpoint: bool = True,
                 process_group_backend: str | None = None,
                 timeout: datetime.timedelta | None = datetime.timedelta(minutes=30),
                 reshard_after_forward: bool | int | str = 'auto', mp_policy: MixedPrecisionPolicy | None = None,
                 offload_policy: OffloadPolicy | None = None, use_master_weights: bool = True) -> None:
        super().__init__(process_group_backend, timeout)
        self._data_parallel_size = data_parallel_size
        self._tensor_parallel_size = tensor_parallel_size
        self.save_distributed_checkpoint = save_distributed_checkpoint
        self.reshard_after_forward = reshard_after_forward
        self._mp_policy = mp_policy
        self._offload_policy = offload_policy or OffloadPolicy()
        self.use_master_weights = use_master_weights

    @property
    def mp_policy(self) -> MixedPrecisionPolicy:
        """The compute / reduce dtypes are a function of ``trainer.precision`` in this engine (gradients live in the
        flat buffer in the parameter dtype and are reduced in it).  An explicit ``mp_policy`` is accepted when it
        says the same thing and REJECTED when it asks for something else (the reference forwards it to
        ``fully_shard``, fsdp2_strategy.py:57,93-103); it is never silently dropped."""
        derived = self.precision_plugin.mp_policy
        given = self._mp_policy
        if given is None:
            return derived
        for field in ('param_dtype', 'reduce_dtype', 'output_dtype'):
            want = getattr(given, field, None)
            if want is not None and want != getattr(derived, field):
                raise ValueError(
                    f'FSDP2Strategy(mp_policy=...): {field}={want} conflicts with precision='
                    f'{self.precision_plugin.precision!r} ({field}={getattr(derived, field)}); this engine derives the '
                    'mixed-precision policy from `trainer.precision` — change that instead')
        return derived

    @property
    def offload_policy(self) -> OffloadPolicy:
        return self._offload_policy

    def _mesh_sizes(self) -> tuple[int, int]:
        world = self.world_size
        dp, tp = self._data_parallel_size, self._tensor_parallel_size
        if dp == 'auto' and tp == 'auto':
            dp = max(1, self.num_nodes)
            tp = world // dp
        elif dp == 'auto':
            dp = world // int(tp)
        elif tp == 'auto':
            tp = world // int(dp)
        dp, tp = int(dp), int(tp)
        if dp * tp != world:
            raise ValueError(f'data_parallel_size ({dp}) x tensor_parallel_size ({tp}) must equal the world size '
                             f'({world})')
        return dp, tp

    @property
    def device_mesh(self) -> Any:
        if self.mesh is None:
            raise RuntimeError('Accessing the device mesh before processes have initialized is not allowed.')
        return self.mesh
