"""Strategy base: process-group setup, the dp x tp mesh, model/optimizer preparation on top of the
flat-parameter engine, collectives for the control plane, and checkpoint I/O.

This is the part of Lightning's ``ParallelStrategy`` + the reference's ``FSDP2Strategy`` glue
(src/llm_training/lightning/strategy/fsdp2/fsdp2_strategy.py:111-442) that every strategy here shares.
"""
import datetime
import logging
import os
from contextlib import contextmanager
from typing import Any, ContextManager

import torch
import torch.distributed as dist
from torch import nn
from torch.optim import Optimizer

from ...optim import ShardedOptimizer
from ...parallel import FlatParamEngine, Mesh, apply_fully_sharded, get_comm, init_mesh
from .precision import FSDP2Precision, MixedPrecisionPolicy, OffloadPolicy

logger = logging.getLogger(__name__)


class Strategy:
    """Single-process / data-parallel behaviour. Subclasses choose mesh sizes and sharding policy."""

    def __init__(self, process_group_backend: str | None = None,
                 timeout: datetime.timedelta | None = datetime.timedelta(minutes=30)) -> None:
        self._process_group_backend = process_group_backend
        self._timeout = timeout
        self._precision = FSDP2Precision('32-true')
        self.mesh: Mesh | None = None
        self.num_nodes = 1
        self.trainer = None
        self.lightning_module = None
        self.engines: list[FlatParamEngine] = []
        self.frozen_engines: list[FlatParamEngine] = []   # parameter-only engines of frozen models (DPO reference)
        self._root_device: torch.device | None = None
        self._grad_norm: torch.Tensor | None = None

    # ---------------------------------------------------------------- precision
    @property
    def precision_plugin(self) -> FSDP2Precision:
        return self._precision

    @precision_plugin.setter
    def precision_plugin(self, value: Any) -> None:
        if isinstance(value, FSDP2Precision):
            self._precision = value
        elif value is not None:
            self._precision = FSDP2Precision(getattr(value, 'precision', value))

    # ---------------------------------------------------------------- topology
    @property
    def root_device(self) -> torch.device:
        if self._root_device is None:
            self._root_device = self._pick_device()
        return self._root_device

    def _pick_device(self) -> torch.device:
        accel = getattr(self.trainer, 'accelerator', 'auto') if self.trainer is not None else 'auto'
        if accel in ('cpu',) or not torch.cuda.is_available():
            return torch.device('cpu')
        return torch.device('cuda', self.local_rank % torch.cuda.device_count())

    @property
    def world_size(self) -> int:
        return int(os.environ.get('WORLD_SIZE', '1')) if not dist.is_initialized() else dist.get_world_size()

    @property
    def global_rank(self) -> int:
        return int(os.environ.get('RANK', '0')) if not dist.is_initialized() else dist.get_rank()

    @property
    def local_rank(self) -> int:
        return int(os.environ.get('LOCAL_RANK', os.environ.get('SLURM_LOCALID', '0')))

    @property
    def is_global_zero(self) -> bool:
        return self.global_rank == 0

    def _mesh_sizes(self) -> tuple[int, int]:
        return self.world_size, 1

    @property
    def dp_mesh(self):
        return None if self.mesh is None else self.mesh.dp

    @property
    def tp_mesh(self):
        return None if self.mesh is None else self.mesh.tp

    @property
    def dp_size(self) -> int:
        return 1 if self.mesh is None else self.mesh.dp.size()

    @property
    def dp_rank(self) -> int:
        return 0 if self.mesh is None else self.mesh.dp.rank

    @property
    def tp_size(self) -> int:
        return 1 if self.mesh is None else self.mesh.tp.size()

    @property
    def tp_rank(self) -> int:
        return 0 if self.mesh is None else self.mesh.tp.rank

    @property
    def distributed_sampler_kwargs(self) -> dict[str, int]:
        """A TP group sees identical batches: the sampler is split over dp only (reference :149-153)."""
        return {'num_replicas': self.dp_size, 'rank': self.dp_rank}

    # ---------------------------------------------------------------- environment
    def setup_environment(self, trainer) -> None:
        self.trainer = trainer
        world = int(os.environ.get('WORLD_SIZE', '1'))
        if world > 1 and not dist.is_initialized():
            backend = self._process_group_backend or ('nccl' if self.root_device.type == 'cuda' else 'gloo')
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            os.environ.setdefault('MASTER_PORT', '29500')
            if self.root_device.type == 'cuda':
                torch.cuda.set_device(self.root_device)
            kwargs = {} if self._timeout is None else {'timeout': self._timeout}
            if backend == 'nccl':
                kwargs['device_id'] = self.root_device
            dist.init_process_group(backend, rank=int(os.environ['RANK']), world_size=world, **kwargs)
        elif self.root_device.type == 'cuda':
            torch.cuda.set_device(self.root_device)
        dp, tp = self._mesh_sizes()
        self.mesh = init_mesh(dp, tp, self.world_size, self.global_rank)
        self.precision_plugin.bind_device(self.root_device.type)
        self.precision_plugin.check_supported()     # GPU + a mode without native kernels: stop here, loudly

    # ---------------------------------------------------------------- model + optimizers
    reshard_after_forward: bool | int | str = 'auto'
    use_master_weights: bool = True

    @property
    def mp_policy(self) -> MixedPrecisionPolicy:
        return self.precision_plugin.mp_policy

    @property
    def offload_policy(self) -> OffloadPolicy:
        return OffloadPolicy()

    def model_init_context(self) -> ContextManager:
        """Build the model directly on the target device in the target dtype (no CPU staging copy)."""
        return _Stack(self.precision_plugin.module_init_context(), torch.device(self.root_device))

    def setup(self, trainer) -> None:
        module = trainer.lightning_module
        self.lightning_module = module
        # anything still on CPU / wrong dtype (metrics, user-built submodules)
        self._materialize(module)
        self.engines = self._build_engines(module)
        for e in self.engines + self.frozen_engines:
            e.sync_shards_from_full()     # weights were loaded / broadcast into the full buffers after sharding
        if getattr(module, 'weights_used_once_per_step', False):
            for e in self.engines:
                e.enable_direct_grad_accumulation(True)
        if self.overlap_optimizer_step and self.root_device.type == 'cuda':
            # the fused AdamW (+ parameter exchange) of step k runs under the forward pass of step k+1
            for e in self.engines:
                e.overlap_optimizer = True
        trainer.optimizers, trainer.lr_scheduler_configs = self._setup_optimizers(trainer, module)
        from ...utils.profiling import enable_nvtx, nvtx_requested
        if nvtx_requested() and self.root_device.type == 'cuda':
            for model in self._sharded_models(module):
                enable_nvtx(model)

    def _materialize(self, module: nn.Module) -> None:
        device = self.root_device
        for m in module.modules():
            has_meta = any(t.is_meta for t in list(m.parameters(recurse=False)) + list(m.buffers(recurse=False)))
            if has_meta:
                m.to_empty(device=device, recurse=False)
        for p in module.parameters():
            if p.device != device:
                p.data = p.data.to(device)
            if p.is_floating_point() and p.dtype != self.precision_plugin.param_dtype:
                p.data = p.data.to(self.precision_plugin.param_dtype)
        for m in module.modules():
            for name, b in list(m.named_buffers(recurse=False)):
                if b is not None and b.device != device:
                    m._buffers[name] = b.to(device)

    def _sharded_models(self, module: nn.Module) -> list[nn.Module]:
        from ...models.base_model import BaseModel
        found = [m for m in module.modules() if isinstance(m, BaseModel)]
        # outermost BaseModels only
        return [m for m in found if not any(m is not o and any(m is c for c in o.modules()) for o in found)]

    def _build_engines(self, module: nn.Module) -> list[FlatParamEngine]:
        engines = []
        self.frozen_engines = []
        for model in self._sharded_models(module):
            engine = getattr(model, '_flat_engine', None)
            if engine is None and not hasattr(model, '_flat_engine'):
                engine = apply_fully_sharded(model, self.dp_mesh, device=self.root_device,
                                             reshard_after_forward=self.reshard_after_forward)
            if engine is None:
                continue                  # frozen model with resident parameters: nothing to manage
            (self.frozen_engines if engine.frozen else engines).append(engine)
        claimed = {id(p) for e in engines for u in e.units for _, p in u.params}
        rest = [(n, p) for n, p in module.named_parameters() if p.requires_grad and id(p) not in claimed]
        if rest:
            engines.append(FlatParamEngine([('_module_rest', rest)], self.dp_mesh, self.root_device))
        return engines

    def _setup_optimizers(self, trainer, module) -> tuple[list[Optimizer], list[dict]]:
        configured = module.configure_optimizers()
        if configured is None:
            return [], []
        if isinstance(configured, dict):
            optimizer, sched = configured['optimizer'], configured.get('lr_scheduler')
        elif isinstance(configured, (tuple, list)):
            optimizer = configured[0][0] if isinstance(configured[0], (list, tuple)) else configured[0]
            sched = configured[1][0] if len(configured) > 1 and configured[1] else None
        else:
            optimizer, sched = configured, None
        extra = []
        if self.tp_size > 1:
            extra.append(get_comm(self.tp_mesh, 'torch'))
        wrapped = ShardedOptimizer(optimizer, self.engines, max_grad_norm=trainer.gradient_clip_val,
                                   norm_extra_axes=extra, offload=self.offload_optimizer_state,
                                   master_weights=bool(getattr(self, 'use_master_weights', True)))
        sched_cfgs = []
        if sched is not None:
            cfg = sched if isinstance(sched, dict) else {'scheduler': sched}
            cfg.setdefault('interval', 'epoch' if not isinstance(sched, dict) else 'step')
            cfg.setdefault('frequency', 1)
            sched_cfgs.append(cfg)
        return [wrapped], sched_cfgs

    # ---------------------------------------------------------------- step plumbing
    def training_step(self, *args: Any, **kwargs: Any) -> Any:
        with self.precision_plugin.forward_context():
            return self.lightning_module.training_step(*args, **kwargs)

    def validation_step(self, *args: Any, **kwargs: Any) -> Any:
        with self.precision_plugin.forward_context():
            return self.lightning_module.validation_step(*args, **kwargs)

    def backward(self, loss: torch.Tensor) -> None:
        self.lightning_module.backward(self.precision_plugin.pre_backward(loss))
        for e in self.engines:
            if not e.require_sync:
                e._rearm()

    @contextmanager
    def block_backward_sync(self):
        """No gradient reduction inside (non-boundary micro-batches; reference :430-442)."""
        for e in self.engines:
            e.set_requires_gradient_sync(False)
        try:
            yield
        finally:
            for e in self.engines:
                e.set_requires_gradient_sync(True)

    def optimizer_step(self, optimizer: Optimizer, closure=None) -> Any:
        out = self.precision_plugin.optimizer_step(optimizer, self.lightning_module, closure)
        for e in self.frozen_engines:
            # units of a frozen model that are read outside a module call (its LM-head weight in the fused loss) stay
            # whole for the step; the step boundary is where they go back to slices
            e.reshard_all()
        self._grad_norm = getattr(optimizer, 'last_grad_norm', None)
        if self.lightning_module is not None:
            self.lightning_module._grad_norm = self._grad_norm
        return out

    # ---------------------------------------------------------------- control-plane collectives
    def barrier(self, name: str | None = None) -> None:
        if dist.is_initialized() and dist.get_world_size() > 1:
            if dist.get_backend() == 'nccl':
                dist.barrier(device_ids=[self.root_device.index])
            else:
                dist.barrier()

    def broadcast(self, obj: Any, src: int = 0) -> Any:
        if not (dist.is_initialized() and dist.get_world_size() > 1):
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src)
        return box[0]

    def reduce(self, tensor: torch.Tensor | float, group: Any | None = None, reduce_op: str = 'mean') -> torch.Tensor:
        t = tensor if isinstance(tensor, torch.Tensor) else torch.tensor(float(tensor))
        if not (dist.is_initialized() and dist.get_world_size(group) > 1):
            return t
        t = t.detach().clone().to(self.root_device, dtype=torch.float32 if t.is_floating_point() else t.dtype)
        op = reduce_op.lower() if isinstance(reduce_op, str) else 'sum'
        dist.all_reduce(t, op=dist.ReduceOp.SUM if op in ('mean', 'avg', 'sum') else dist.ReduceOp.MAX, group=group)
        if op in ('mean', 'avg'):
            t = t / dist.get_world_size(group)
        return t

    # ---------------------------------------------------------------- checkpoint I/O (see checkpoint_io.py)
    save_distributed_checkpoint: bool = False
    # measured on 1, 2 and 8 B200s (profiles/bench_8gpu*_r1_final.log): no gain beyond noise, the step is power /
    # HBM bound and the push is already short; available with LLMT_OVERLAP_OPT=1 or by setting the attribute
    @property
    def offload_optimizer_state(self) -> bool:
        """fp32 masters + Adam moments in host memory (FSDP2 ``CPUOffloadPolicy`` / DeepSpeed ``offload_optimizer``)."""
        policy = getattr(self, 'offload_policy', None)
        return type(policy).__name__ == 'CPUOffloadPolicy' or bool(getattr(self, 'offload_optimizer', False))

    overlap_optimizer_step: bool = os.environ.get('LLMT_OVERLAP_OPT', '0') == '1'

    def save_checkpoint(self, checkpoint: dict[str, Any], filepath: str) -> None:
        from ..checkpoint_io import save_checkpoint
        self.synchronize_optimizer()
        save_checkpoint(self, checkpoint, filepath)

    def load_checkpoint(self, filepath: str) -> dict[str, Any]:
        from ..checkpoint_io import load_checkpoint
        return load_checkpoint(self, filepath)

    def remove_checkpoint(self, filepath: str) -> None:
        import shutil
        if self.is_global_zero:
            if os.path.isdir(filepath):
                shutil.rmtree(filepath, ignore_errors=True)
            elif os.path.exists(filepath):
                os.remove(filepath)

    # ---------------------------------------------------------------- Lightning strategy interface (names the
    # reference's FSDP2Strategy exposes, fsdp2_strategy.py:140-167, 241-263, 315-360, 422-428)
    @property
    def num_processes(self) -> int:
        """Processes (GPUs) of this node."""
        return max(1, self.world_size // max(1, self.num_nodes))

    @property
    def is_distributed(self) -> bool:
        return True

    @property
    def process_group_backend(self) -> str | None:
        return self._process_group_backend

    @property
    def restore_checkpoint_after_setup(self) -> bool:
        return True                      # weights are loaded into the flat buffers the engine builds in setup()

    @property
    def lightning_restore_optimizer(self) -> bool:
        return False                     # optimizer slices are restored by `load_checkpoint`, not by the trainer

    def setup_optimizers(self, trainer) -> None:
        trainer.optimizers, trainer.lr_scheduler_configs = self._setup_optimizers(trainer, trainer.lightning_module)

    def model_to_device(self) -> None:
        if self.lightning_module is not None:
            self._materialize(self.lightning_module)

    def tensor_init_context(self, empty_init: bool | None = None) -> ContextManager:
        """Dtype (and, with ``empty_init``, meta-device) context for building modules; materialisation is ``setup``."""
        cms = [self.precision_plugin.tensor_init_context()]
        if empty_init:
            cms.append(torch.device('meta'))
        return _Stack(*cms)

    def lightning_module_state_dict(self) -> dict[str, Any]:
        """Full-shape CPU state dict (ZeRO-3 units are gathered, TP shards are gathered over the tp axis)."""
        from ..checkpoint_io import module_state_dict
        return module_state_dict(self, balanced=False)

    def load_model_state_dict(self, checkpoint: Any, strict: bool = True) -> None:
        """Nothing to do: ``load_checkpoint`` already wrote the weights into the flat buffers."""

    def optimizer_state(self, optimizer: Optimizer) -> dict[str, Any]:
        return optimizer.state_dict()

    def load_optimizer_state_dict(self, checkpoint: Any) -> None:
        """Nothing to do: ``load_checkpoint`` already restored this rank's optimizer slices."""

    def set_world_ranks(self) -> None:
        """Ranks come from the launcher's environment (RANK / LOCAL_RANK / WORLD_SIZE); nothing to assign."""

    def synchronize_optimizer(self) -> None:
        """The current stream waits for in-flight (overlapped) optimizer work; call before touching parameters or
        optimizer state outside a model forward, and at timing boundaries."""
        for e in getattr(self, 'engines', []):
            e.synchronize_optimizer()

    def teardown(self) -> None:
        self.synchronize_optimizer()


class _Stack:
    def __init__(self, *cms: Any) -> None:
        self.cms = cms

    def __enter__(self):
        for cm in self.cms:
            cm.__enter__()

    def __exit__(self, *exc):
        for cm in reversed(self.cms):
            cm.__exit__(*exc)


class SingleDeviceStrategy(Strategy):
    """One process, one device (the default when no strategy is configured)."""

    def _mesh_sizes(self) -> tuple[int, int]:
        return self.world_size, 1
