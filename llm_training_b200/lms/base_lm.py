"""Shared LM lifecycle (reference: src/llm_training/lms/base_lm.py:32-313): model construction under the
right init context, freezing by regex, parallelisation, (pre-trained) weight loading with rank-0 read +
broadcast, optimizer/scheduler construction, gradient-norm logging, required checkpoint keys."""
import inspect
import logging
import re
from typing import Any

import torch

from ..lightning.module import LightningModule
from ..models.base_model import BaseModel
from .base_lm_config import BaseLightningModuleConfig

logger = logging.getLogger(__name__)


class BaseLightningModule(LightningModule):
    def __init__(self, config: BaseLightningModuleConfig) -> None:
        super().__init__()
        self.config = config
        self._grad_norm: torch.Tensor | float | None = None
        user_configure = self.configure_model

        def wrapped_configure_model() -> None:
            with BaseModel.init_weights_context(self.should_initialize_weights):
                user_configure()
            self.on_after_configure_model()
            if self.strategy is not None:
                self.strategy.barrier()

        self.configure_model = wrapped_configure_model   # type: ignore[method-assign]

    # ---------------------------------------------------------------- state queries
    @property
    def strategy(self):
        return None if self._trainer is None else self._trainer.strategy

    @property
    def has_pre_trained_weights(self) -> bool:
        return self.config.pre_trained_weights is not None

    @property
    def is_loading_from_checkpoint(self) -> bool:
        return self._trainer is not None and self._trainer.ckpt_path is not None

    @property
    def should_load_pre_trained_weights(self) -> bool:
        return self.has_pre_trained_weights and self.config.load_weights and not self.is_loading_from_checkpoint

    @property
    def should_initialize_weights(self) -> bool:
        return self.config.init_weights and not self.is_loading_from_checkpoint and not self.config.load_weights

    @property
    def grad_norm(self):
        return self._grad_norm

    # ---- names of the reference's strategy queries / contexts (lms/base_lm.py:48-57, 91-93, 102-114, 195-222).
    #      Both strategies run on the same native engine here, so the queries describe the sharding MODE.
    @property
    def is_using_ds_z3(self) -> bool:
        from ..lightning.strategy import DeepSpeedStrategy
        return isinstance(self.strategy, DeepSpeedStrategy) and self.strategy.zero_stage_3

    @property
    def is_using_fsdp(self) -> bool:
        from ..lightning.strategy import FSDP2Strategy
        return isinstance(self.strategy, FSDP2Strategy)

    @property
    def should_sync_weights(self) -> bool:
        return self.should_load_pre_trained_weights or self.should_initialize_weights

    @staticmethod
    def shutdown_ds_init_context():
        """DeepSpeed's ZeRO-3 init context has no counterpart: parameters are built whole and sharded in ``setup``."""
        import contextlib
        return contextlib.nullcontext()

    def get_pre_trained_weights_context(self):
        from ..utils.context_managers import ContextManagers
        return ContextManagers([])

    def configure_model_context(self):
        """Context ``configure_model`` runs in: weight-init switch (+ the strategy's device / dtype context, which
        the trainer enters around it)."""
        return BaseModel.init_weights_context(self.should_initialize_weights)

    def fsdp_parallelize_context(self):
        """Records tied weights before parallelisation and re-ties them afterwards (reference :203-222)."""
        import contextlib

        @contextlib.contextmanager
        def ctx():
            tied = self._record_tied_weights()
            yield
            self._retie(tied)
        return ctx()

    def get_model(self) -> BaseModel:
        raise NotImplementedError

    def get_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        import safetensors.torch
        return safetensors.torch.load_file(self.config.pre_trained_weights)

    # ---------------------------------------------------------------- after configure_model
    def on_fsdp_parallelize_model(self, **kwargs: Any) -> None:
        raise NotImplementedError

    def on_after_configure_model(self) -> None:
        if self.config.frozen_modules:
            for name, module in self.named_modules():
                if any(re.search(pattern, name) for pattern in self.config.frozen_modules):
                    if self.global_rank == 0:
                        logger.info(f'Freeze `{name}`')
                    module.eval().requires_grad_(False)
        strategy = self.strategy
        if strategy is not None:
            strategy._materialize(self)
            tied = self._record_tied_weights()
            self.on_fsdp_parallelize_model(
                dp_mesh=strategy.dp_mesh, tp_mesh=strategy.tp_mesh,
                reshard_after_forward=strategy.reshard_after_forward, mp_policy=strategy.mp_policy,
                offload_policy=strategy.offload_policy)
            self._retie(tied)
        if self.should_load_pre_trained_weights:
            self.load_pre_trained_weights()
        elif strategy is not None and strategy.world_size > 1 and not self.is_loading_from_checkpoint:
            self._broadcast_initial_weights()

    def _record_tied_weights(self) -> list[tuple[str, str]]:
        seen: dict[int, str] = {}
        tied = []
        for name, p in self.named_parameters(remove_duplicate=False):
            if id(p) in seen:
                tied.append((name, seen[id(p)]))
            else:
                seen[id(p)] = name
        return tied

    def _retie(self, tied: list[tuple[str, str]]) -> None:
        for alias, original in tied:
            module_name, _, attr = alias.rpartition('.')
            self.get_submodule(module_name).register_parameter(attr, self.get_parameter(original))

    def _broadcast_initial_weights(self) -> None:
        """Random init differs per rank: rank 0 of every dp group is the source of truth."""
        import torch.distributed as dist
        if not dist.is_initialized():
            return
        group = self.strategy.dp_mesh.group if self.strategy.dp_size > 1 else None
        if group is None:
            return
        src = self.strategy.dp_mesh.ranks[0]
        for p in self.parameters():
            dist.broadcast(p.data, src=src, group=group)

    def load_pre_trained_weights(self) -> None:
        """Rank 0 reads, every tensor is broadcast (and sliced for TP) one at a time
        (reference: base_lm.py:147-193)."""
        import torch.distributed as dist
        strategy = self.strategy
        distributed = strategy is not None and strategy.world_size > 1 and dist.is_initialized()
        state = self.get_pre_trained_weights() if (not distributed or self.global_rank == 0) else None
        own = self.state_dict(keep_vars=True)
        if not distributed:
            missing = [k for k in own if k not in state and own[k].requires_grad]
            with torch.no_grad():
                for name, target in own.items():
                    if name in state:
                        target.copy_(state[name].to(target.device, dtype=target.dtype))
            if missing:
                logger.warning(f'pre-trained weights lack {len(missing)} tensors, e.g. {missing[:3]}')
            return
        names = strategy.broadcast(sorted(state.keys()) if state is not None else None)
        available = set(names)
        device = strategy.root_device
        with torch.no_grad():
            for name, target in own.items():
                if name not in available:
                    continue
                info = getattr(target, '_tp_info', None)
                shape = info.full_shape if info is not None and strategy.tp_size > 1 else target.shape
                buf = torch.empty(shape, dtype=target.dtype, device=device)
                if self.global_rank == 0:
                    buf.copy_(state[name].to(device, dtype=target.dtype))
                dist.broadcast(buf, src=0)
                if info is not None and strategy.tp_size > 1:
                    buf = info.scatter(buf)
                target.copy_(buf)

    # ---------------------------------------------------------------- optimisation
    def configure_optimizers(self):
        optim = self.config.optim
        assert optim is not None, '`optim` must be configured for training'
        params = [p for p in self.parameters() if p.requires_grad]
        optimizer = optim.optimizer_class(params, **optim.optimizer_kwargs)
        sched_class = optim.lr_scheduler_class
        kwargs: dict[str, Any] = {}
        if 'num_total_steps' in inspect.signature(sched_class).parameters:
            kwargs['num_total_steps'] = self.trainer.estimated_stepping_batches
        kwargs |= optim.lr_scheduler_kwargs
        scheduler = sched_class(optimizer, **kwargs)
        return {'optimizer': optimizer, 'lr_scheduler': {'scheduler': scheduler, 'interval': 'step'}}

    def optimizer_step(self, epoch: int, batch_idx: int, optimizer, optimizer_closure=None) -> None:
        super().optimizer_step(epoch, batch_idx, optimizer, optimizer_closure)
        if self.config.log_grad_norm and self.grad_norm is not None:
            self.log('Gradient Norm', self.grad_norm)

    @property
    def required_keys(self) -> set[str]:
        model = self.get_model()
        for name, child in self.named_children():
            if child is model:
                return set(model.state_dict(prefix=name + '.').keys())
        raise RuntimeError('failed to infer the model prefix')
