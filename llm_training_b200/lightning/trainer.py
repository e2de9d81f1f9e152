"""The training runtime: fit loop, gradient accumulation, validation, callbacks, logging, checkpoints.

Replaces ``lightning.Trainer`` + the reference's subclass (src/llm_training/lightning/cli/trainer.py:4-11)
with the subset of behaviour the reference's configs use (``strategy, precision, logger,
max_epochs, val_check_interval, accumulate_grad_batches, gradient_clip_val, callbacks, num_nodes``).
Hot-loop rules: no per-step host synchronisation — logged tensors are kept on device and
only converted when a logger / progress bar actually consumes them.
"""
import logging
import math
import os
import random
import subprocess
import sys
from typing import Any, Iterable

import numpy as np
import torch

from .callbacks.base import Callback
from .module import LightningDataModule, LightningModule
from .strategy.base import SingleDeviceStrategy, Strategy
from .strategy.precision import FSDP2Precision

logger = logging.getLogger(__name__)


def seed_everything(seed: int | None = None, workers: bool = False) -> int:
    if seed is None:
        seed = int(os.environ.get('PL_GLOBAL_SEED', random.randrange(0, 2 ** 31 - 1)))
    os.environ['PL_GLOBAL_SEED'] = str(seed)
    random.seed(seed)
    np.random.seed(seed % (2 ** 32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    return seed


class FitLoop:
    """Progress counters that are checkpointed (``trainer.fit_loop.batch_idx`` drives the resumable sampler)."""

    def __init__(self) -> None:
        self.epoch = 0
        self.batch_idx = 0            # number of batches already consumed in the current epoch
        self.global_step = 0          # optimizer steps
        self.total_batches = 0

    def state_dict(self) -> dict[str, int]:
        return dict(epoch=self.epoch, batch_idx=self.batch_idx, global_step=self.global_step,
                    total_batches=self.total_batches)

    def load_state_dict(self, state: dict[str, int]) -> None:
        self.epoch = int(state.get('epoch', 0))
        self.batch_idx = int(state.get('batch_idx', 0))
        self.global_step = int(state.get('global_step', 0))
        self.total_batches = int(state.get('total_batches', 0))


class Trainer:
    def __init__(self, strategy: Strategy | str | None = None, precision: str | int | None = '32-true',
                 logger: Any | bool | Iterable[Any] | None = None, callbacks: list[Callback] | Callback | None = None,
                 max_epochs: int | None = None, max_steps: int = -1, val_check_interval: int | float | None = None,
                 check_val_every_n_epoch: int | None = 1, accumulate_grad_batches: int = 1,
                 gradient_clip_val: float | None = None, gradient_clip_algorithm: str | None = None,
                 num_nodes: int = 1, devices: int | str | list[int] = 'auto', accelerator: str = 'auto',
                 limit_train_batches: int | float | None = None, limit_val_batches: int | float | None = None,
                 log_every_n_steps: int = 50, enable_checkpointing: bool = True, enable_progress_bar: bool = True,
                 default_root_dir: str | None = None, num_sanity_val_steps: int = 0,
                 enable_model_summary: bool = False, deterministic: bool | str | None = None, benchmark: bool | None = None,
                 use_distributed_sampler: bool = True, **unused: Any) -> None:
        if unused:
            logger_ = logging.getLogger(__name__)
            logger_.warning(f'Trainer: ignoring unsupported arguments {sorted(unused)}')
        self.accelerator = accelerator
        self.num_nodes = num_nodes
        self.devices = devices
        if strategy is None or isinstance(strategy, str):
            strategy = SingleDeviceStrategy()
        self.strategy: Strategy = strategy
        self.strategy.num_nodes = num_nodes
        self.strategy.trainer = self
        self.precision = str(precision if precision is not None else '32-true')
        self.strategy.precision_plugin = FSDP2Precision(self.precision)

        if logger is None or logger is True:
            from .loggers import CSVLogger
            self.loggers = [CSVLogger(default_root_dir or os.getcwd())] if logger is True else []
        elif logger is False:
            self.loggers = []
        else:
            self.loggers = list(logger) if isinstance(logger, (list, tuple)) else [logger]

        cbs = [] if callbacks is None else (list(callbacks) if isinstance(callbacks, (list, tuple)) else [callbacks])
        self.callbacks: list[Callback] = cbs
        self.enable_checkpointing = enable_checkpointing
        self.enable_progress_bar = enable_progress_bar
        if enable_progress_bar and not any(type(c).__name__ == 'TQDMProgressBar' for c in self.callbacks):
            from .callbacks import TQDMProgressBar
            self.callbacks.append(TQDMProgressBar())

        self.max_epochs = max_epochs if max_epochs is not None else (1000 if max_steps == -1 else -1)
        self.max_steps = max_steps
        self.val_check_interval = val_check_interval
        self.check_val_every_n_epoch = check_val_every_n_epoch
        self.accumulate_grad_batches = max(1, int(accumulate_grad_batches))
        self.gradient_clip_val = gradient_clip_val
        self.limit_train_batches = limit_train_batches
        self.limit_val_batches = limit_val_batches
        self.log_every_n_steps = max(1, log_every_n_steps)
        self.default_root_dir = default_root_dir or os.getcwd()
        self.use_distributed_sampler = use_distributed_sampler
        # Lightning semantics: True -> torch.use_deterministic_algorithms(True), 'warn' -> warn_only.  The in-tree
        # kernels read the same switch: attention backward orders its dQ reduce-adds, the gradient reduce-scatter sums
        # the ranks in rank order instead of inside the switch; everything else is order-independent by construction.
        self.deterministic = deterministic
        if deterministic:
            os.environ.setdefault('CUBLAS_WORKSPACE_CONFIG', ':4096:8')
            torch.use_deterministic_algorithms(True, warn_only=(deterministic == 'warn'))
        if benchmark is not None:
            torch.backends.cudnn.benchmark = bool(benchmark)

        self.fit_loop = FitLoop()
        self.lightning_module: LightningModule | None = None
        self.datamodule: LightningDataModule | None = None
        self.optimizers: list = []
        self.lr_scheduler_configs: list[dict] = []
        self.ckpt_path: str | None = None
        self.should_stop = False
        self.training = False
        self.validating = False
        self.callback_metrics: dict[str, Any] = {}
        self.progress_bar_metrics: dict[str, Any] = {}
        self._pending_logs: dict[str, Any] = {}
        self._val_accum: dict[str, list] = {}
        self.train_dataloader = None
        self.num_training_batches: int | float = float('inf')
        self.state_fn = 'fit'

    # ------------------------------------------------------------------ properties
    @property
    def logger(self):
        return self.loggers[0] if self.loggers else None

    @property
    def log_dir(self) -> str:
        lg = self.logger
        if lg is not None and getattr(lg, 'log_dir', None):
            return lg.log_dir
        return self.default_root_dir

    @property
    def global_rank(self) -> int:
        return self.strategy.global_rank

    @property
    def local_rank(self) -> int:
        return self.strategy.local_rank

    @property
    def world_size(self) -> int:
        return self.strategy.world_size

    @property
    def is_global_zero(self) -> bool:
        return self.strategy.is_global_zero

    @property
    def num_devices(self) -> int:
        return max(1, self.world_size // max(1, self.num_nodes))

    @property
    def global_step(self) -> int:
        return self.fit_loop.global_step

    @property
    def current_epoch(self) -> int:
        return self.fit_loop.epoch

    @property
    def checkpoint_callbacks(self) -> list:
        return [c for c in self.callbacks if type(c).__name__ == 'ModelCheckpoint']

    @property
    def estimated_stepping_batches(self) -> int | float:
        """Total optimizer steps of this fit (used for ``num_total_steps`` of the LR schedules).  Builds a
        throw-away dataloader so nothing created before checkpoint restore is left behind
        (reference: src/llm_training/lightning/cli/trainer.py:4-11)."""
        if self.max_steps and self.max_steps > 0:
            return self.max_steps
        n = self.num_training_batches
        if n == float('inf'):
            loader = self._make_train_dataloader()
            n = self._limit(len(loader), self.limit_train_batches)
            del loader
        if self.max_epochs is None or self.max_epochs < 0:
            return float('inf')
        return math.ceil(n / self.accumulate_grad_batches) * max(self.max_epochs, 1)

    # ------------------------------------------------------------------ logging
    def _log_metric(self, name: str, value: Any, prog_bar: bool = False, logger: bool = True,
                    sync_dist: bool = False, batch_size: int | None = None) -> None:
        if isinstance(value, torch.Tensor):
            value = value.detach()
        if self.validating:
            self._val_accum.setdefault(name, []).append((value, batch_size or 1, sync_dist))
            return
        self.callback_metrics[name] = value
        if prog_bar:
            self.progress_bar_metrics[name] = value
        if logger:
            self._pending_logs[name] = value

    @staticmethod
    def _resolve(value: Any) -> float:
        from ..metrics import Metric
        if isinstance(value, Metric):
            value = value.compute()
        if isinstance(value, torch.Tensor):
            return float(value.float().mean().item())
        return float(value)

    def _flush_logs(self, force: bool = False) -> None:
        if not self._pending_logs:
            return
        if not force and (self.global_step % self.log_every_n_steps != 0):
            return
        metrics = {k: self._resolve(v) for k, v in self._pending_logs.items()}
        self._pending_logs.clear()
        if self.is_global_zero:
            for lg in self.loggers:
                lg.log_metrics(metrics, step=self.global_step)

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _limit(n: int, limit: int | float | None) -> int:
        if limit is None:
            return n
        if isinstance(limit, float) and limit <= 1.0:
            return max(1, int(n * limit))
        return min(n, int(limit))

    def _call(self, hook: str, *args: Any, **kwargs: Any) -> None:
        for cb in self.callbacks:
            fn = getattr(cb, hook, None)
            if fn is not None:
                fn(self, self.lightning_module, *args, **kwargs)

    def _make_train_dataloader(self):
        source = self.datamodule if self.datamodule is not None else self.lightning_module
        return source.train_dataloader()

    def _make_val_dataloader(self):
        source = self.datamodule if self.datamodule is not None else self.lightning_module
        fn = getattr(source, 'val_dataloader', None)
        return fn() if fn is not None else None

    # ------------------------------------------------------------------ process launch
    def _maybe_launch(self) -> bool:
        """One process per GPU.  If several devices are requested and we were not started by torchrun /
        srun, re-execute this command once per local rank (reference: Lightning's subprocess launcher,
        fsdp2_strategy.py:169-173).  Returns True in the parent after the children finished."""
        if 'RANK' in os.environ or 'LLMT_LAUNCHED' in os.environ:
            return False
        if 'SLURM_PROCID' in os.environ and int(os.environ.get('SLURM_NTASKS', '1')) > 1:
            os.environ.setdefault('RANK', os.environ['SLURM_PROCID'])
            os.environ.setdefault('WORLD_SIZE', os.environ['SLURM_NTASKS'])
            os.environ.setdefault('LOCAL_RANK', os.environ.get('SLURM_LOCALID', '0'))
            return False
        if self.devices == 'auto' or self.devices == -1:
            n = torch.cuda.device_count() if (torch.cuda.is_available() and self.accelerator != 'cpu') else 1
        elif isinstance(self.devices, (list, tuple)):
            n = len(self.devices)
        else:
            n = int(self.devices)
        if n <= 1 or self.num_nodes > 1:
            return False
        port = str(29500 + random.randrange(0, 2000))
        procs = []
        for local_rank in range(n):
            env = dict(os.environ, RANK=str(local_rank), LOCAL_RANK=str(local_rank), WORLD_SIZE=str(n),
                       MASTER_ADDR='127.0.0.1', MASTER_PORT=port, LLMT_LAUNCHED='1')
            procs.append(subprocess.Popen([sys.executable] + sys.argv, env=env))
        codes = [p.wait() for p in procs]
        if any(codes):
            raise SystemExit(max(codes))
        return True

    # ------------------------------------------------------------------ fit
    def fit(self, model: LightningModule, datamodule: LightningDataModule | None = None,
            ckpt_path: str | None = None, train_dataloaders: Any | None = None) -> None:
        if self._maybe_launch():
            return
        self.lightning_module = model
        self.datamodule = datamodule
        self.ckpt_path = ckpt_path
        model.trainer = self
        if datamodule is not None:
            datamodule.trainer = self
        if train_dataloaders is not None:
            model.train_dataloader = lambda: train_dataloaders   # type: ignore[assignment]
        self.strategy.setup_environment(self)

        if datamodule is not None:
            if self.local_rank == 0 or not getattr(datamodule, 'prepare_data_per_node', True) and self.is_global_zero:
                datamodule.prepare_data()
            self.strategy.barrier()
            datamodule.setup('fit')

        self._call('setup', stage='fit')
        with self.strategy.model_init_context():
            model.configure_model()
        self.strategy.setup(self)
        if ckpt_path is not None:
            self._restore(ckpt_path)

        self._call('on_fit_start')
        model.on_fit_start()
        try:
            self._run_training()
        finally:
            if hasattr(self.strategy, 'synchronize_optimizer'):
                self.strategy.synchronize_optimizer()
            self._flush_logs(force=True)
            self._call('on_fit_end')
            model.on_fit_end()
            self._call('teardown', stage='fit')
            for lg in self.loggers:
                if hasattr(lg, 'finalize'):
                    lg.finalize('success')
            self.strategy.teardown()

    def _reached_max_steps(self) -> bool:
        return self.max_steps is not None and self.max_steps > 0 and self.global_step >= self.max_steps

    def _run_training(self) -> None:
        model = self.lightning_module
        self.training = True
        self._call('on_train_start')
        model.on_train_start()
        while not self.should_stop and not self._reached_max_steps():
            if self.max_epochs is not None and self.max_epochs >= 0 and self.fit_loop.epoch >= self.max_epochs:
                break
            self._run_epoch()
            if self.should_stop or self._reached_max_steps():
                break
            self.fit_loop.epoch += 1
            self.fit_loop.batch_idx = 0
        self._call('on_train_end')
        model.on_train_end()
        self.training = False

    def _run_epoch(self) -> None:
        model = self.lightning_module
        model.train()
        loader = self._make_train_dataloader()
        self.train_dataloader = loader
        sampler = getattr(loader, 'sampler', None)
        if hasattr(sampler, 'set_epoch'):
            sampler.set_epoch(self.fit_loop.epoch)
        n_batches = self._limit(len(loader), self.limit_train_batches)
        self.num_training_batches = n_batches
        self._call('on_train_epoch_start')
        model.on_train_epoch_start()
        accum = self.accumulate_grad_batches
        device = self.strategy.root_device
        start_idx = self.fit_loop.batch_idx         # > 0 when resuming mid-epoch: the sampler skips those
        optimizer = self.optimizers[0] if self.optimizers else None
        val_every = None
        if isinstance(self.val_check_interval, int) and not isinstance(self.val_check_interval, bool):
            val_every = self.val_check_interval
        elif isinstance(self.val_check_interval, float):
            val_every = max(1, int(n_batches * self.val_check_interval))

        for offset, batch in enumerate(loader):
            batch_idx = start_idx + offset
            if batch_idx >= n_batches:
                break
            batch = model.transfer_batch_to_device(batch, device)
            self._call('on_train_batch_start', batch, batch_idx)
            model.on_train_batch_start(batch, batch_idx)
            boundary = ((batch_idx + 1) % accum == 0) or (batch_idx + 1 == n_batches)
            if boundary:
                loss = self.strategy.training_step(batch, batch_idx)
                self.strategy.backward(loss / accum if accum > 1 else loss)
            else:
                with self.strategy.block_backward_sync():
                    loss = self.strategy.training_step(batch, batch_idx)
                    self.strategy.backward(loss / accum)
            self.fit_loop.batch_idx = batch_idx + 1
            self.fit_loop.total_batches += 1
            if boundary and optimizer is not None:
                self._call('on_before_optimizer_step', optimizer)
                model.optimizer_step(self.fit_loop.epoch, batch_idx, _StrategyStepper(self.strategy, optimizer))
                model.optimizer_zero_grad(self.fit_loop.epoch, batch_idx, optimizer)
                self.fit_loop.global_step += 1
                for cfg in self.lr_scheduler_configs:
                    if cfg.get('interval', 'step') == 'step' and self.global_step % cfg.get('frequency', 1) == 0:
                        cfg['scheduler'].step()
            self._call('on_train_batch_end', loss, batch, batch_idx)
            model.on_train_batch_end(loss, batch, batch_idx)
            if boundary:
                self._flush_logs()
            if val_every is not None and (batch_idx + 1) % val_every == 0:
                self._run_validation()
                model.train()
            if self.should_stop or self._reached_max_steps():
                break

        for cfg in self.lr_scheduler_configs:
            if cfg.get('interval') == 'epoch':
                cfg['scheduler'].step()
        if val_every is None and self.check_val_every_n_epoch and \
                (self.fit_loop.epoch + 1) % self.check_val_every_n_epoch == 0 and not self.should_stop:
            self._run_validation()
        self._call('on_train_epoch_end')
        model.on_train_epoch_end()

    # ------------------------------------------------------------------ validation
    @torch.no_grad()
    def _run_validation(self) -> None:
        loader = self._make_val_dataloader()
        if loader is None:
            return
        model = self.lightning_module
        if hasattr(self.strategy, 'synchronize_optimizer'):
            self.strategy.synchronize_optimizer()
        model.eval()
        self.validating = True
        self._val_accum = {}
        self._call('on_validation_start')
        model.on_validation_epoch_start()
        n = self._limit(len(loader), self.limit_val_batches)
        device = self.strategy.root_device
        for i, batch in enumerate(loader):
            if i >= n:
                break
            batch = model.transfer_batch_to_device(batch, device)
            self.strategy.validation_step(batch, i)
        self.validating = False
        metrics: dict[str, float] = {}
        from ..metrics import Metric
        for name, entries in self._val_accum.items():
            last = entries[-1][0]
            if isinstance(last, Metric):
                metrics[name] = self._resolve(last)
                last.reset()
                continue
            total = sum(self._resolve(v) * bs for v, bs, _ in entries)
            weight = sum(bs for _, bs, _ in entries)
            value = torch.tensor([total, float(weight)], dtype=torch.float64)
            if any(sd for _, _, sd in entries):
                value = self.strategy.reduce(value.float(), reduce_op='sum').double().cpu()
            metrics[name] = float(value[0] / max(value[1], 1e-12))
        self.callback_metrics.update(metrics)
        if self.is_global_zero:
            for lg in self.loggers:
                lg.log_metrics(metrics, step=self.global_step)
        model.on_validation_epoch_end()
        self._call('on_validation_end')

    def validate(self, model: LightningModule, datamodule: LightningDataModule | None = None) -> dict[str, Any]:
        self.lightning_module, self.datamodule = model, datamodule
        model.trainer = self
        if datamodule is not None:
            datamodule.trainer = self
        self.strategy.setup_environment(self)
        if datamodule is not None:
            datamodule.prepare_data()
            datamodule.setup('validate')
        with self.strategy.model_init_context():
            model.configure_model()
        self.strategy._materialize(model)
        self._run_validation()
        return dict(self.callback_metrics)

    # ------------------------------------------------------------------ checkpoints
    def dump_checkpoint(self) -> dict[str, Any]:
        model = self.lightning_module
        if hasattr(self.strategy, 'synchronize_optimizer'):
            self.strategy.synchronize_optimizer()
        ckpt: dict[str, Any] = {
            'epoch': self.fit_loop.epoch,
            'global_step': self.fit_loop.global_step,
            'loops': {'fit_loop': self.fit_loop.state_dict()},
            'callbacks': {self._cb_key(cb): cb.state_dict() for cb in self.callbacks if cb.state_dict()},
            'lr_schedulers': [cfg['scheduler'].state_dict() for cfg in self.lr_scheduler_configs],
            'precision': self.strategy.precision_plugin.state_dict(),
            'llm_training_b200_version': __import__('llm_training_b200').__version__,
        }
        if self.datamodule is not None:
            ckpt['datamodule'] = self.datamodule.state_dict()
        model.on_save_checkpoint(ckpt)
        self._call('on_save_checkpoint', ckpt)
        return ckpt

    @staticmethod
    def _cb_key(cb: Callback) -> str:
        return getattr(cb, 'state_key', type(cb).__qualname__)

    def save_checkpoint(self, filepath: str, weights_only: bool = False) -> None:
        ckpt = self.dump_checkpoint()
        self.strategy.save_checkpoint(ckpt, filepath)
        self.strategy.barrier()

    def _restore(self, path: str) -> None:
        ckpt = self.strategy.load_checkpoint(path)
        self.fit_loop.load_state_dict(ckpt.get('loops', {}).get('fit_loop', {
            'epoch': ckpt.get('epoch', 0), 'global_step': ckpt.get('global_step', 0)}))
        for cfg, state in zip(self.lr_scheduler_configs, ckpt.get('lr_schedulers', [])):
            cfg['scheduler'].load_state_dict(state)
        self.strategy.precision_plugin.load_state_dict(ckpt.get('precision', {}))
        saved_cb = ckpt.get('callbacks', {})
        for cb in self.callbacks:
            key = self._cb_key(cb)
            if key in saved_cb:
                cb.load_state_dict(saved_cb[key])
        if self.datamodule is not None and 'datamodule' in ckpt:
            self.datamodule.load_state_dict(ckpt['datamodule'])
        self.lightning_module.on_load_checkpoint(ckpt)
        self._call('on_load_checkpoint', ckpt)
        logger.info(f'Restored checkpoint {path} (epoch {self.fit_loop.epoch}, step {self.fit_loop.global_step}, '
                    f'batch {self.fit_loop.batch_idx})')


class _StrategyStepper:
    """What ``LightningModule.optimizer_step`` receives: ``.step()`` goes through the strategy so that the
    gradient norm is captured (reference: fsdp2_strategy.py:301-304)."""

    def __init__(self, strategy: Strategy, optimizer: Any) -> None:
        self._strategy, self._optimizer = strategy, optimizer

    def step(self, closure=None):
        return self._strategy.optimizer_step(self._optimizer, closure)

    def __getattr__(self, name: str) -> Any:
        return getattr(self._optimizer, name)
