"""CUDA-event step timing, max over ranks (the reference's ``TrainingTimeEstimator`` uses wall clock
without a device synchronisation, training_time_estimator.py:63-77)."""
from typing import Any

import torch

from .base import Callback


class StepTimeMonitor(Callback):
    """Trainer callback: device time of every ``every_n_steps``-th window of optimizer steps, CUDA events on the
    training stream, reduced with MAX over ranks (the measurement rule of BASELINE.md §3), logged as
    ``step_time_ms`` / ``tokens_per_sec``."""

    def __init__(self, every_n_steps: int = 10, tokens_per_step: int | None = None) -> None:
        self.every = max(1, every_n_steps)
        self.tokens_per_step = tokens_per_step
        self._start: Any = None
        self._count = 0
        self.last_ms: float | None = None

    def on_train_batch_start(self, trainer, pl_module, batch, batch_idx) -> None:
        if trainer.strategy.root_device.type != 'cuda':
            return
        if self._start is None:
            self._start = torch.cuda.Event(enable_timing=True)
            self._start.record()
            self._count = 0

    def on_train_batch_end(self, trainer, pl_module, outputs, batch, batch_idx) -> None:
        if self._start is None:
            return
        self._count += 1
        if self._count < self.every:
            return
        if hasattr(trainer.strategy, 'synchronize_optimizer'):
            trainer.strategy.synchronize_optimizer()
        end = torch.cuda.Event(enable_timing=True)
        end.record()
        end.synchronize()
        ms = self._start.elapsed_time(end) / self._count
        ms = float(trainer.strategy.reduce(torch.tensor(ms), reduce_op='max'))
        self.last_ms = ms
        metrics = {'step_time_ms': ms}
        if self.tokens_per_step:
            metrics['tokens_per_sec'] = self.tokens_per_step * trainer.strategy.dp_size / ms * 1e3
        pl_module.log_dict(metrics) if hasattr(pl_module, 'log_dict') else None
        self._start = None

    def state_dict(self) -> dict:
        return {}

    def load_state_dict(self, state: dict) -> None:
        pass
