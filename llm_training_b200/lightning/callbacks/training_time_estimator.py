"""Times optimizer steps ``[num_warmup_steps, num_test_steps)``, extrapolates the full run, stops training
(reference: src/llm_training/lightning/callbacks/training_time_estimator.py:12-83).  Unlike the reference's
wall clock around un-synchronised steps, the interval is bracketed by device synchronisation and reduced as
the max over ranks; tokens/s is reported when the LM tracks consumed tokens."""
import logging
import time

import torch

from .base import Callback

logger = logging.getLogger(__name__)


class TrainingTimeEstimator(Callback):
    def __init__(self, num_test_steps: int, num_warmup_steps: int = 2, enable_checkpointing: bool = False) -> None:
        assert 0 <= num_warmup_steps < num_test_steps
        self.num_test_steps = num_test_steps
        self.num_warmup_steps = num_warmup_steps
        self.enable_checkpointing = enable_checkpointing
        self.start_time = self.end_time = 0.0
        self._started = False
        self.steps_per_second: float | None = None

    def setup(self, trainer, pl_module, stage: str) -> None:
        if not self.enable_checkpointing:
            if trainer.is_global_zero:
                logger.info('Disabling checkpointing')
            trainer.callbacks = [c for c in trainer.callbacks if type(c).__name__ != 'ModelCheckpoint']

    @staticmethod
    def _sync(trainer) -> None:
        if trainer.strategy.root_device.type == 'cuda':
            torch.cuda.synchronize(trainer.strategy.root_device)

    def on_train_batch_start(self, trainer, pl_module, batch, batch_idx) -> None:
        if trainer.global_step == self.num_warmup_steps and not self._started:
            self._sync(trainer)
            self._started = True
            self.start_time = time.perf_counter()

    def on_train_batch_end(self, trainer, pl_module, outputs, batch, batch_idx) -> None:
        if trainer.global_step == self.num_test_steps and not trainer.should_stop:
            self._sync(trainer)
            self.end_time = time.perf_counter()
            trainer.should_stop = True

    def on_fit_end(self, trainer, pl_module) -> None:
        seconds = float(trainer.strategy.reduce(torch.tensor(self.end_time - self.start_time), reduce_op='max'))
        if seconds <= 0:
            return
        self.print_estimated_training_time(trainer, seconds)

    def print_estimated_training_time(self, trainer, seconds: float) -> None:
        steps = self.num_test_steps - self.num_warmup_steps
        self.steps_per_second = steps / seconds
        total_steps = trainer.estimated_stepping_batches
        estimate = total_steps / self.steps_per_second
        for unit_seconds, unit in ((86400, 'days'), (3600, 'hours'), (60, 'minutes'), (1, 'seconds')):
            if estimate >= unit_seconds or unit == 'seconds':
                pretty = f'{estimate / unit_seconds:.2f} {unit}'
                break
        if trainer.is_global_zero:
            from tabulate import tabulate
            print(tabulate([['Running Time', f'{seconds:.2f} seconds'],
                            ['Steps per second', f'{self.steps_per_second:.2f} steps'],
                            ['Estimated training time', pretty]], tablefmt='fancy_grid'))
