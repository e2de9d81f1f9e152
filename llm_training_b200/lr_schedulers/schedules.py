"""Learning-rate schedules with linear warm-up.

Same class names and constructor arguments as the reference
(src/llm_training/lr_schedulers/{warmup,constant,cosine,linear}.py) but every schedule is a
*closed-form* function of the step counter instead of a wrapper that drives a second torch
scheduler.  That makes ``state_dict`` trivially resumable (only ``last_epoch`` matters) and lets
the fused optimizer kernel read the next LR without extra Python objects.

Value parity with the reference (checked in tests/test_core.py::test_lr_schedules_values_and_resume and the warm-up cases of tests/test_reference_oracle.py):
  * warm-up:  lr(t) = base * (t + 1) / W                 for t <  W      (warmup.py:25-28)
  * constant: lr(t) = base * factor if (t-W) < total_iters else base     (constant.py:7-26)
  * cosine:   lr(t) = min + (base-min) * (1+cos(pi*(t-W)/(N-W)))/2       (cosine.py:7-26)
  * linear:   warm-up uses (t+1)/(W+1); decay is linear from base at t=W towards
              ``min_lr`` at t=N                                          (linear.py:25-39)
"""
import math
from typing import Callable

from torch.optim import Optimizer
from torch.optim.lr_scheduler import LRScheduler


class WarmupLR(LRScheduler):
    """Linear warm-up for ``num_warmup_epochs`` steps, then ``after_warmup(step - W, base_lr)``.

    ``lr_scheduler`` may be a torch ``LRScheduler`` (reference-compatible call style); it is then
    stepped after warm-up exactly like the reference does.
    """

    def __init__(
        self,
        optimizer: Optimizer,
        lr_scheduler: LRScheduler | Callable[[int, float], float] | None = None,
        num_warmup_epochs: int = 0,
        last_epoch: int = -1,
    ) -> None:
        self.num_warmup_epochs = int(num_warmup_epochs)
        self._delegate = lr_scheduler if isinstance(lr_scheduler, LRScheduler) else None
        self._fn = lr_scheduler if (lr_scheduler is not None and self._delegate is None) else None
        super().__init__(optimizer, last_epoch)

    # -- schedule -----------------------------------------------------------------------------
    def after_warmup(self, t: int, base_lr: float) -> float:
        if self._fn is not None:
            return self._fn(t, base_lr)
        return base_lr

    def lr_at(self, step: int, base_lr: float) -> float:
        w = self.num_warmup_epochs
        if step < w:
            return base_lr * (step + 1) / w
        return self.after_warmup(step - w, base_lr)

    def get_lr(self) -> list[float]:
        if self._delegate is not None and self.last_epoch >= self.num_warmup_epochs:
            return self._delegate.get_last_lr()
        return [self.lr_at(self.last_epoch, lr) for lr in self.base_lrs]

    def step(self, epoch: int | None = None) -> None:
        if self._delegate is not None and self.last_epoch >= self.num_warmup_epochs:
            if self.last_epoch == self.num_warmup_epochs:
                self._delegate.base_lrs = list(self.base_lrs)
            self._delegate.step()
        super().step(epoch)

    # -- (de)serialisation ----------------------------------------------------------------------
    def state_dict(self) -> dict:
        skip = {'optimizer', '_delegate', '_fn'}
        state = {k: v for k, v in self.__dict__.items() if k not in skip}
        if self._delegate is not None:
            state['lr_scheduler_state_dict'] = self._delegate.state_dict()
        return state

    def load_state_dict(self, state_dict: dict) -> None:
        state_dict = dict(state_dict)
        inner = state_dict.pop('lr_scheduler_state_dict', None)
        if inner is not None and self._delegate is not None:
            self._delegate.load_state_dict(inner)
        self.__dict__.update(state_dict)


class ConstantWarmupLR(WarmupLR):
    def __init__(
        self,
        optimizer: Optimizer,
        factor: float = 1.0,
        total_iters: int = 0,
        num_warmup_steps: int = 0,
        last_epoch: int = -1,
    ) -> None:
        self.factor = factor
        self.total_iters = total_iters
        super().__init__(optimizer, None, num_warmup_steps, last_epoch)

    def after_warmup(self, t: int, base_lr: float) -> float:
        return base_lr * self.factor if t < self.total_iters else base_lr


class CosineAnnealingWarmupLR(WarmupLR):
    def __init__(
        self,
        optimizer: Optimizer,
        num_warmup_steps: int,
        num_total_steps: int,
        min_lr: float,
        last_epoch: int = -1,
    ) -> None:
        self.num_total_steps = int(num_total_steps)
        self.min_lr = float(min_lr)
        super().__init__(optimizer, None, num_warmup_steps, last_epoch)

    def after_warmup(self, t: int, base_lr: float) -> float:
        t_max = max(1, self.num_total_steps - self.num_warmup_epochs)
        return self.min_lr + (base_lr - self.min_lr) * (1.0 + math.cos(math.pi * t / t_max)) / 2.0


class LinearWarmupLR(LRScheduler):
    def __init__(
        self,
        optimizer: Optimizer,
        num_warmup_steps: int,
        num_total_steps: int,
        min_lr: float | list[float],
        last_epoch: int = -1,
    ) -> None:
        self.num_warmup_steps = int(num_warmup_steps)
        self.num_total_steps = int(num_total_steps)
        self.min_lr = min_lr
        super().__init__(optimizer, last_epoch=last_epoch)
        assert isinstance(min_lr, (int, float)) or len(min_lr) == len(self.base_lrs)

    @property
    def min_lrs(self) -> list[float]:
        if isinstance(self.min_lr, (int, float)):
            return [float(self.min_lr)] * len(self.base_lrs)
        return list(self.min_lr)

    def get_lr(self) -> list[float]:
        t, w, n = self.last_epoch, self.num_warmup_steps, self.num_total_steps
        if t < w:
            return [(t + 1) / (w + 1) * lr for lr in self.base_lrs]
        remaining = (n - t) / max(1, n - w)
        return [floor + (lr - floor) * remaining for lr, floor in zip(self.base_lrs, self.min_lrs)]
