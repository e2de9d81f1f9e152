"""Logs the learning rate of every optimizer / param group each step (Lightning's ``LearningRateMonitor``,
used by every example config of the reference)."""
from .base import Callback


class LearningRateMonitor(Callback):
    def __init__(self, logging_interval: str | None = 'step', log_momentum: bool = False,
                 log_weight_decay: bool = False) -> None:
        self.logging_interval = logging_interval
        self.log_weight_decay = log_weight_decay

    def _names(self, trainer):
        out = []
        for i, opt in enumerate(trainer.optimizers):
            base = 'lr-' + getattr(opt, 'inner_class_name', type(opt).__name__)
            if len(trainer.optimizers) > 1:
                base += f'-{i}'
            groups = opt.param_groups
            for g_idx, g in enumerate(groups):
                out.append((base if len(groups) == 1 else f'{base}/pg{g_idx + 1}', g))
        return out

    def on_train_batch_start(self, trainer, pl_module, batch, batch_idx) -> None:
        if self.logging_interval == 'epoch':
            return
        for name, group in self._names(trainer):
            trainer._log_metric(name, group['lr'])
            if self.log_weight_decay:
                trainer._log_metric(name + '-weight_decay', group.get('weight_decay', 0.0))

    def on_train_epoch_start(self, trainer, pl_module) -> None:
        if self.logging_interval == 'epoch':
            for name, group in self._names(trainer):
                trainer._log_metric(name, group['lr'])
