"""Progress bar whose counter starts at the restored batch index on resume
(reference: src/llm_training/lightning/callbacks/tqdm_progress.py:6-11)."""
import sys

from .base import Callback


class TQDMProgressBar(Callback):
    def __init__(self, refresh_rate: int = 1, process_position: int = 0, leave: bool = False) -> None:
        self.refresh_rate = max(0, refresh_rate)
        self.process_position = process_position
        self.leave = leave
        self.bar = None

    def on_train_epoch_start(self, trainer, pl_module) -> None:
        if not trainer.is_global_zero or self.refresh_rate == 0:
            return
        from tqdm.auto import tqdm
        total = trainer.num_training_batches
        self.bar = tqdm(total=None if total == float('inf') else total, initial=trainer.fit_loop.batch_idx,
                        desc=f'Epoch {trainer.current_epoch}', position=self.process_position, leave=self.leave,
                        dynamic_ncols=True, file=sys.stdout, disable=not sys.stdout.isatty() and False)

    def on_train_batch_end(self, trainer, pl_module, outputs, batch, batch_idx) -> None:
        if self.bar is None:
            return
        if (batch_idx + 1) % self.refresh_rate == 0:
            postfix = {k: f'{trainer._resolve(v):.4g}' for k, v in trainer.progress_bar_metrics.items()}
            self.bar.set_postfix(postfix, refresh=False)
        self.bar.update(1)

    def on_train_epoch_end(self, trainer, pl_module) -> None:
        if self.bar is not None:
            self.bar.close()
            self.bar = None
