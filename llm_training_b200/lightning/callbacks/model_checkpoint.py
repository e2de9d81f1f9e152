"""``ModelCheckpoint`` with Lightning's naming (``epoch=E-step=S.ckpt``) and the reference's directory rule:
checkpoints live in ``<log_dir>/checkpoints`` (src/llm_training/lightning/callbacks/model_checkpoint.py:13-18).
Supports the options the example configs use: ``save_on_train_epoch_end``, ``save_top_k`` (-1 / 0 / k),
``every_n_train_steps``, ``every_n_epochs``, ``monitor`` + ``mode``, ``save_last``, ``filename``."""
import os
import re
from typing import Any

from .base import Callback


class ModelCheckpoint(Callback):
    CHECKPOINT_EQUALS_CHAR = '='
    FILE_EXTENSION = '.ckpt'
    _instances = 0

    def __init__(self, dirpath: str | None = None, filename: str | None = None, monitor: str | None = None,
                 save_last: bool | None = None, save_top_k: int = 1, mode: str = 'min',
                 every_n_train_steps: int | None = None, every_n_epochs: int | None = None,
                 save_on_train_epoch_end: bool | None = None, save_weights_only: bool = False,
                 auto_insert_metric_name: bool = True, verbose: bool = False, **_: Any) -> None:
        self.dirpath = dirpath
        self.filename = filename
        self.monitor = monitor
        self.save_last = save_last
        self.save_top_k = save_top_k
        self.mode = mode
        self.every_n_train_steps = every_n_train_steps or 0
        self.every_n_epochs = every_n_epochs if every_n_epochs is not None else (0 if every_n_train_steps else 1)
        self.save_on_train_epoch_end = save_on_train_epoch_end
        self.auto_insert_metric_name = auto_insert_metric_name
        self.best_k: dict[str, float] = {}
        self.saved: list[str] = []
        self.best_model_path = ''
        self.last_model_path = ''
        self._last_step_saved = -1
        ModelCheckpoint._instances += 1
        self.state_key = f'ModelCheckpoint{{every_n_train_steps={self.every_n_train_steps},' \
                         f'every_n_epochs={self.every_n_epochs},monitor={monitor},top_k={save_top_k}}}'

    def setup(self, trainer, pl_module, stage: str) -> None:
        if self.dirpath is None:
            self.dirpath = os.path.join(trainer.log_dir, 'checkpoints')
        self.dirpath = trainer.strategy.broadcast(self.dirpath)

    def format_checkpoint_name(self, trainer, metrics: dict[str, Any]) -> str:
        template = self.filename or '{epoch}-{step}'
        values = dict(metrics, epoch=trainer.current_epoch, step=trainer.global_step)

        def repl(match: re.Match) -> str:
            key, fmt = match.group(1), match.group(2) or ''
            val = values.get(key, 0)
            text = format(val, fmt[1:]) if fmt else str(val)
            return f'{key}{self.CHECKPOINT_EQUALS_CHAR}{text}' if self.auto_insert_metric_name else text

        name = re.sub(r'\{([^{}:]+)(:[^{}]*)?\}', repl, template)
        return os.path.join(self.dirpath, name + self.FILE_EXTENSION)

    def _save(self, trainer) -> None:
        if self.save_top_k == 0 or trainer.global_step == self._last_step_saved:
            return
        metrics = {k: trainer._resolve(v) for k, v in trainer.callback_metrics.items()
                   if self.monitor is None or k == self.monitor}
        path = self.format_checkpoint_name(trainer, metrics)
        score = metrics.get(self.monitor) if self.monitor else None
        if self.monitor and self.save_top_k > 0 and len(self.best_k) >= self.save_top_k and score is not None:
            worst = (max if self.mode == 'min' else min)(self.best_k.values())
            if (score >= worst) if self.mode == 'min' else (score <= worst):
                return
        os.makedirs(self.dirpath, exist_ok=True) if trainer.is_global_zero else None
        trainer.save_checkpoint(path)
        self._last_step_saved = trainer.global_step
        self.last_model_path = path
        self.saved.append(path)
        if self.save_last:
            # Lightning's `last.ckpt`: always the most recent state, whatever top-k keeps
            last = os.path.join(self.dirpath, 'last' + self.FILE_EXTENSION)
            trainer.strategy.remove_checkpoint(last)
            trainer.strategy.barrier()
            trainer.save_checkpoint(last)
            self.last_model_path = last
        if self.monitor and score is not None:
            self.best_k[path] = score
            pick = min if self.mode == 'min' else max
            self.best_model_path = pick(self.best_k, key=self.best_k.get)
            while self.save_top_k > 0 and len(self.best_k) > self.save_top_k:
                drop = (max if self.mode == 'min' else min)(self.best_k, key=self.best_k.get)
                self.best_k.pop(drop)
                trainer.strategy.remove_checkpoint(drop)
        elif self.save_top_k > 0:
            self.best_model_path = path
            while len(self.saved) > self.save_top_k:
                trainer.strategy.remove_checkpoint(self.saved.pop(0))

    def on_train_batch_end(self, trainer, pl_module, outputs, batch, batch_idx) -> None:
        n = self.every_n_train_steps
        if n > 0 and trainer.global_step > 0 and trainer.global_step % n == 0:
            self._save(trainer)

    def on_train_epoch_end(self, trainer, pl_module) -> None:
        if self.save_on_train_epoch_end is False and self.monitor:
            return
        if self.every_n_epochs > 0 and (trainer.current_epoch + 1) % self.every_n_epochs == 0:
            self._save(trainer)

    def on_validation_end(self, trainer, pl_module) -> None:
        if self.save_on_train_epoch_end:
            return
        if self.monitor and self.every_n_train_steps == 0 and trainer.training is False:
            self._save(trainer)

    def state_dict(self) -> dict:
        return dict(best_k=self.best_k, saved=self.saved, best_model_path=self.best_model_path,
                    last_model_path=self.last_model_path, dirpath=self.dirpath)

    def load_state_dict(self, state_dict: dict) -> None:
        self.best_k = dict(state_dict.get('best_k', {}))
        self.saved = list(state_dict.get('saved', []))
        self.best_model_path = state_dict.get('best_model_path', '')
        self.last_model_path = state_dict.get('last_model_path', '')
