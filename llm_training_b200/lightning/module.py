"""Minimal stand-ins for ``lightning.LightningModule`` / ``LightningDataModule``: just the contract the
reference's LMs and datamodules use (hooks, ``self.log``, ``self.trainer``), with no dependency on
Lightning."""
from typing import Any

import torch
from torch import nn


class LightningModule(nn.Module):
    def __init__(self) -> None:
        super().__init__()
        self._trainer = None

    # -- wiring ----------------------------------------------------------------------------------------
    @property
    def trainer(self):
        if self._trainer is None:
            raise RuntimeError(f'{type(self).__name__} is not attached to a Trainer')
        return self._trainer

    @trainer.setter
    def trainer(self, trainer) -> None:
        self._trainer = trainer

    @property
    def global_rank(self) -> int:
        return 0 if self._trainer is None else self._trainer.global_rank

    @property
    def local_rank(self) -> int:
        return 0 if self._trainer is None else self._trainer.local_rank

    @property
    def global_step(self) -> int:
        return 0 if self._trainer is None else self._trainer.global_step

    @property
    def current_epoch(self) -> int:
        return 0 if self._trainer is None else self._trainer.current_epoch

    @property
    def logger(self):
        return None if self._trainer is None else self._trainer.logger

    @property
    def device(self) -> torch.device:
        for p in self.parameters():
            if not p.is_meta:
                return p.device
        return torch.device('cpu') if self._trainer is None else self._trainer.strategy.root_device

    # -- logging ---------------------------------------------------------------------------------------
    def log(self, name: str, value: Any, prog_bar: bool = False, logger: bool = True, batch_size: int | None = None,
            sync_dist: bool = False, **_: Any) -> None:
        if self._trainer is not None:
            self._trainer._log_metric(name, value, prog_bar=prog_bar, logger=logger, sync_dist=sync_dist,
                                      batch_size=batch_size)

    def log_dict(self, metrics: dict[str, Any], **kwargs: Any) -> None:
        for k, v in metrics.items():
            self.log(k, v, **kwargs)

    # -- hooks (no-ops by default) -----------------------------------------------------------------------
    def configure_model(self) -> None: ...
    def configure_optimizers(self): ...
    def training_step(self, batch: Any, batch_idx: int): ...
    def validation_step(self, batch: Any, batch_idx: int, dataloader_idx: int = 0): ...
    def on_fit_start(self) -> None: ...
    def on_fit_end(self) -> None: ...
    def on_train_start(self) -> None: ...
    def on_train_end(self) -> None: ...
    def on_train_epoch_start(self) -> None: ...
    def on_train_epoch_end(self) -> None: ...
    def on_validation_epoch_start(self) -> None: ...
    def on_validation_epoch_end(self) -> None: ...
    def on_train_batch_start(self, batch: Any, batch_idx: int) -> None: ...
    def on_train_batch_end(self, outputs: Any, batch: Any, batch_idx: int) -> None: ...
    def on_save_checkpoint(self, checkpoint: dict) -> None: ...
    def on_load_checkpoint(self, checkpoint: dict) -> None: ...

    def backward(self, loss: torch.Tensor, *args: Any, **kwargs: Any) -> None:
        loss.backward(*args, **kwargs)

    def optimizer_step(self, epoch: int, batch_idx: int, optimizer, optimizer_closure=None) -> None:
        optimizer.step(closure=optimizer_closure)

    def optimizer_zero_grad(self, epoch: int, batch_idx: int, optimizer) -> None:
        optimizer.zero_grad()

    def transfer_batch_to_device(self, batch: Any, device: torch.device) -> Any:
        return move_to_device(batch, device)


class LightningDataModule:
    prepare_data_per_node: bool = True

    def __init__(self) -> None:
        self.trainer = None

    def prepare_data(self) -> None: ...
    def setup(self, stage: str | None = None) -> None: ...
    def teardown(self, stage: str | None = None) -> None: ...
    def train_dataloader(self): return None
    def val_dataloader(self): return None
    def test_dataloader(self): return None
    def predict_dataloader(self): return None
    def state_dict(self) -> dict: return {}
    def load_state_dict(self, state_dict: dict) -> None: ...


def move_to_device(obj: Any, device: torch.device) -> Any:
    if isinstance(obj, torch.Tensor):
        return obj.to(device, non_blocking=True)
    if isinstance(obj, dict):
        return {k: move_to_device(v, device) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(move_to_device(v, device) for v in obj)
    return obj
