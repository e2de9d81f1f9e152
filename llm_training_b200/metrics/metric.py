"""Minimal device-resident metric base (replaces torchmetrics, which is not a dependency).

Semantics kept from the reference (src/llm_training/metrics/metric.py:5-21):
  * states are tensors registered as buffers; *persistent* states land in ``state_dict`` and are
    restored on resume;
  * states are reduced over a process group with ``sum``.

Difference: the reference all-gathers metric state **every training step**.
Here ``update`` only touches the local device tensor; the cross-rank reduction happens in
``compute(sync=True)``, which the trainer calls only when it is about to log.
"""
from typing import Any

import torch
import torch.distributed as dist
from torch import nn


class Metric(nn.Module):
    higher_is_better: bool | None = None

    def __init__(self, process_group: Any | None = None, **_: Any) -> None:
        super().__init__()
        self.process_group = process_group
        self._defaults: dict[str, torch.Tensor] = {}
        self._reductions: dict[str, str] = {}

    def add_state(self, name: str, default: torch.Tensor, dist_reduce_fx: str = 'sum',
                  persistent: bool = False) -> None:
        assert dist_reduce_fx == 'sum', 'only sum-reduced states are needed here'
        self._defaults[name] = default.clone()
        self._reductions[name] = dist_reduce_fx
        self.register_buffer(name, default.clone(), persistent=persistent)

    def reset(self) -> None:
        for name, default in self._defaults.items():
            getattr(self, name).copy_(default)

    def update(self, *args: Any, **kwargs: Any) -> None:
        raise NotImplementedError

    def _compute(self, states: dict[str, torch.Tensor]) -> torch.Tensor:
        raise NotImplementedError

    def _synced_states(self, sync: bool) -> dict[str, torch.Tensor]:
        states = {name: getattr(self, name) for name in self._defaults}
        if sync and dist.is_available() and dist.is_initialized():
            group = self.process_group
            if dist.get_world_size(group) > 1:
                names = list(states)
                packed = torch.stack([states[n].double().reshape(()) for n in names])
                dist.all_reduce(packed, group=group)
                states = {n: packed[i].to(states[n].dtype) for i, n in enumerate(names)}
        return states

    def compute(self, sync: bool = True) -> torch.Tensor:
        return self._compute(self._synced_states(sync))

    def forward(self, *args: Any, **kwargs: Any) -> torch.Tensor:
        """Update, then return the value of *this batch only* (torchmetrics ``forward`` semantics)."""
        saved = {name: getattr(self, name).clone() for name in self._defaults}
        self.reset()
        self.update(*args, **kwargs)
        batch_value = self._compute({name: getattr(self, name).clone() for name in self._defaults})
        for name, old in saved.items():
            getattr(self, name).add_(old)
        return batch_value
