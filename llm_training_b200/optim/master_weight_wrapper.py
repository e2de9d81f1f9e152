"""fp32 master weights around any optimizer (reference: src/llm_training/optim/master_weight_wrapper.py:10-96).

Kept for API parity and for optimizers the fused path does not cover.  Differences: the class is an
ordinary wrapper (no dynamic subclass per wrapped type — ``type(opt).__name__`` still reports the inner
name through ``__class__`` shims used by LearningRateMonitor), and the master copies ARE saved in
``state_dict`` (``master_params``) so a resume does not lose the low bits; old
checkpoints without that key still load.
"""
from typing import Any

import torch
from torch.optim import Optimizer


class MasterWeightsOptimizer(Optimizer):
    def __init__(self, optimizer: Optimizer) -> None:   # noqa: super().__init__ deliberately not called
        self._optimizer = optimizer
        self._params = [p for g in optimizer.param_groups for p in g['params']]
        self._masters = [p if p.dtype == torch.float32 else p.detach().float().requires_grad_() for p in self._params]
        self._to_master = {id(p): m for p, m in zip(self._params, self._masters)}
        # the inner optimizer permanently works on the fp32 copies
        for group in optimizer.param_groups:
            group['params'] = [self._to_master[id(p)] for p in group['params']]

    # -- attribute plumbing ---------------------------------------------------------------------------
    @property
    def param_groups(self):
        return self._optimizer.param_groups

    @param_groups.setter
    def param_groups(self, value):
        self._optimizer.param_groups = value

    @property
    def state(self):
        return self._optimizer.state

    @property
    def defaults(self):
        return self._optimizer.defaults

    def __getattr__(self, name: str) -> Any:
        if name.startswith('_'):
            raise AttributeError(name)
        return getattr(self._optimizer, name)

    def __repr__(self) -> str:
        return f'MasterWeights({self._optimizer!r})'

    @property
    def inner_class_name(self) -> str:
        return type(self._optimizer).__name__

    # -- optimisation ---------------------------------------------------------------------------------
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for p, m in zip(self._params, self._masters):
            if p is m:
                continue
            m.grad = None if p.grad is None else p.grad.float()
        self._optimizer.step()
        with torch.no_grad():
            for p, m in zip(self._params, self._masters):
                if p is not m:
                    p.copy_(m)
        return loss

    def zero_grad(self, set_to_none: bool = True) -> None:
        self._optimizer.zero_grad(set_to_none)
        for p in self._params:
            if p.grad is None:
                continue
            if set_to_none:
                p.grad = None
            else:
                p.grad.zero_()

    def state_dict(self) -> dict:
        sd = self._optimizer.state_dict()
        sd['master_params'] = [m.detach().clone() for m in self._masters]
        return sd

    def load_state_dict(self, state_dict: dict) -> None:
        state_dict = dict(state_dict)
        masters = state_dict.pop('master_params', None)
        self._optimizer.load_state_dict(state_dict)
        with torch.no_grad():
            if masters is not None:
                for m, saved in zip(self._masters, masters):
                    m.copy_(saved)
            else:   # reference-format checkpoint: rebuild from the low-precision weights
                for p, m in zip(self._params, self._masters):
                    if p is not m:
                        m.copy_(p)
