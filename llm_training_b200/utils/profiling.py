"""Tracing helpers (the reference has none): NVTX ranges around sharding units and collectives, (the CUDA-event step timer
is ``lightning.callbacks.StepTimeMonitor``).

``LLMT_NVTX=1`` (or ``enable_nvtx(model)``) wraps every sharding unit's forward in an NVTX range so that a timeline
(nsys, or ``ncu --nvtx --nvtx-include``) can be cut per layer; the ranges cost one push/pop per unit and are off by
default.
"""
import contextlib
import os

import torch
from torch import nn


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


def nvtx_requested() -> bool:
    return os.environ.get('LLMT_NVTX', '0') == '1'


def enable_nvtx(model: nn.Module) -> int:
    """One NVTX range per sharding unit forward (``embed_tokens``, ``layers.N``, ``norm``, ``lm_head``)."""
    units = model.sharding_units() if hasattr(model, 'sharding_units') else [('model', model)]
    count = 0
    for name, module in units:
        if getattr(module, '_llmt_nvtx', False):
            continue
        module._llmt_nvtx = True
        module.register_forward_pre_hook(lambda m, a, n=name: torch.cuda.nvtx.range_push(f'fwd:{n}'))
        module.register_forward_hook(lambda m, a, o: torch.cuda.nvtx.range_pop())
        count += 1
    return count
