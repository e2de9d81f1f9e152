"""Meta-device construction (reference: src/llm_training/models/utils/utils.py:8-64, which patches
``nn.Module.register_parameter``).  ``torch.device`` as a context manager is the supported way to
do the same thing today; buffers can be kept on a real device so RoPE tables stay usable."""
from contextlib import contextmanager

import torch
from torch import nn


@contextmanager
def init_on_device(device: torch.device | str, include_buffers: bool = False):
    device = torch.device(device)
    if include_buffers:
        with device:
            yield
        return
    original = nn.Module.register_buffer

    def register_buffer(self, name, tensor, persistent=True):
        if tensor is not None and tensor.device.type == device.type:
            tensor = torch.empty_like(tensor, device='cpu') if tensor.is_meta else tensor.to('cpu')
        return original(self, name, tensor, persistent=persistent)

    nn.Module.register_buffer = register_buffer
    try:
        with device:
            yield
    finally:
        nn.Module.register_buffer = original


def init_empty_weights(include_buffers: bool = False):
    return init_on_device('meta', include_buffers=include_buffers)
