"""Meta-device construction (reference: src/llm_training/models/utils/utils.py:8-64).

Parameters are moved to ``device`` (usually ``meta``) the moment a module registers them, so no real storage
survives construction; buffers are left alone unless ``include_buffers`` — RoPE tables and other computed,
non-persistent buffers therefore hold real values.  Initialisers from ``torch.nn.init`` are turned into no-ops
inside the context: the freshly allocated CPU pages of a parameter are never touched before the parameter is
replaced, which keeps building a multi-billion-parameter skeleton cheap.
"""
from contextlib import contextmanager

import torch
from torch import nn

_INIT_FNS = ('uniform_', 'normal_', 'trunc_normal_', 'constant_', 'ones_', 'zeros_', 'eye_', 'dirac_',
             'xavier_uniform_', 'xavier_normal_', 'kaiming_uniform_', 'kaiming_normal_', 'orthogonal_', 'sparse_')


@contextmanager
def init_on_device(device: torch.device | str, include_buffers: bool = False):
    device = torch.device(device)
    original_register_parameter = nn.Module.register_parameter
    original_register_buffer = nn.Module.register_buffer

    def register_parameter(module, name, param):
        original_register_parameter(module, name, param)
        if param is not None and param.device != device:
            cls = type(module._parameters[name])
            kwargs = {k: v for k, v in module._parameters[name].__dict__.items()}
            moved = cls(module._parameters[name].to(device), requires_grad=param.requires_grad)
            moved.__dict__.update(kwargs)
            module._parameters[name] = moved

    def register_buffer(module, name, tensor, persistent=True):
        original_register_buffer(module, name, tensor, persistent=persistent)
        if tensor is not None and tensor.device != device:
            module._buffers[name] = module._buffers[name].to(device)

    if device.type == 'meta':
        # transformers snapshots the torch.nn.init functions when its modelling modules are first imported; make
        # sure that happens BEFORE they are replaced below, or every later Hugging Face model of the process would
        # be "initialised" by the no-ops
        for mod in ('transformers.modeling_utils', 'transformers.initialization'):
            try:
                __import__(mod)
            except Exception:  # noqa: BLE001 -- transformers absent or the module does not exist in this version
                pass
    saved_inits = {fn: getattr(nn.init, fn) for fn in _INIT_FNS if hasattr(nn.init, fn)}

    def _noop(tensor, *args, **kwargs):
        return tensor

    nn.Module.register_parameter = register_parameter
    if include_buffers:
        nn.Module.register_buffer = register_buffer
    if device.type == 'meta':
        for fn in saved_inits:
            setattr(nn.init, fn, _noop)
    try:
        yield
    finally:
        nn.Module.register_parameter = original_register_parameter
        nn.Module.register_buffer = original_register_buffer
        for fn, impl in saved_inits.items():
            setattr(nn.init, fn, impl)


def init_empty_weights(include_buffers: bool = False):
    return init_on_device('meta', include_buffers=include_buffers)
