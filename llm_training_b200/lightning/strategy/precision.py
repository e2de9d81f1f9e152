"""Precision handling (reference: src/llm_training/lightning/strategy/fsdp2/fsdp2_precision.py:24-194).

``32-true | 16-true | bf16-true | 16-mixed | bf16-mixed`` -> parameter dtype, autocast dtype, reduce dtype.
``bf16-true`` is the hot path (bf16 parameters + fp32 master weights inside the sharded optimizer, every
op on the in-tree kernels); the others run through PyTorch ops.  The fp16 GradScaler is created for the
``16-*`` modes (the reference reads ``self.precision`` before assigning it, so its scaler is never
auto-created — SURVEY Appendix B7).
"""
from contextlib import nullcontext
from dataclasses import dataclass
from typing import Any, ContextManager, Literal

import torch

_PRECISION = Literal['32-true', '16-true', 'bf16-true', '16-mixed', 'bf16-mixed']
_ALIASES = {'32': '32-true', 32: '32-true', '16': '16-mixed', 16: '16-mixed', 'bf16': 'bf16-mixed'}


@dataclass
class MixedPrecisionPolicy:
    param_dtype: torch.dtype | None = None
    reduce_dtype: torch.dtype | None = None
    output_dtype: torch.dtype | None = None
    cast_forward_inputs: bool = True


@dataclass
class OffloadPolicy:
    pin_memory: bool | None = None


class CPUOffloadPolicy(OffloadPolicy):
    pass


class FSDP2Precision:
    def __init__(self, precision: str | int = '32-true', scaler: Any | None = None) -> None:
        precision = _ALIASES.get(precision, precision)
        supported = ('32-true', '16-true', 'bf16-true', '16-mixed', 'bf16-mixed')
        if precision not in supported:
            raise ValueError(f'`precision={precision!r}` is not supported; choose from {supported}')
        self.precision: str = precision
        self.scaler = scaler
        if precision == '16-mixed' and scaler is None:
            self.scaler = torch.amp.GradScaler('cuda' if torch.cuda.is_available() else 'cpu')
        table = {
            '32-true': (torch.float32, None),
            '16-true': (torch.float16, None),
            'bf16-true': (torch.bfloat16, None),
            '16-mixed': (torch.float32, torch.float16),
            'bf16-mixed': (torch.float32, torch.bfloat16),
        }
        self.param_dtype, self.autocast_dtype = table[precision]

    @property
    def mp_policy(self) -> MixedPrecisionPolicy:
        reduce = self.param_dtype if self.autocast_dtype is None else torch.float32
        return MixedPrecisionPolicy(param_dtype=self.autocast_dtype or self.param_dtype, reduce_dtype=reduce,
                                    output_dtype=self.autocast_dtype or self.param_dtype)

    def convert_module(self, module: torch.nn.Module) -> torch.nn.Module:
        """Cast floating-point parameters/buffers that live in a real device; meta tensors are converted too."""
        return module.to(dtype=self.param_dtype)

    def convert_input(self, data: Any) -> Any:
        return data

    def module_init_context(self) -> ContextManager:
        return _default_dtype(self.param_dtype)

    def forward_context(self) -> ContextManager:
        if self.autocast_dtype is None:
            return nullcontext()
        device_type = 'cuda' if torch.cuda.is_available() else 'cpu'
        return torch.autocast(device_type, dtype=self.autocast_dtype)

    def pre_backward(self, loss: torch.Tensor) -> torch.Tensor:
        return self.scaler.scale(loss) if self.scaler is not None else loss

    def state_dict(self) -> dict:
        return self.scaler.state_dict() if self.scaler is not None else {}

    def load_state_dict(self, state_dict: dict) -> None:
        if self.scaler is not None and state_dict:
            self.scaler.load_state_dict(state_dict)


class _default_dtype:
    def __init__(self, dtype: torch.dtype) -> None:
        self.dtype = dtype

    def __enter__(self):
        self.previous = torch.get_default_dtype()
        torch.set_default_dtype(self.dtype)

    def __exit__(self, *exc):
        torch.set_default_dtype(self.previous)
