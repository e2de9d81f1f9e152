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
    def __init__(self, precision: str | int = '32-true', device: str | None = None, scaler: Any | None = None
                 ) -> None:
        """Same arguments as the reference (fsdp2_precision.py:25-30); ``device`` only selects the GradScaler /
        autocast device type and defaults to CUDA when available."""
        precision = _ALIASES.get(precision, precision)
        self.device = device or ('cuda' if torch.cuda.is_available() else 'cpu')
        supported = ('32-true', '16-true', 'bf16-true', '16-mixed', 'bf16-mixed')
        if precision not in supported:
            raise ValueError(f'`precision={precision!r}` is not supported; choose from {supported}')
        self.precision: str = precision
        self.scaler = scaler
        if precision == '16-mixed' and scaler is None:
            self.scaler = torch.amp.GradScaler(self.device)
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

    def tensor_init_context(self) -> ContextManager:
        """Default dtype while modules are built (reference: fsdp2_precision.py:104-106)."""
        return self.module_init_context()

    # ---- optimizer step with dynamic loss scaling (fp16-mixed).  The reference delegates to torch's GradScaler
    #      (fsdp2_precision.py:130-163); the sharded optimizer unscales inside its fused update instead, so the only
    #      host work is reading the norm's finiteness.
    growth_factor, backoff_factor, growth_interval = 2.0, 0.5, 2000
    min_scale: float = 1.0
    raise_error_at_min_scale: bool = False

    def optimizer_step(self, optimizer: Any, model: Any = None, closure: Any = None, **kwargs: Any) -> Any:
        if closure is not None:
            closure()
        if self.scaler is None:
            return optimizer.step(**kwargs)
        scale = float(self.scaler.get_scale())
        optimizer.grad_scale, optimizer.skip_nonfinite = 1.0 / scale, True
        out = optimizer.step(**kwargs)
        if getattr(optimizer, 'last_step_skipped', False):
            self.skipped_steps = getattr(self, 'skipped_steps', 0) + 1
            self._good_steps = 0
            new_scale = scale * self.backoff_factor
            if new_scale < self.min_scale:
                if self.raise_error_at_min_scale:
                    raise RuntimeError(f'fp16 loss scale fell below the minimum ({self.min_scale}): the gradients '
                                       'overflow even at the smallest scale')
                new_scale = self.min_scale
        else:
            self._good_steps = getattr(self, '_good_steps', 0) + 1
            new_scale = scale * self.growth_factor if self._good_steps % self.growth_interval == 0 else scale
        self.scaler.update(new_scale)
        return out

    def clip_grad_by_norm(self, optimizer: Any, clip_val: float) -> None:
        """The clip runs inside the fused optimizer kernel from the device-side global norm; this only sets it."""
        optimizer.max_grad_norm = float(clip_val) if clip_val else None

    def clip_gradients(self, optimizer: Any, clip_val: int | float = 0.0, gradient_clip_algorithm: Any = 'norm'
                       ) -> None:
        if str(getattr(gradient_clip_algorithm, 'value', gradient_clip_algorithm)).lower() not in ('norm', 'none'):
            raise ValueError('only gradient clipping by global norm is supported (the reference clips by norm)')
        self.clip_grad_by_norm(optimizer, clip_val)

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
