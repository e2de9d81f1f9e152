"""Precision handling (reference: src/llm_training/lightning/strategy/fsdp2/fsdp2_precision.py:24-194).

``32-true | 16-true | bf16-true | 16-mixed | bf16-mixed`` -> parameter dtype, autocast dtype, reduce dtype.

On a GPU the sm_100a kernels are bf16 kernels (fp32 accumulation, fp32 softmax / norm / loss statistics inside):

* ``bf16-true``  bf16 parameters + fp32 master weights inside the sharded optimizer — the hot path;
* ``bf16-mixed`` runs on the SAME kernels: what the reference expresses as "fp32 parameters + bf16 autocast + bf16
  gradient reduction" (fsdp2_precision.py:63-69) is, in this engine, "fp32 master slices in the optimizer + a bf16
  compute copy that the modules see"; the ops autocast keeps in fp32 (norm statistics, softmax, loss) are fp32 inside
  the kernels.  The two bf16 modes therefore differ only in name here (and in the reference only in which copy is
  checkpointed: the bf16 one here);
* ``32-true``, ``16-true``, ``16-mixed`` have no native kernels.  They raise on a GPU instead of silently running
  cuBLAS / eager PyTorch; ``LLMT_ALLOW_TORCH_PRECISION=1`` opts into the PyTorch path (logged once).

On CPU every mode is plain PyTorch (the oracle / gloo test path).  The fp16 GradScaler is created for the ``16-*``
modes (the reference reads ``self.precision`` before assigning it, so its scaler is never auto-created — SURVEY
Appendix B7).
"""
import logging
import os
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
        self._table = table
        self.bind_device(self.device)

    def bind_device(self, device_type: str) -> None:
        """(Re)derive the dtypes for the device training actually runs on (the strategy calls this once the root
        device is known: `accelerator='cpu'` on a GPU box must get the CPU semantics)."""
        self.device = device_type
        self.param_dtype, self.autocast_dtype = self._table[self.precision]
        # GPU: which modes own a native kernel path (see module docstring)
        self.native = self.device == 'cuda' and self.precision in ('bf16-true', 'bf16-mixed')
        if self.device == 'cuda' and self.precision == 'bf16-mixed':
            self.param_dtype, self.autocast_dtype = torch.bfloat16, None

    def check_supported(self) -> None:
        """Called by the strategy when training is set up on a device: a GPU run in a mode without native kernels
        stops here (or continues on PyTorch ops when ``LLMT_ALLOW_TORCH_PRECISION=1``)."""
        if self.device != 'cuda' or self.native:
            return
        if os.environ.get('LLMT_ALLOW_TORCH_PRECISION', '0') != '1':
            raise NotImplementedError(
                f'precision={self.precision!r} has no sm_100a kernel path (the kernels are bf16 with fp32 '
                'accumulation and fp32 master weights): use bf16-true or bf16-mixed, or set '
                'LLMT_ALLOW_TORCH_PRECISION=1 to run this mode on PyTorch / cuBLAS ops instead')
        logging.getLogger(__name__).warning(
            f'precision={self.precision!r}: running on PyTorch ops (cuBLAS / eager), NOT on the in-tree kernels')

    @property
    def mp_policy(self) -> MixedPrecisionPolicy:
        """The reference's table (fsdp2_precision.py:55-90): compute dtype, gradient-reduction dtype, output dtype.
        Gradients are reduced in the compute dtype for every mode (bf16 for both bf16 modes, fp16 for both fp16
        modes, fp32 for 32-true)."""
        compute = self.autocast_dtype or self.param_dtype
        return MixedPrecisionPolicy(param_dtype=compute, reduce_dtype=compute, output_dtype=compute,
                                    cast_forward_inputs=False)

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
        return torch.autocast(self.device, dtype=self.autocast_dtype)

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
