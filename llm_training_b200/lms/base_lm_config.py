"""(reference: src/llm_training/lms/base_lm_config.py:13-43)"""
from typing import Any

from pydantic import BaseModel as PydanticModel
from pydantic import ConfigDict, Field, field_validator

from ..lr_schedulers import ConstantWarmupLR
from ..models.base_model.base_model_config import _mentions_dtype
from ..utils.imports import resolve_class


class BaseOptimizerConfig(PydanticModel):
    optimizer_class: Any
    optimizer_kwargs: dict[str, Any] = Field(default_factory=dict)
    lr_scheduler_class: Any = ConstantWarmupLR
    lr_scheduler_kwargs: dict[str, Any] = Field(default_factory=dict)

    model_config = ConfigDict(arbitrary_types_allowed=True)

    @field_validator('optimizer_class', 'lr_scheduler_class', mode='before')
    @classmethod
    def _resolve(cls, value: Any) -> Any:
        """Dotted strings (``torch.optim.AdamW``, ``deepspeed.ops.adam.FusedAdam`` -> native fused AdamW)."""
        resolved = resolve_class(value)
        if not isinstance(resolved, type):
            raise ValueError(f'{value!r} does not name a class')
        return resolved


class BaseLightningModuleConfig(PydanticModel):
    init_weights: bool = False
    load_weights: bool = True
    pre_trained_weights: str | None = None
    optim: BaseOptimizerConfig | None = None
    frozen_modules: list[str] | None = None
    log_grad_norm: bool = True

    model_config = ConfigDict(arbitrary_types_allowed=True, protected_namespaces=())

    @field_validator('*', mode='before')
    @classmethod
    def _coerce_torch_dtype(cls, value: Any, info) -> Any:
        import torch
        field = cls.model_fields.get(info.field_name)
        if field is not None and _mentions_dtype(field.annotation) and isinstance(value, str) and value != 'auto':
            return getattr(torch, value.removeprefix('torch.'))
        return value
