"""Pydantic base for model configs (reference: src/llm_training/models/base_model/base_model_config.py:8-21)."""
import types
import typing
from typing import Any

import torch
from pydantic import BaseModel, ConfigDict, field_validator


def _mentions_dtype(annotation: Any) -> bool:
    if annotation is torch.dtype:
        return True
    if isinstance(annotation, types.UnionType) or typing.get_origin(annotation) is typing.Union:
        return any(_mentions_dtype(a) for a in typing.get_args(annotation))
    return False


class BaseModelConfig(BaseModel):
    pre_trained_weights: str | None = None

    model_config = ConfigDict(arbitrary_types_allowed=True, protected_namespaces=())

    @field_validator('*', mode='before')
    @classmethod
    def _coerce_torch_dtype(cls, value: Any, info) -> Any:
        """'bfloat16' / 'torch.bfloat16' -> torch.bfloat16 for dtype-typed fields ('auto' passes through)."""
        field = cls.model_fields.get(info.field_name)
        if field is not None and _mentions_dtype(field.annotation) and isinstance(value, str) and value != 'auto':
            return getattr(torch, value.removeprefix('torch.'))
        return value
