"""(reference: src/llm_training/lms/utils.py:6-12)"""
from typing import Any, Callable

from ..models.base_model import BaseModel
from .model_provider import ModelProvider

ModelType = ModelProvider | BaseModel | Callable[[], BaseModel]


def coerce_model_type(value: Any) -> Any:
    """YAML gives ``{model_class: ..., model_config: ...}``; turn it into a ModelProvider."""
    if isinstance(value, dict) and 'model_class' in value:
        return ModelProvider(value['model_class'], value.get('model_config', {}))
    return value


def get_model(model_or_provider: ModelType) -> BaseModel:
    if isinstance(model_or_provider, BaseModel):
        return model_or_provider
    return model_or_provider()
