"""``{model_class, model_config}`` -> lazy model factory (reference: src/llm_training/lms/model_provider.py:8-30)."""
from typing import Any

from ..models.base_model import BaseModel, BaseModelConfig
from ..utils.imports import resolve_class


class ModelProvider:
    def __init__(self, model_class: type[BaseModel] | str, model_config: dict[str, Any] | BaseModelConfig) -> None:
        self.model_class: type[BaseModel] = resolve_class(model_class)
        if isinstance(model_config, dict):
            self.model_config = self.model_class.config_class.model_validate(model_config)
        else:
            self.model_config = model_config

    def __call__(self) -> BaseModel:
        return self.model_class(self.model_config)

    def __repr__(self) -> str:
        return f'{type(self).__name__}(\n   model_class={self.model_class},\n   model_config={self.model_config!r}\n)'
