"""(reference: src/llm_training/lms/orpo/orpo_config.py:5-9)"""
from typing import Any

from pydantic import field_validator

from ..base_lm_config import BaseLightningModuleConfig
from ..utils import coerce_model_type


class ORPOConfig(BaseLightningModuleConfig):
    model: Any
    beta: float = 0.1
    ignore_index: int = -100
    empty_cache_threshold: int | None = None

    @field_validator('model', mode='before')
    @classmethod
    def _model(cls, value: Any) -> Any:
        return coerce_model_type(value)
