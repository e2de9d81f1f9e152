"""(reference: src/llm_training/lms/dpo/dpo_config.py:5-10)"""
from typing import Any

from pydantic import field_validator

from ..base_lm_config import BaseLightningModuleConfig
from ..utils import coerce_model_type


class DPOConfig(BaseLightningModuleConfig):
    model: Any
    ref_model: Any | None = None
    beta: float = 0.1
    label_smoothing: float = 0.0
    ignore_index: int = -100

    @field_validator('model', 'ref_model', mode='before')
    @classmethod
    def _model(cls, value: Any) -> Any:
        return coerce_model_type(value)
