"""(reference: src/llm_training/lms/clm/clm_config.py:5-9)"""
from typing import Any

from pydantic import field_validator

from ..base_lm_config import BaseLightningModuleConfig
from ..utils import coerce_model_type


class CLMConfig(BaseLightningModuleConfig):
    model: Any
    ignore_index: int = -100
    neftune_alpha: float | None = None
    log_perplexity: bool = True
    # extension: 'mean' = per-micro-batch mean like the reference (README.md:117-125)
    fused_loss: bool = True

    @field_validator('model', mode='before')
    @classmethod
    def _model(cls, value: Any) -> Any:
        return coerce_model_type(value)
