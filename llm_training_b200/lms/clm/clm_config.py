"""(reference: src/llm_training/lms/clm/clm_config.py:5-9)"""
from typing import Any, Literal

from pydantic import field_validator

from ..base_lm_config import BaseLightningModuleConfig
from ..utils import coerce_model_type


class CLMConfig(BaseLightningModuleConfig):
    model: Any
    ignore_index: int = -100
    neftune_alpha: float | None = None
    log_perplexity: bool = True
    fused_loss: bool = True
    # extension.  'micro_batch' (default) = every rank's micro-batch mean, averaged over ranks, exactly like the
    # reference (its README.md:117-125 discusses the bias).  'global_tokens' weights every token of the data-parallel
    # step equally: each rank's loss is rescaled by its share of the valid tokens (one scalar all-reduce per step).
    loss_normalization: Literal['micro_batch', 'global_tokens'] = 'micro_batch'

    @field_validator('model', mode='before')
    @classmethod
    def _model(cls, value: Any) -> Any:
        return coerce_model_type(value)
