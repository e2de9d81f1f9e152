"""(reference: src/llm_training/data/dummy/dummy_datamodule_config.py:8-26)"""
import random

from pydantic import Field, model_validator

from ..base_datamodule_config import BaseDataModuleConfig


class DummyDataModuleConfig(BaseDataModuleConfig):
    vocab_size: int
    max_length: int
    num_samples: int | None = None
    num_tokens: int | None = None
    base_seed: int | None = Field(None, validate_default=True)

    @model_validator(mode='after')
    def _check(self) -> 'DummyDataModuleConfig':
        if (self.num_samples is None) == (self.num_tokens is None):
            raise ValueError('give exactly one of `num_samples` and `num_tokens`')
        if self.base_seed is None:
            self.base_seed = random.randrange(0, 999_999)
        return self
