"""(reference: src/llm_training/data/preference_tuning/preference_tuning_datamodule_config.py:15-31)"""
from typing import Any

from pydantic import field_validator

from ...utils.str_enum import StrEnum
from ..chat_templates import get_chat_template
from ..hf_based import HFBasedDataModuleConfig


class OverlongHandlingMethod(StrEnum):
    DROP = 'drop'


class PreferenceTuningDataModuleConfig(HFBasedDataModuleConfig):
    tokenizer: Any
    chat_template: str | None = None
    max_length: int | None = None
    overlong_handling_method: OverlongHandlingMethod = OverlongHandlingMethod.DROP
    pad_to_multiple_of: int | None = None

    @field_validator('chat_template')
    @classmethod
    def _template(cls, value: str | None) -> str | None:
        return None if value is None else get_chat_template(value)
