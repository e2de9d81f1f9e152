"""(reference: src/llm_training/data/instruction_tuning/instruction_tuning_datamodule_config.py:15-57)"""
from typing import Any

from pydantic import field_validator, model_validator

from ...utils.str_enum import StrEnum
from ..chat_templates import get_chat_template
from ..hf_based import HFBasedDataModuleConfig


class OverlongHandlingMethod(StrEnum):
    DROP = 'drop'
    TRUNCATE = 'truncate'


class PackingMethod(StrEnum):
    NO_PACKING = 'no_packing'
    GROUP_BY_LENGTH = 'group_by_length'


class InstructionTuningDataModuleConfig(HFBasedDataModuleConfig):
    tokenizer: Any
    chat_template: str | None = None
    max_length: int | None = None
    overlong_handling_method: OverlongHandlingMethod = OverlongHandlingMethod.DROP
    packing_method: PackingMethod = PackingMethod.NO_PACKING
    pad_to_multiple_of: int | None = None
    add_default_system_prompt_rate: float | None = None
    default_system_prompt: str | None = None

    @field_validator('chat_template')
    @classmethod
    def _template(cls, value: str | None) -> str | None:
        return None if value is None else get_chat_template(value)

    @model_validator(mode='after')
    def _check(self) -> 'InstructionTuningDataModuleConfig':
        if self.add_default_system_prompt_rate is not None and self.default_system_prompt is None:
            raise ValueError('`default_system_prompt` must be set to use `add_default_system_prompt_rate`')
        return self
