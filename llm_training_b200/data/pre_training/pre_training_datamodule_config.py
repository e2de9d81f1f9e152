"""(reference: src/llm_training/data/pre_training/pre_training_datamodule_config.py:10-44)"""
from typing import Any

from pydantic import Field, model_validator

from ...utils.str_enum import StrEnum
from ..hf_based import HFBasedDataModuleConfig


class PackingMethod(StrEnum):
    NO_PACKING = 'no_packing'
    NAIVE_PACKING = 'naive_packing'
    BEST_FIT_BIN_PACKING = 'best_fit_bin_packing'


class PreTrainingDataModuleConfig(HFBasedDataModuleConfig):
    tokenizer: Any
    max_length: int | None = None
    stride: int | None = None
    packing_method: PackingMethod = PackingMethod.NAIVE_PACKING
    sample_rate: dict[str, float] = Field(default_factory=dict)
    pre_processing_batch_size: int = 1000
    pad_to_multiple_of: int | None = None
    # extension (default off = reference behaviour): keep the packer's per-document
    # segment ids in the batch so documents of one pack do not attend to each other
    block_cross_document_attention: bool = False

    @model_validator(mode='after')
    def _check(self) -> 'PreTrainingDataModuleConfig':
        # packing needs a capacity (the reference's assertion is inverted, Appendix B14)
        if self.packing_method != PackingMethod.NO_PACKING and self.max_length is None:
            raise ValueError('You must set `max_length` to pack data')
        if self.stride is None:
            self.stride = self.max_length
        else:
            if self.max_length is None:
                raise ValueError('You must also set `max_length` to use `stride`')
            if self.stride > self.max_length:
                raise ValueError('`stride` must be <= `max_length`')
        return self
