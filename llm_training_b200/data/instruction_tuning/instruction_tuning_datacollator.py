"""(reference: src/llm_training/data/instruction_tuning/instruction_tuning_datacollator.py:12-72)
Keeps the per-document segment ids of a pack in ``attention_mask`` (0 = padding) and numbers positions
``0..n-1`` over the whole pack, like the reference."""
from typing import Any

import torch

from ..base_datacollator import BaseDataCollator
from ..chat_utils import pad_rows
from .instruction_tuning_datamodule_config import InstructionTuningDataModuleConfig, PackingMethod


class InstructionTuningDataCollator(BaseDataCollator):
    config: InstructionTuningDataModuleConfig

    def __init__(self, config: InstructionTuningDataModuleConfig) -> None:
        super().__init__(config)
        assert config.tokenizer.pad_token_id is not None, '`pad_token` is not specified. Please set it manually.'

    def __call__(self, batch: list[dict[str, Any]]) -> dict[str, torch.Tensor]:
        tok = self.config.tokenizer
        left = tok.padding_side == 'left'
        m = self.config.pad_to_multiple_of
        packed = self.config.packing_method == PackingMethod.GROUP_BY_LENGTH
        ids = [list(x['input_ids']) for x in batch]
        masks = [list(x['attention_mask']) if packed else [1] * len(x['input_ids']) for x in batch]
        positions = [list(range(len(r))) for r in ids]
        labels = [list(x['labels']) for x in batch]
        return {
            'input_ids': pad_rows(ids, tok.pad_token_id, left, m),
            'attention_mask': pad_rows(masks, 0, left, m),
            'position_ids': pad_rows(positions, 0, left, m),
            'labels': pad_rows(labels, -100, left, m),
        }
