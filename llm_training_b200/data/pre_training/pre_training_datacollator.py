"""Pads a batch of (packed) token rows; labels mask BOS and padding
(reference: src/llm_training/data/pre_training/pre_training_datacollator.py:9-46).

Parity note: like the reference, the default batch carries an all-ones-minus-padding
mask and ``arange`` positions, i.e. documents inside a pack attend to each other.  With
``block_cross_document_attention: true`` the packer's segment ids are kept instead."""
from typing import Any

import torch

from ..base_datacollator import BaseDataCollator
from .pre_training_datamodule_config import PreTrainingDataModuleConfig


class PreTrainingDataCollator(BaseDataCollator):
    config: PreTrainingDataModuleConfig

    def __init__(self, config: PreTrainingDataModuleConfig) -> None:
        super().__init__(config)
        assert config.tokenizer.pad_token_id is not None, '`pad_token` is not specified. Please set it manually.'

    @property
    def tokenizer(self):
        return self.config.tokenizer

    def _target_length(self, rows: list[list[int]]) -> int:
        n = max(len(r) for r in rows)
        m = self.config.pad_to_multiple_of
        return n if m is None else ((n // m) + 1) * m

    def _pad(self, rows: list[list[int]], width: int, value: int) -> torch.Tensor:
        out = torch.full((len(rows), width), value, dtype=torch.long)
        left = self.tokenizer.padding_side == 'left'
        for i, r in enumerate(rows):
            if not r:
                continue
            t = torch.as_tensor(r, dtype=torch.long)
            if left:
                out[i, width - len(r):] = t
            else:
                out[i, :len(r)] = t
        return out

    def __call__(self, batch: list[dict[str, Any]]) -> dict[str, torch.Tensor]:
        rows = [list(x['input_ids']) for x in batch]
        width = self._target_length(rows)
        input_ids = self._pad(rows, width, -1)
        padding = input_ids == -1
        input_ids[padding] = self.tokenizer.pad_token_id
        bos = input_ids == self.tokenizer.bos_token_id if self.tokenizer.bos_token_id is not None \
            else torch.zeros_like(padding)
        if self.config.block_cross_document_attention and all('attention_mask' in x for x in batch):
            attention_mask = self._pad([list(x['attention_mask']) for x in batch], width, 0)
        else:
            attention_mask = (~padding).long()
        return {
            'input_ids': input_ids,
            'attention_mask': attention_mask,
            'position_ids': torch.arange(width).unsqueeze(0),
            'labels': input_ids.masked_fill(bos | padding, -100),
        }
