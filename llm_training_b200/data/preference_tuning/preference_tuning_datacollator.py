"""Pads chosen and rejected sequences independently
(reference: src/llm_training/data/preference_tuning/preference_tuning_datacollator.py:12-69)."""
from typing import Any

import torch

from ..base_datacollator import BaseDataCollator
from ..chat_utils import pad_rows
from .preference_tuning_datamodule_config import PreferenceTuningDataModuleConfig


class PreferenceTuningDataCollator(BaseDataCollator):
    config: PreferenceTuningDataModuleConfig

    def __init__(self, config: PreferenceTuningDataModuleConfig) -> None:
        super().__init__(config)
        assert config.tokenizer.pad_token_id is not None, '`pad_token` is not specified. Please set it manually.'

    def __call__(self, batch: list[dict[str, Any]]) -> dict[str, torch.Tensor]:
        tok = self.config.tokenizer
        left = tok.padding_side == 'left'
        m = self.config.pad_to_multiple_of
        out: dict[str, torch.Tensor] = {}
        for side in ('chosen', 'rejected'):
            ids = [list(x[f'{side}_input_ids']) for x in batch]
            out[f'{side}_input_ids'] = pad_rows(ids, tok.pad_token_id, left, m)
            out[f'{side}_attention_mask'] = pad_rows([[1] * len(r) for r in ids], 0, left, m)
            out[f'{side}_labels'] = pad_rows([list(x[f'{side}_labels']) for x in batch], -100, left, m)
            out[f'{side}_position_ids'] = torch.arange(out[f'{side}_input_ids'].size(1)).unsqueeze(0)
        return out
