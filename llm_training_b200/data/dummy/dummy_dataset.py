"""Deterministic random tokens, sample ``i`` seeded by ``base_seed + i``
(reference: src/llm_training/data/dummy/dummy_dataset.py:9-33)."""
import math

import torch
from torch.utils.data import Dataset

from .dummy_datamodule_config import DummyDataModuleConfig


class DummyDataset(Dataset):
    def __init__(self, config: DummyDataModuleConfig) -> None:
        self.config = config
        self.base_seed = config.base_seed
        if config.num_samples is not None:
            self.num_samples = config.num_samples
        else:
            self.num_samples = math.ceil(config.num_tokens / config.max_length)

    def __len__(self) -> int:
        return self.num_samples

    def __getitem__(self, index: int) -> dict[str, torch.Tensor]:
        g = torch.Generator()
        g.manual_seed(self.base_seed + index)
        input_ids = torch.randint(0, self.config.vocab_size, (self.config.max_length,), generator=g)
        return dict(input_ids=input_ids, attention_mask=torch.ones_like(input_ids),
                    position_ids=torch.arange(input_ids.size(0)), labels=input_ids)
