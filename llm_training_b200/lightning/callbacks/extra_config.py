"""Process-wide settings applied before anything else runs
(reference: src/llm_training/lightning/callbacks/extra_config.py:13-45).  The reference also gives every
rank its own TRITON_CACHE_DIR; there is no Triton here, the in-tree kernels are prebuilt."""
import logging
from typing import Literal

import torch

from .base import Callback


class ExtraConfig(Callback):
    def __init__(self, float32_matmul_precision: Literal['medium', 'high', 'highest'] | None = None,
                 logging_level: int | str = logging.INFO) -> None:
        self.float32_matmul_precision = float32_matmul_precision
        self.logging_level = logging_level
        if float32_matmul_precision is not None:
            torch.set_float32_matmul_precision(float32_matmul_precision)
        level = getattr(logging, logging_level.upper()) if isinstance(logging_level, str) else logging_level
        for name in ('llm_training', 'llm_training_b200'):
            logging.getLogger(name).setLevel(level)
