"""Phi-3 / 3.5 / 4 hyper-parameters (reference: src/llm_training/models/phi3/phi3_config.py:9-79)."""
from typing import Any, Literal

import torch
from pydantic import ValidationInfo, field_validator

from ..hf_compat_model import HFCompatModelConfig


class Phi3Config(HFCompatModelConfig):
    vocab_size: int = 32064
    hidden_size: int = 3072
    intermediate_size: int = 8192
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int | None = None
    resid_pdrop: float = 0.0
    embd_pdrop: float = 0.0
    attention_dropout: float = 0.0
    max_position_embeddings: int = 4096
    original_max_position_embeddings: int = 4096
    initializer_range: float = 0.02
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    rope_scaling: dict | None = None
    bos_token_id: int | None = 1
    eos_token_id: int | list[int] | None = 32000
    pad_token_id: int | None = 32000
    sliding_window: int | None = None
    tie_word_embeddings: bool = False

    enable_gradient_checkpointing: bool = False
    recompute_granularity: Literal['full', 'selective'] = 'full'
    attention_compute_dtype: torch.dtype | str | None = None

    @field_validator('rope_scaling')
    @classmethod
    def validate_rope_scaling(cls, rope_scaling: dict[str, Any] | None, info: ValidationInfo):
        if rope_scaling is None:
            return None
        kind = rope_scaling.get('type', rope_scaling.get('rope_type'))
        if kind not in ('longrope', 'su', 'yarn'):
            raise ValueError(f"`rope_scaling` type must be 'longrope', got {kind!r}")
        half = info.data['hidden_size'] // info.data['num_attention_heads'] // 2
        for key in ('short_factor', 'long_factor'):
            values = rope_scaling.get(key)
            if not (isinstance(values, list) and all(isinstance(x, (int, float)) for x in values)):
                raise ValueError(f'`rope_scaling.{key}` must be a list of numbers, got {values!r}')
            if len(values) != half:
                raise ValueError(f'`rope_scaling.{key}` must have length {half}, got {len(values)}')
        normalised = dict(rope_scaling)
        normalised['type'] = 'longrope'
        normalised.pop('rope_type', None)
        return normalised
