"""Config shared by models that interoperate with Hugging Face checkpoints
(reference: src/llm_training/models/hf_compat_model/hf_compat_config.py:9-26)."""
from typing import Literal

import torch
from pydantic import Field

from ..base_model.base_model_config import BaseModelConfig


class HFCompatModelConfig(BaseModelConfig):
    hf_path: str | None = None
    hf_tokenizer_path: str | None = None

    torch_dtype: str | torch.dtype = 'auto'
    trust_remote_code: bool = False
    low_cpu_mem_usage: bool = True
    revision: str = 'main'
    attn_implementation: Literal['eager', 'sdpa', 'flash_attention_2'] | None = None
    hf_extra_kwargs: dict = Field(default_factory=dict)

    load_hf_weights: bool = True

    @property
    def _attn_implementation(self) -> str:
        """``flash_attention_2`` names the in-tree packed attention kernel (CUDA) — on a machine
        without CUDA the default is ``sdpa`` (the reference calls ``get_device_capability`` here and
        therefore cannot even build a model on CPU, hf_compat_config.py:22-26)."""
        if self.attn_implementation is None:
            return 'flash_attention_2' if torch.cuda.is_available() else 'sdpa'
        return self.attn_implementation
