"""(reference: src/llm_training/models/hf_causal_lm/hf_causal_lm_config.py:8-17)"""
from typing import Any

from ..hf_compat_model import HFCompatModelConfig


class HFCausalLMConfig(HFCompatModelConfig):
    enable_gradient_checkpointing: bool = False
    enable_liger_kernel: bool = False   # route the HF model's RMSNorm / SwiGLU MLP modules through the in-tree ops

    hf_config: Any | None = None

    def __getattr__(self, name: str) -> Any:
        hf_config = self.__dict__.get('hf_config')
        if hf_config is not None and hasattr(hf_config, name):
            return getattr(hf_config, name)
        return super().__getattr__(name)
