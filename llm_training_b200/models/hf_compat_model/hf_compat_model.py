"""Bridge between in-tree models and Hugging Face: config merge, weight import (rank 0), export
(reference: src/llm_training/models/hf_compat_model/hf_compat_model.py:16-145)."""
from contextlib import contextmanager
from typing import Any, Generator

import torch

from ..base_model.base_model import BaseModel
from ..utils.utils import init_empty_weights
from .hf_compat_config import HFCompatModelConfig


class HFCompatModel(BaseModel):
    config: HFCompatModelConfig
    config_class = HFCompatModelConfig

    hf_config_class: Any = None   # default: transformers.AutoConfig
    hf_model_class: Any = None    # default: transformers.AutoModel

    def __init__(self, config: HFCompatModelConfig) -> None:
        super().__init__(config)
        self.hf_config = None
        if self.config.hf_path is not None:
            self.hf_config = self.load_hf_config()
            self.merge_hf_config(self.hf_config)

    # -- HF classes are resolved lazily so that importing the package never needs transformers -------
    @classmethod
    def _hf_config_class(cls):
        if cls.hf_config_class is not None:
            return cls.hf_config_class
        from transformers import AutoConfig
        return AutoConfig

    @classmethod
    def _hf_model_class(cls):
        if cls.hf_model_class is not None:
            return cls.hf_model_class
        from transformers import AutoModel
        return AutoModel

    @property
    def has_pre_trained_weights(self) -> bool:
        return super().has_pre_trained_weights or (self.config.hf_path is not None and self.config.load_hf_weights)

    def merge_hf_config(self, hf_config: Any) -> None: ...

    def _common_hf_kwargs(self) -> dict[str, Any]:
        return {'trust_remote_code': self.config.trust_remote_code, 'revision': self.config.revision}

    def load_hf_config(self, **kwargs: Any):
        merged = self._common_hf_kwargs() | dict(self.config.hf_extra_kwargs) | kwargs
        return self._hf_config_class().from_pretrained(self.config.hf_path, **merged)

    def load_hf_model(self, **kwargs: Any):
        merged = self._common_hf_kwargs() | {
            'low_cpu_mem_usage': self.config.low_cpu_mem_usage,
            'torch_dtype': self.config.torch_dtype,
        } | dict(self.config.hf_extra_kwargs) | kwargs
        return self._hf_model_class().from_pretrained(self.config.hf_path, **merged)

    def load_hf_tokenizer(self, **kwargs: Any):
        from transformers import AutoTokenizer
        path = self.config.hf_tokenizer_path or self.config.hf_path
        return AutoTokenizer.from_pretrained(path, **(self._common_hf_kwargs() | kwargs))

    @contextmanager
    def torch_dtype_context(self) -> Generator[None, None, None]:
        previous = torch.get_default_dtype()
        wanted = self.config.torch_dtype
        torch.set_default_dtype(previous if wanted == 'auto' else wanted)
        try:
            yield
        finally:
            torch.set_default_dtype(previous)

    def build_hf_config(self):
        """HF config describing THIS model (used for export when no ``hf_path`` was given)."""
        raise NotImplementedError

    def construct_hf_model(self, **kwargs: Any):
        hf_config = self.hf_config if self.hf_config is not None else self.build_hf_config()
        model_class = self._hf_model_class()
        with self.torch_dtype_context():
            if hasattr(model_class, 'from_config'):
                model = model_class.from_config(hf_config, trust_remote_code=self.config.trust_remote_code, **kwargs)
            else:
                model = model_class(hf_config)
        if self.config.hf_path is not None:
            try:
                from transformers import GenerationConfig
                model.generation_config = GenerationConfig.from_pretrained(self.config.hf_path)
            except Exception:  # no generation config published with the checkpoint
                pass
        return model

    def convert_state_dict_from_hf(self, hf_state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return hf_state_dict

    def convert_state_dict_to_hf(self, state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return state_dict

    def get_hf_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        return self.convert_state_dict_from_hf(self.load_hf_model().state_dict())

    def get_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        if self.config.hf_path is not None:
            return self.get_hf_pre_trained_weights()
        return super().get_pre_trained_weights()

    def get_hf_model(self):
        with init_empty_weights(include_buffers=False):
            hf_model = self.construct_hf_model()
        state_dict = self.convert_state_dict_to_hf({k: v.detach() for k, v in self.state_dict().items()})
        hf_model.load_state_dict(state_dict, assign=True, strict=False)
        hf_model.tie_weights()
        return hf_model
