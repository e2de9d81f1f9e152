from .base_model import BaseModel, BaseModelConfig
from .hf_causal_lm import HFCausalLM, HFCausalLMConfig
from .hf_compat_model import HFCompatModel, HFCompatModelConfig
from .llama import Llama, LlamaConfig
from .phi3 import Phi3, Phi3Config
