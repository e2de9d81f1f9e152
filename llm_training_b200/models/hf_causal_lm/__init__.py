from .hf_causal_lm import HFCausalLM
from .hf_causal_lm_config import HFCausalLMConfig
