from .llama_config import LlamaConfig
from .llama_model import Llama
