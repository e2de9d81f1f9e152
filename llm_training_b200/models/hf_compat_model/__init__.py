from .hf_compat_config import HFCompatModelConfig
from .hf_compat_model import HFCompatModel
