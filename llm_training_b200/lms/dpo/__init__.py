from .dpo import DPO
from .dpo_config import DPOConfig
