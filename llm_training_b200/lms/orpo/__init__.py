from .orpo import ORPO
from .orpo_config import ORPOConfig
