"""Training objectives (reference: src/llm_training/lms/__init__.py:1-5)."""
from .base_lm import BaseLightningModule
from .base_lm_config import BaseLightningModuleConfig, BaseOptimizerConfig
from .clm import CLM, CLMConfig
from .dpo import DPO, DPOConfig
from .model_provider import ModelProvider
from .orpo import ORPO, ORPOConfig

__all__ = ['BaseLightningModule', 'BaseLightningModuleConfig', 'BaseOptimizerConfig', 'CLM', 'CLMConfig', 'DPO',
           'DPOConfig', 'ORPO', 'ORPOConfig', 'ModelProvider']
