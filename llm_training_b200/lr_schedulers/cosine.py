"""Import-path parity with the reference (``llm_training.lr_schedulers.cosine.CosineAnnealingWarmupLR``); the implementation lives in
``schedules.py`` (closed-form schedules)."""
from .schedules import CosineAnnealingWarmupLR

__all__ = ['CosineAnnealingWarmupLR']
