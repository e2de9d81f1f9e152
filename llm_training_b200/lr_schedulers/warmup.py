"""Import-path parity with the reference (``llm_training.lr_schedulers.warmup.WarmupLR``); the implementation lives in
``schedules.py`` (closed-form schedules)."""
from .schedules import WarmupLR

__all__ = ['WarmupLR']
