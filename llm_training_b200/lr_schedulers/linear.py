"""Import-path parity with the reference (``llm_training.lr_schedulers.linear.LinearWarmupLR``); the implementation lives in
``schedules.py`` (closed-form schedules)."""
from .schedules import LinearWarmupLR

__all__ = ['LinearWarmupLR']
