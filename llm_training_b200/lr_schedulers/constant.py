"""Import-path parity with the reference (``llm_training.lr_schedulers.constant.ConstantWarmupLR``); the implementation lives in
``schedules.py`` (closed-form schedules)."""
from .schedules import ConstantWarmupLR

__all__ = ['ConstantWarmupLR']
