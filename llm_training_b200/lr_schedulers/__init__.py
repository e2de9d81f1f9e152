from .schedules import (ConstantWarmupLR, CosineAnnealingWarmupLR,
                        LinearWarmupLR, WarmupLR)

__all__ = ['WarmupLR', 'ConstantWarmupLR', 'CosineAnnealingWarmupLR', 'LinearWarmupLR']
