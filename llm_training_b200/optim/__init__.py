from .fused_adamw import FusedAdamW
from .master_weight_wrapper import MasterWeightsOptimizer
from .sharded_optimizer import ShardedOptimizer

__all__ = ['FusedAdamW', 'MasterWeightsOptimizer', 'ShardedOptimizer']
