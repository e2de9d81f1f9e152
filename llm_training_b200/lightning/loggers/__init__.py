from .csv_logger import CSVLogger
from .wandb import WandbLogger

__all__ = ['CSVLogger', 'WandbLogger']
