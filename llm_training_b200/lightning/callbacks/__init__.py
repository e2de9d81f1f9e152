from .base import Callback
from .extra_config import ExtraConfig
from .lr_monitor import LearningRateMonitor
from .model_checkpoint import ModelCheckpoint
from .output_redirection import OutputRedirection
from .save_config_callback import SaveConfigCallback
from .step_time_monitor import StepTimeMonitor
from .tqdm_progress import TQDMProgressBar
from .training_time_estimator import TrainingTimeEstimator

__all__ = ['Callback', 'ExtraConfig', 'LearningRateMonitor', 'ModelCheckpoint', 'OutputRedirection',
           'SaveConfigCallback', 'StepTimeMonitor', 'TQDMProgressBar', 'TrainingTimeEstimator']
