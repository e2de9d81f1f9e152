"""Saves the resolved YAML next to the logs, logs it as hyper-parameters (+ SLURM ids) and — the part
``convert_to_hf.py`` relies on — embeds it in EVERY checkpoint under ``checkpoint['config']``
(reference: src/llm_training/lightning/callbacks/save_config_callback.py:12-49)."""
import os
from typing import Any

import yaml

from .base import Callback


class SaveConfigCallback(Callback):
    def __init__(self, config: dict[str, Any] | None = None, config_filename: str = 'config.yaml',
                 overwrite: bool = True, save_to_log_dir: bool = True, **_: Any) -> None:
        self.config = config or {}
        self.config_filename = config_filename
        self.overwrite = overwrite
        self.save_to_log_dir = save_to_log_dir
        self.yaml_config: dict[str, Any] | None = None
        self._done = False

    def setup(self, trainer, pl_module, stage: str) -> None:
        if self._done:
            return
        self._done = True
        self.yaml_config = yaml.safe_load(yaml.safe_dump(self.config, sort_keys=False)) or {}
        self.yaml_config['world_size'] = trainer.world_size
        if 'SLURM_JOB_ID' in os.environ:
            self.yaml_config['slurm_job_id'] = os.environ['SLURM_JOB_ID']
            self.yaml_config['slurm_job_name'] = os.environ.get('SLURM_JOB_NAME')
            self.yaml_config['slurm_num_nodes'] = os.environ.get('SLURM_NNODES')
            self.yaml_config['slurm_ntasks'] = os.environ.get('SLURM_NTASKS')
        self.save_config(trainer, pl_module, stage)

    def save_config(self, trainer, pl_module, stage: str) -> None:
        """Write the YAML next to the logs and hand it to the logger (reference: save_config_callback.py:20-46)."""
        if trainer.is_global_zero and self.save_to_log_dir:
            logger = trainer.logger
            save_dir = getattr(logger, 'save_dir', None) or trainer.log_dir
            os.makedirs(save_dir, exist_ok=True)
            path = os.path.join(save_dir, self.config_filename)
            if self.overwrite or not os.path.exists(path):
                with open(path, 'w') as f:
                    yaml.safe_dump(self.config, f, sort_keys=False)
            if logger is not None and hasattr(logger, 'log_hyperparams'):
                logger.log_hyperparams(self.yaml_config)
            if logger is not None and hasattr(logger, 'save_file'):
                logger.save_file(path)
            if logger is not None and hasattr(logger, 'log_code'):
                logger.log_code()          # source tree upload (reference :40)

    def on_save_checkpoint(self, trainer, pl_module, checkpoint: dict[str, Any]) -> None:
        checkpoint['config'] = trainer.strategy.broadcast(self.yaml_config if trainer.is_global_zero else None, src=0)
