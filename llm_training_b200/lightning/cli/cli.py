"""``llm-training fit --config cfg.yaml [--trainer.num_nodes 2] [--ckpt_path ...]``.

Own implementation of the reference's ``LightningCLI`` subclass (src/llm_training/lightning/cli/cli.py:17-83):
top-level ``seed_everything``, ``float32_matmul_precision``, ``logging_level``, ``trainer``, ``model``, ``data``,
``output_redirection``, ``tqdm_progress``, ``ckpt_path``; ``ExtraConfig`` is callback 0; the resolved config
is saved and embedded in checkpoints by ``SaveConfigCallback``; ``SLURM_JOB_ID``/``SLURM_NTASKS`` are dropped
when ``SLURM_NTASKS == 1`` so an interactive allocation launches its own local ranks (cli.py:76-83).
"""
import logging
import os
import sys
from typing import Any

from ..callbacks import ExtraConfig, OutputRedirection, SaveConfigCallback, TQDMProgressBar
from ..trainer import Trainer, seed_everything
from .config import instantiate, instantiate_nested, load_config, parse_overrides, set_by_path

logger = logging.getLogger(__name__)

SUBCOMMANDS = ('fit', 'validate')


class LightningCLI:
    def __init__(self, model_class: Any | None = None, datamodule_class: Any | None = None,
                 save_config_callback: type[SaveConfigCallback] | None = SaveConfigCallback,
                 save_config_kwargs: dict[str, Any] | None = None, trainer_class: type[Trainer] = Trainer,
                 trainer_defaults: dict[str, Any] | None = None, seed_everything_default: bool | int = True,
                 parser_kwargs: dict[str, Any] | None = None, subclass_mode_model: bool = False,
                 subclass_mode_data: bool = False, args: list[str] | None = None, run: bool = True,
                 auto_configure_optimizers: bool = True) -> None:
        """Same keyword arguments as the reference's ``LightningCLI`` (lightning/cli/cli.py:18-33).  ``model_class`` /
        ``datamodule_class`` fix the class when the YAML section has no ``class_path``; ``seed_everything_default``
        is used when the config has no ``seed_everything``; ``save_config_kwargs`` go to the save-config callback;
        ``parser_kwargs`` / ``subclass_mode_*`` / ``auto_configure_optimizers`` are jsonargparse-specific and
        accepted for compatibility (every section is already resolved in subclass mode here)."""
        if isinstance(model_class, (list, tuple)):      # LightningCLI(argv): argument list passed positionally
            args, model_class = list(model_class), None
        self.model_class, self.datamodule_class = model_class, datamodule_class
        self.save_config_kwargs = save_config_kwargs or {}
        self.seed_everything_default = seed_everything_default
        argv = list(sys.argv[1:] if args is None else args)
        configs, overrides, positional = parse_overrides(argv)
        self.subcommand = positional[0] if positional and positional[0] in SUBCOMMANDS else None
        if run and self.subcommand is None:
            raise SystemExit(f'usage: llm-training {{{",".join(SUBCOMMANDS)}}} --config CONFIG [--key.subkey VALUE ...]')
        if not configs:
            raise SystemExit('at least one --config file is required')
        self.config = load_config(configs)
        for key, value in overrides.items():
            set_by_path(self.config, key, value)
        self.trainer_class = trainer_class
        self.save_config_callback = save_config_callback
        self.trainer_defaults = trainer_defaults or {}

        self.seed = None
        seed = self.config.get('seed_everything', self.seed_everything_default)
        if seed not in (None, False):
            self.seed = seed_everything(None if seed is True else int(seed))
        self.model = self._instantiate_section('model', self.model_class)
        self.datamodule = self._instantiate_section('data', self.datamodule_class) if self.config.get('data') else None
        self.trainer = self._instantiate_trainer()
        if run:
            self.run()

    def _instantiate_section(self, key: str, fixed_class: Any | None) -> Any:
        """A section is ``{class_path, init_args}``; with a fixed class (``model_class`` / ``datamodule_class``) it may
        also be the bare ``init_args`` mapping, as in Lightning's non-subclass mode."""
        spec = self.config[key]
        if fixed_class is not None and 'class_path' not in spec:
            from .config import instantiate_nested as _nested
            return fixed_class(**_nested(spec.get('init_args', spec)))
        return instantiate(spec)

    def add_arguments_to_parser(self, parser: Any = None) -> None:
        """Hook of the reference's jsonargparse-based CLI (cli.py:52-77) where it registers ``--float32-matmul-
        precision``, ``--logging-level`` and the ``output_redirection`` / ``tqdm_progress`` argument groups.  The YAML
        loader of this CLI reads those keys directly, so there is nothing to register; subclasses may still override
        this to post-process ``self.config`` before the trainer is built."""

    def _instantiate_trainer(self) -> Trainer:
        cfg = dict(self.trainer_defaults)
        cfg.update(self.config.get('trainer') or {})
        if os.environ.get('SLURM_NTASKS') == '1':
            os.environ.pop('SLURM_NTASKS', None)
            os.environ.pop('SLURM_JOB_ID', None)
        kwargs = instantiate_nested(cfg)
        callbacks = list(kwargs.pop('callbacks', None) or [])
        extra = ExtraConfig(float32_matmul_precision=self.config.get('float32_matmul_precision'),
                            logging_level=self.config.get('logging_level', logging.INFO))
        callbacks.insert(0, extra)
        redirect = self.config.get('output_redirection')
        callbacks.append(OutputRedirection(**(redirect or {})) if not isinstance(redirect, OutputRedirection)
                         else redirect)
        progress = self.config.get('tqdm_progress')
        if progress is not None or kwargs.get('enable_progress_bar', True):
            callbacks.append(TQDMProgressBar(**(progress or {})))
        self.add_arguments_to_parser(None)
        if self.save_config_callback is not None:
            callbacks.append(self.save_config_callback(config=self.config, **self.save_config_kwargs))
        return self.trainer_class(callbacks=callbacks, **kwargs)

    def run(self) -> None:
        ckpt_path = self.config.get('ckpt_path')
        if self.subcommand == 'validate':
            self.trainer.validate(self.model, self.datamodule)
        else:
            self.trainer.fit(self.model, self.datamodule, ckpt_path=ckpt_path)
