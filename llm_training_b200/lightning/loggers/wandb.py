"""W&B logger with the reference's directory convention: ``save_dir == log_dir ==
<save_dir>/<project>/<name>`` created lazily (src/llm_training/lightning/loggers/wandb.py:10-72).
When ``wandb`` cannot be imported or no network/API key is available the run is recorded offline
(``WANDB_MODE=offline``) and metrics are mirrored to ``metrics.jsonl`` in the same directory."""
import json
import logging
import os
from typing import Any

logger = logging.getLogger(__name__)


class WandbLogger:
    def __init__(self, name: str | None = None, save_dir: str = '.', version: str | None = None,
                 offline: bool = False, dir: str | None = None, id: str | None = None, anonymous: bool | None = None,
                 project: str | None = None, log_model: bool | str = False, experiment: Any | None = None,
                 prefix: str = '', checkpoint_name: str | None = None, job_type: str | None = None,
                 save_code: bool | None = None, tags: list[str] | None = None, entity: str | None = None,
                 notes: str | None = None, group: str | None = None, resume: str | bool | None = None,
                 **kwargs: Any) -> None:
        self._name = name
        self._project = project or os.environ.get('WANDB_PROJECT', 'lightning_logs')
        self._root = dir or save_dir
        self._version = version or id
        self._offline = offline or os.environ.get('WANDB_MODE') == 'offline'
        self._prefix = prefix
        self._experiment = experiment
        self._init_kwargs = dict(job_type=job_type, save_code=save_code, tags=tags, entity=entity, notes=notes,
                                 group=group, resume=resume, **kwargs)
        self._jsonl = None

    @property
    def name(self) -> str | None:
        return self._name

    @property
    def version(self) -> str | None:
        return self._version

    @property
    def save_dir(self) -> str:
        path = os.path.join(self._root, self._project, self._name or 'run')
        os.makedirs(path, exist_ok=True)
        return path

    @property
    def log_dir(self) -> str:
        return self.save_dir

    @property
    def experiment(self) -> Any:
        if self._experiment is None:
            try:
                import wandb
                if self._offline or not (os.environ.get('WANDB_API_KEY') or os.path.exists(os.path.expanduser('~/.netrc'))):
                    os.environ.setdefault('WANDB_MODE', 'offline')
                kw = {k: v for k, v in self._init_kwargs.items() if v is not None}
                self._experiment = wandb.init(name=self._name, project=self._project, dir=self.save_dir,
                                              id=self._version, **kw)
            except Exception as e:   # wandb missing / cannot start: local mirror only
                logger.warning(f'wandb unavailable ({e!r}); logging to {self.save_dir}/metrics.jsonl only')
                self._experiment = _NullRun()
        return self._experiment

    def _mirror(self, record: dict) -> None:
        if self._jsonl is None:
            self._jsonl = open(os.path.join(self.save_dir, 'metrics.jsonl'), 'a', buffering=1)
        self._jsonl.write(json.dumps(record) + '\n')

    def log_hyperparams(self, params: dict[str, Any]) -> None:
        try:
            self.experiment.config.update(params, allow_val_change=True)
        except Exception:
            pass
        self._mirror({'hparams': params})

    def log_metrics(self, metrics: dict[str, float], step: int | None = None) -> None:
        metrics = {self._prefix + k: v for k, v in metrics.items()}
        self._mirror({'step': step, **metrics})
        try:
            self.experiment.log(dict(metrics, **({'trainer/global_step': step} if step is not None else {})))
        except Exception:
            pass

    def save_file(self, path: str) -> None:
        try:
            self.experiment.save(path, policy='now')
        except Exception:
            pass

    def log_code(self, root: str | None = None) -> list[str]:
        """Upload the training source tree with the run (reference: save_config_callback.py:39-49 uploads ``src/`` and
        ``scripts/``, ``.py/.sh/.j2`` only).  Here the tree is this package, ``scripts/`` and the kernel sources
        (``csrc/``, ``include/``: ``.cu/.cuh/.cpp/.h``).  Returns the relative paths that were selected (also written
        to ``code_files.txt`` beside the metrics mirror, so offline runs keep the record)."""
        root = root or os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
        tops = ('llm_training_b200', 'llm_training', 'scripts', 'csrc', 'include')
        suffixes = ('.py', '.sh', '.j2', '.cu', '.cuh', '.cpp', '.h')

        def include(path: str, base: str = root) -> bool:
            rel = os.path.relpath(path, base)
            return rel.split(os.sep)[0] in tops and rel.endswith(suffixes)

        selected = sorted(os.path.relpath(os.path.join(d, f), root) for top in tops
                          for d, _, files in os.walk(os.path.join(root, top)) for f in files
                          if include(os.path.join(d, f)))
        with open(os.path.join(self.save_dir, 'code_files.txt'), 'w') as f:
            f.write('\n'.join(selected) + '\n')
        try:
            self.experiment.log_code(root=root, include_fn=include)
        except Exception as e:   # offline / stub run
            logger.debug(f'wandb log_code skipped: {e!r}')
        return selected

    def finalize(self, status: str) -> None:
        if self._jsonl is not None:
            self._jsonl.close()
            self._jsonl = None
        if self._experiment is not None and hasattr(self._experiment, 'finish'):
            try:
                self._experiment.finish()
            except Exception:
                pass


class _NullRun:
    class _Cfg(dict):
        def update(self, *a, **k):   # noqa: D401
            dict.update(self, *a)
    config = _Cfg()

    def log(self, *a, **k): ...
    def save(self, *a, **k): ...
    def log_code(self, *a, **k): ...
    def finish(self): ...
