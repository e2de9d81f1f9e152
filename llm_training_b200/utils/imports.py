"""Dotted-path → object resolution with an alias table.

The reference relies on jsonargparse to turn ``optimizer_class: torch.optim.AdamW`` into a class
(reference: src/llm_training/lms/base_lm_config.py:13-19).  Here the same strings are resolved by
hand; third-party classes whose job is done natively are mapped to the in-tree implementation.
"""
import importlib
from typing import Any

# third-party path  ->  in-tree replacement
CLASS_ALIASES: dict[str, str] = {
    'deepspeed.ops.adam.FusedAdam': 'llm_training_b200.optim.FusedAdamW',
    'deepspeed.ops.adam.DeepSpeedCPUAdam': 'llm_training_b200.optim.FusedAdamW',
    'lightning.pytorch.callbacks.LearningRateMonitor': 'llm_training_b200.lightning.LearningRateMonitor',
    'lightning.pytorch.callbacks.ModelCheckpoint': 'llm_training_b200.lightning.ModelCheckpoint',
    'lightning.pytorch.loggers.WandbLogger': 'llm_training_b200.lightning.WandbLogger',
    'lightning.pytorch.loggers.CSVLogger': 'llm_training_b200.lightning.CSVLogger',
}

# modules searched for un-dotted names such as ``class_path: LearningRateMonitor``
SHORT_NAME_MODULES: list[str] = [
    'llm_training_b200.lightning',
    'llm_training_b200.lms',
    'llm_training_b200.data',
    'llm_training_b200.models',
    'llm_training_b200.lr_schedulers',
    'llm_training_b200.optim',
    'torch.optim',
    'torch.optim.lr_scheduler',
]


def import_object(path: str) -> Any:
    """Import ``pkg.mod.Name`` (also ``pkg.mod.Class.attr``)."""
    path = CLASS_ALIASES.get(path, path)
    if path.startswith('llm_training.'):
        path = 'llm_training_b200.' + path[len('llm_training.'):]
    if '.' not in path:
        for module_name in SHORT_NAME_MODULES:
            module = importlib.import_module(module_name)
            if hasattr(module, path):
                return getattr(module, path)
        raise ImportError(f'cannot resolve short name `{path}`; searched {SHORT_NAME_MODULES}')
    parts = path.split('.')
    last_error: Exception | None = None
    for split in range(len(parts) - 1, 0, -1):
        module_name = '.'.join(parts[:split])
        try:
            obj = importlib.import_module(module_name)
        except ImportError as e:
            last_error = e
            continue
        try:
            for attr in parts[split:]:
                obj = getattr(obj, attr)
        except AttributeError as e:
            last_error = e
            continue
        return obj
    raise ImportError(f'cannot import `{path}`: {last_error}')


def resolve_class(value: Any) -> Any:
    """Strings become the object they name; everything else passes through."""
    if isinstance(value, str):
        return import_object(value)
    return value
