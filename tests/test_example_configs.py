"""Every example YAML — this repository's and, when the read-only reference tree is present, the reference's
ORIGINAL files — must resolve against this framework without touching the network: every ``class_path`` imports, every
``init_args`` key is a parameter of the class it configures, and every key of an objective / datamodule ``config``
block is a field of its pydantic model.  This is the "a reference user can switch over with their YAML" check."""
import glob
import inspect
import os
import typing

import pytest

from llm_training_b200.lightning.cli.config import _pydantic_in, load_config
from llm_training_b200.utils.imports import import_object

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS = sorted(glob.glob(os.path.join(ROOT, 'config', 'examples', '**', '*.yaml'), recursive=True))
THEIRS = sorted(glob.glob('/root/reference/config/examples/**/*.yaml', recursive=True))
SHORT = {   # names Lightning's CLI resolves without a module path
    'LearningRateMonitor': 'llm_training.lightning.LearningRateMonitor',
    'HFTokenizer': 'llm_training.lightning.HFTokenizer',
}


def _resolve(path: str):
    return import_object(SHORT.get(path, path))


def _check_spec(spec: dict, where: str, problems: list) -> None:
    try:
        target = _resolve(spec['class_path'])
    except Exception as e:  # noqa: BLE001
        problems.append(f'{where}: class_path {spec["class_path"]!r} does not import ({type(e).__name__}: {e})')
        return
    init_args = spec.get('init_args') or {}
    fn = target.__init__ if inspect.isclass(target) else target
    try:
        sig = inspect.signature(fn)
        hints = typing.get_type_hints(fn)
    except Exception:  # noqa: BLE001
        sig, hints = None, {}
    accepts_kwargs = sig is None or any(p.kind is p.VAR_KEYWORD for p in sig.parameters.values())
    for key, value in init_args.items():
        if sig is not None and key not in sig.parameters and not accepts_kwargs:
            problems.append(f'{where}: {spec["class_path"]} has no argument {key!r}')
            continue
        model_cls = _pydantic_in(hints.get(key))
        if model_cls is not None and isinstance(value, dict):
            unknown = set(value) - set(model_cls.model_fields)
            if unknown and model_cls.model_config.get('extra') != 'allow':
                problems.append(f'{where}.{key}: unknown fields {sorted(unknown)} for {model_cls.__name__}')
        _walk(value, f'{where}.{key}', problems)


def _walk(node, where: str, problems: list) -> None:
    if isinstance(node, dict):
        if 'class_path' in node:
            _check_spec(node, where, problems)
        else:
            for k, v in node.items():
                _walk(v, f'{where}.{k}', problems)
    elif isinstance(node, list):
        for i, v in enumerate(node):
            _walk(v, f'{where}[{i}]', problems)


def _check_dotted_classes(node, where: str, problems: list) -> None:
    """``model_class`` / ``optimizer_class`` / ``lr_scheduler_class`` strings inside objective configs."""
    if isinstance(node, dict):
        for k, v in node.items():
            if k in ('model_class', 'optimizer_class', 'lr_scheduler_class') and isinstance(v, str):
                if v.startswith('deepspeed.'):
                    continue            # mapped onto the fused AdamW by the sharded optimizer (optim/sharded_optimizer.py)
                try:
                    import_object(v)
                except Exception as e:  # noqa: BLE001
                    problems.append(f'{where}.{k}: {v!r} does not import ({type(e).__name__})')
            else:
                _check_dotted_classes(v, f'{where}.{k}', problems)
    elif isinstance(node, list):
        for i, v in enumerate(node):
            _check_dotted_classes(v, f'{where}[{i}]', problems)


@pytest.mark.parametrize('path', OURS + THEIRS, ids=lambda p: '/'.join(p.split('/')[-3:]))
def test_example_config_resolves(path):
    cfg = load_config([path])
    assert {'trainer', 'model', 'data'} <= set(cfg), sorted(cfg)
    problems: list = []
    _walk(cfg['model'], 'model', problems)
    _walk(cfg['data'], 'data', problems)
    _walk(cfg['trainer'].get('strategy'), 'trainer.strategy', problems)
    _walk(cfg['trainer'].get('logger'), 'trainer.logger', problems)
    _walk(cfg['trainer'].get('callbacks'), 'trainer.callbacks', problems)
    _check_dotted_classes(cfg['model'], 'model', problems)
    from llm_training_b200.lightning import Trainer
    trainer_args = set(inspect.signature(Trainer.__init__).parameters)
    for key in cfg['trainer']:
        if key not in trainer_args:
            problems.append(f'trainer: unknown argument {key!r}')
    assert not problems, '\n'.join(problems)
