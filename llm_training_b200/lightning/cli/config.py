"""YAML + command-line configuration without jsonargparse / omegaconf.

Supported, because the reference's configs use it:
  * ``class_path`` / ``init_args`` objects, nested anywhere; short names (``LearningRateMonitor``, ``HFTokenizer``)
  * dotted keys inside YAML (``init_args.config:``) — expanded into nested mappings
  * ``--a.b.c value`` / ``--a.b.c=value`` overrides with YAML-typed values, several ``--config`` files merged
  * constructor arguments annotated with a pydantic model are validated from plain dicts
"""
import copy
import inspect
import types
import typing
from typing import Any

import yaml
from pydantic import BaseModel as PydanticModel

from ...utils.imports import import_object


class _Loader(yaml.SafeLoader):
    pass


# YAML 1.1 reads ``1e-5`` as a string; the reference's configs rely on it being a float
_Loader.add_implicit_resolver(
    'tag:yaml.org,2002:float',
    __import__('re').compile(r'''^(?:[-+]?(?:[0-9][0-9_]*)\.[0-9_]*(?:[eE][-+]?[0-9]+)?
                                 |[-+]?(?:[0-9][0-9_]*)(?:[eE][-+]?[0-9]+)
                                 |\.[0-9_]+(?:[eE][-+][0-9]+)?
                                 |[-+]?\.(?:inf|Inf|INF)|\.(?:nan|NaN|NAN))$''', __import__('re').X),
    list('-+0123456789.'))


def parse_value(text: str) -> Any:
    try:
        return yaml.load(text, Loader=_Loader)
    except yaml.YAMLError:
        return text


def expand_dotted(obj: Any) -> Any:
    """{'init_args.config': x} -> {'init_args': {'config': x}} (recursively, merging siblings)."""
    if isinstance(obj, list):
        return [expand_dotted(v) for v in obj]
    if not isinstance(obj, dict):
        return obj
    out: dict[str, Any] = {}
    for key, value in obj.items():
        value = expand_dotted(value)
        parts = str(key).split('.') if isinstance(key, str) else [key]
        node = out
        for part in parts[:-1]:
            node = node.setdefault(part, {})
        leaf = parts[-1]
        if isinstance(node.get(leaf), dict) and isinstance(value, dict):
            node[leaf] = deep_merge(node[leaf], value)
        else:
            node[leaf] = value
    return out


def deep_merge(base: Any, override: Any) -> Any:
    if isinstance(base, dict) and isinstance(override, dict):
        # replacing an object by one of another class drops the old init_args
        if 'class_path' in override and override.get('class_path') != base.get('class_path'):
            return copy.deepcopy(override)
        merged = dict(base)
        for k, v in override.items():
            merged[k] = deep_merge(base[k], v) if k in base else copy.deepcopy(v)
        return merged
    return copy.deepcopy(override)


def set_by_path(cfg: dict, dotted: str, value: Any) -> None:
    parts = dotted.split('.')
    node: Any = cfg
    for i, part in enumerate(parts[:-1]):
        if isinstance(node, list):
            node = node[int(part)]
            continue
        nxt = node.get(part)
        if not isinstance(nxt, (dict, list)):
            nxt = {}
            node[part] = nxt
        node = nxt
    last = parts[-1]
    if isinstance(node, list):
        node[int(last)] = value
    else:
        node[last] = value


def load_config(paths: list[str]) -> dict[str, Any]:
    cfg: dict[str, Any] = {}
    for path in paths:
        with open(path) as f:
            cfg = deep_merge(cfg, expand_dotted(yaml.load(f, Loader=_Loader) or {}))
    return cfg


def parse_overrides(argv: list[str]) -> tuple[list[str], dict[str, Any], list[str]]:
    """-> (config files, {dotted.key: value}, positional words)."""
    configs: list[str] = []
    overrides: dict[str, Any] = {}
    positional: list[str] = []
    i = 0
    while i < len(argv):
        arg = argv[i]
        if arg.startswith('--'):
            key, eq, val = arg[2:].partition('=')
            if not eq:
                if i + 1 < len(argv) and not argv[i + 1].startswith('--'):
                    val = argv[i + 1]
                    i += 1
                else:
                    val = 'true'
            key = key.replace('-', '_') if '.' not in key else key
            if key in ('config', 'c'):
                configs.append(val)
            else:
                overrides[key] = parse_value(val)
        elif arg in ('-c',) and i + 1 < len(argv):
            configs.append(argv[i + 1])
            i += 1
        else:
            positional.append(arg)
        i += 1
    return configs, overrides, positional


# ---------------------------------------------------------------------------------------------------
# object construction
# ---------------------------------------------------------------------------------------------------
def _pydantic_in(annotation: Any) -> type[PydanticModel] | None:
    if inspect.isclass(annotation) and issubclass(annotation, PydanticModel):
        return annotation
    if isinstance(annotation, types.UnionType) or typing.get_origin(annotation) is typing.Union:
        for a in typing.get_args(annotation):
            found = _pydantic_in(a)
            if found is not None:
                return found
    return None


def instantiate_nested(value: Any) -> Any:
    """Replace every ``{class_path: ..., init_args: ...}`` mapping (at any depth) by the object it describes."""
    if isinstance(value, dict):
        if 'class_path' in value and set(value) <= {'class_path', 'init_args', 'dict_kwargs'}:
            return instantiate(value)
        return {k: instantiate_nested(v) for k, v in value.items()}
    if isinstance(value, list):
        return [instantiate_nested(v) for v in value]
    return value


def instantiate(spec: dict[str, Any] | str, **extra: Any) -> Any:
    if isinstance(spec, str):
        spec = {'class_path': spec}
    target = import_object(spec['class_path'])
    kwargs = dict(instantiate_nested(spec.get('init_args') or {}))
    kwargs.update(spec.get('dict_kwargs') or {})
    kwargs.update(extra)
    try:
        hints = typing.get_type_hints(target.__init__ if inspect.isclass(target) else target)
    except Exception:
        hints = {}
    for name, value in list(kwargs.items()):
        model_cls = _pydantic_in(hints.get(name))
        if model_cls is not None and isinstance(value, dict):
            kwargs[name] = model_cls.model_validate(value)
    return target(**kwargs)
