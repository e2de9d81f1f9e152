"""Name-compatibility shim.

Existing YAML configs written for the reference say ``class_path: llm_training.lms.CLM``
(reference: config/examples/llama-3.1/llama-3.1-8b_pt_example.yaml:24).  The implementation
lives in :mod:`llm_training_b200`; this package installs an import hook so that every
``llm_training.X`` module *is* the ``llm_training_b200.X`` module object (no double import,
class identity is preserved).
"""
import importlib
import importlib.abc
import importlib.util
import sys

_ALIAS = 'llm_training'
_REAL = 'llm_training_b200'


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, real_name: str) -> None:
        self.real_name = real_name

    def create_module(self, spec):
        module = importlib.import_module(self.real_name)
        self._real_spec = module.__spec__
        return module

    def exec_module(self, module) -> None:
        # already executed under its real name; undo importlib's __spec__ overwrite
        module.__spec__ = self._real_spec


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(_ALIAS + '.'):
            return None
        real_name = _REAL + fullname[len(_ALIAS):]
        if importlib.util.find_spec(real_name) is None:
            return None
        return importlib.util.spec_from_loader(fullname, _AliasLoader(real_name))


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())

_real = importlib.import_module(_REAL)
# keep the module's real identity (its __spec__/__name__ must stay consistent for
# submodule imports made under the real name), just register it under the alias too
sys.modules[_ALIAS] = _real
