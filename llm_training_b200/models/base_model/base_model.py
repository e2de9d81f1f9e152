"""Base class of every model (reference: src/llm_training/models/base_model/base_model.py:14-74)."""
from contextlib import contextmanager
from typing import Any, ClassVar, Generator

import torch
from torch import nn

from .base_model_config import BaseModelConfig


class BaseModel(nn.Module):
    _init_weights: ClassVar[bool] = True

    config_class: type[BaseModelConfig] = BaseModelConfig
    no_split_modules: list[str] = []

    def __init__(self, config: BaseModelConfig) -> None:
        super().__init__()
        self.config = config

    def __init_subclass__(cls, **kwargs) -> None:
        """Run ``init_weights`` after the *outermost* subclass constructor (the reference calls it
        from ``BaseModel.__init__`` before any submodule exists, so it initialises nothing there and
        relies on nn defaults; here ``init_weights: true`` really applies the model's initialiser)."""
        super().__init_subclass__(**kwargs)
        original_init = cls.__init__

        def wrapped_init(self, *args, **kw):
            depth = getattr(self, '_ctor_depth', 0)
            object.__setattr__(self, '_ctor_depth', depth + 1)
            try:
                original_init(self, *args, **kw)
            finally:
                object.__setattr__(self, '_ctor_depth', depth)
            if depth == 0 and BaseModel._init_weights and type(self)._init_weights:
                if not any(p.is_meta for p in self.parameters()):
                    self.init_weights()

        wrapped_init.__wrapped__ = original_init
        cls.__init__ = wrapped_init

    @property
    def has_pre_trained_weights(self) -> bool:
        return self.config.pre_trained_weights is not None

    def get_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        import safetensors.torch
        return safetensors.torch.load_file(self.config.pre_trained_weights)

    def _init_weights_impl(self, module: nn.Module) -> None: ...

    @torch.no_grad()
    def init_weights(self) -> None:
        self.apply(self._init_weights_impl)

    @classmethod
    @contextmanager
    def init_weights_context(cls, init_weights: bool) -> Generator[None, None, None]:
        previous = BaseModel._init_weights
        BaseModel._init_weights = bool(init_weights)
        try:
            yield
        finally:
            BaseModel._init_weights = previous

    # -- parallelisation hooks (reference :48-74); ``dp_mesh``/``tp_mesh`` are parallel.Mesh axes ------
    def configure_tensor_parallel(self, tp_mesh: Any) -> None:
        if tp_mesh is None or tp_mesh.size() == 1:
            return
        raise NotImplementedError(f'`{type(self).__name__}` does not support tensor parallel.')

    def configure_fully_sharded_data_parallel(self, dp_mesh: Any, **kwargs: Any) -> None:
        if dp_mesh is None or dp_mesh.size() == 1:
            return
        raise NotImplementedError(f'`{type(self).__name__}` does not support fully sharded data parallel.')

    def sharding_units(self) -> list[tuple[str, nn.Module]]:
        """(name, module) groups whose parameters are gathered / reduced together (one flat buffer each)."""
        return [('', self)]

    def parallelize(self, dp_mesh: Any, tp_mesh: Any, **fsdp_kwargs: Any) -> None:
        self.configure_tensor_parallel(tp_mesh)
        self.configure_fully_sharded_data_parallel(dp_mesh, **fsdp_kwargs)
