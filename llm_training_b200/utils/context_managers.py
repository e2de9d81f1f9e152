"""Enter several context managers as one (reference: src/llm_training/utils/context_managers.py)."""
from contextlib import ExitStack
from typing import ContextManager, Iterable


class ContextManagers:
    def __init__(self, context_managers: Iterable[ContextManager]) -> None:
        self.context_managers = list(context_managers)
        self._stack: ExitStack | None = None

    def __enter__(self) -> "ContextManagers":
        self._stack = ExitStack()
        self._stack.__enter__()
        try:
            for cm in self.context_managers:
                self._stack.enter_context(cm)
        except BaseException:
            self._stack.close()
            raise
        return self

    def __exit__(self, *exc) -> bool | None:
        assert self._stack is not None
        return self._stack.__exit__(*exc)
