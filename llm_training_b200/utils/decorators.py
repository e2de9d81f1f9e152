"""Typing helper: declare that a method (typically ``__call__``) has the signature of another
(typically ``forward``) while still deferring to the parent class implementation
(reference: src/llm_training/utils/decorators.py)."""
import functools
from typing import Callable, ParamSpec, TypeVar

P = ParamSpec('P')
R = TypeVar('R')


def copy_method_signature(source: Callable[P, R]) -> Callable[[Callable], Callable[P, R]]:
    def decorator(stub: Callable) -> Callable[P, R]:
        name = stub.__name__
        owner_qualname = stub.__qualname__.rsplit('.', 1)[0]

        def passthrough(self, *args, **kwargs):
            # find the class that declared the stub, then call the next one in the MRO
            for i, klass in enumerate(type(self).__mro__):
                if klass.__qualname__ == owner_qualname and name in vars(klass):
                    return getattr(super(klass, self), name)(*args, **kwargs)
            raise AttributeError(name)

        functools.update_wrapper(passthrough, source, assigned=('__doc__', '__annotations__'), updated=())
        passthrough.__name__ = name
        passthrough.__qualname__ = stub.__qualname__
        return passthrough
    return decorator
