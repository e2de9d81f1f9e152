from .context_managers import ContextManagers
from .decorators import copy_method_signature
from .imports import import_object, resolve_class
from .str_enum import StrEnum
