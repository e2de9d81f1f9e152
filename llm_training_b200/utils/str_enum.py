"""Case-insensitive string enum (reference: src/llm_training/utils/str_enum.py)."""
from enum import Enum


class StrEnum(str, Enum):
    def __str__(self) -> str:
        return str(self.value)

    @classmethod
    def _missing_(cls, value):
        if isinstance(value, str):
            folded = value.lower()
            for member in cls:
                if member.value.lower() == folded or member.name.lower() == folded:
                    return member
        return None
