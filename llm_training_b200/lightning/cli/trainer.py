"""Import location kept from the reference (src/llm_training/lightning/cli/trainer.py)."""
from ..trainer import Trainer  # noqa: F401
