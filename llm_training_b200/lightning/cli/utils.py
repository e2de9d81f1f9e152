"""``HFTokenizer``: ``AutoTokenizer.from_pretrained`` exposed as a class path for YAML configs
(reference: src/llm_training/lightning/cli/utils.py:7-22)."""
from typing import Any


def HFTokenizer(path: str, pad_token: str | None = None, padding_side: str | None = None, **kwargs: Any):
    from transformers import AutoTokenizer
    tokenizer = AutoTokenizer.from_pretrained(path, **kwargs)
    if pad_token is not None:
        tokenizer.pad_token = pad_token
    if padding_side is not None:
        tokenizer.padding_side = padding_side
    return tokenizer
