"""Import-path parity with the reference (``llm_training.metrics.consumed_tokens.ConsumedTokens``)."""
from .consumed import ConsumedTokens

__all__ = ['ConsumedTokens']
