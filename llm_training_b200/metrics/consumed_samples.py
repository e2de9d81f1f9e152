"""Import-path parity with the reference (``llm_training.metrics.consumed_samples.ConsumedSamples``)."""
from .consumed import ConsumedSamples

__all__ = ['ConsumedSamples']
