from .metric import Metric
from .consumed import ConsumedSamples, ConsumedTokens
from .perplexity import Perplexity

__all__ = ['Metric', 'ConsumedSamples', 'ConsumedTokens', 'Perplexity']
