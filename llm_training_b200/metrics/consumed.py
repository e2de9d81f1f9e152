"""Running counters of samples / supervised tokens seen, saved in checkpoints
(reference: src/llm_training/metrics/consumed_samples.py:5-20, consumed_tokens.py:5-21)."""
import torch

from .metric import Metric


class ConsumedSamples(Metric):
    higher_is_better = True

    def __init__(self, **kwargs) -> None:
        super().__init__(**kwargs)
        self.add_state('n', torch.tensor(0, dtype=torch.int64), persistent=True)

    def update(self, target: torch.Tensor) -> None:
        self.n += target.size(0)

    def _compute(self, states):
        return states['n']


class ConsumedTokens(Metric):
    higher_is_better = True

    def __init__(self, ignore_index: int = -100, **kwargs) -> None:
        super().__init__(**kwargs)
        self.ignore_index = ignore_index
        self.add_state('n', torch.tensor(0, dtype=torch.int64), persistent=True)

    def update(self, target: torch.Tensor) -> None:
        self.n += target.ne(self.ignore_index).sum()   # stays on device: no host sync

    def _compute(self, states):
        return states['n']
