"""Perplexity from either logits+targets or an already-computed mean loss
(reference: src/llm_training/metrics/perplexity.py:8-39)."""
import torch
import torch.nn.functional as F

from .metric import Metric


class Perplexity(Metric):
    higher_is_better = False

    def __init__(self, ignore_index: int | None = None, **kwargs) -> None:
        super().__init__(**kwargs)
        if ignore_index is not None and not isinstance(ignore_index, int):
            raise ValueError(f'`ignore_index` must be None or int, got {ignore_index!r}')
        self.ignore_index = ignore_index
        self.add_state('total_log_probs', torch.tensor(0.0, dtype=torch.float64))
        self.add_state('count', torch.tensor(0.0, dtype=torch.float64))

    def update(self, preds_or_loss: torch.Tensor, target: torch.Tensor | None = None) -> None:
        if preds_or_loss.dim() == 0:
            self.total_log_probs += preds_or_loss.detach().double()
            self.count += 1
            return
        assert target is not None
        logits = preds_or_loss.detach().float().flatten(end_dim=-2)
        target = target.flatten()
        ignore = -100 if self.ignore_index is None else self.ignore_index
        nll = F.cross_entropy(logits, target, ignore_index=ignore, reduction='sum')
        self.total_log_probs += nll.double()
        self.count += target.ne(ignore).sum().double()

    def _compute(self, states):
        return torch.exp(states['total_log_probs'] / states['count']).float()
