"""Odds-Ratio Preference Optimisation (reference: src/llm_training/lms/orpo/orpo.py:35-249): one model,
length-normalised log-probs, odds-ratio term + CE on the chosen sequence.  Both terms are derived from the
same fused per-token log-prob pass (CE on chosen == -mean log p over supervised tokens), so the chosen
sequence costs one LM-head pass instead of the reference's log-softmax + separate cross-entropy."""
import gc
import logging
from typing import Any

import torch
import torch.nn.functional as F

from ...models.base_model import BaseModel
from ...ops import shift_labels
from ..base_lm import BaseLightningModule
from ..dpo.dpo import sequence_logps
from ..utils import get_model
from .orpo_config import ORPOConfig

logger = logging.getLogger(__name__)


class ORPO(BaseLightningModule):
    config: ORPOConfig

    def __init__(self, config: ORPOConfig) -> None:
        super().__init__(config)
        self.model: BaseModel | None = None

    @property
    def has_pre_trained_weights(self) -> bool:
        return self.model is not None and self.model.has_pre_trained_weights

    def get_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        return {f'model.{k}': v for k, v in self.model.get_pre_trained_weights().items()}

    def configure_model(self) -> None:
        self.model = get_model(self.config.model)
        if self.global_rank == 0:
            logger.info(f'Config:\n{self.config!r}')

    def on_fsdp_parallelize_model(self, **kwargs: Any) -> None:
        self.model.parallelize(**kwargs)

    def get_model(self) -> BaseModel:
        return self.model

    def forward_batch(self, batch: dict[str, torch.Tensor | Any]) -> dict[str, Any]:
        ii = self.config.ignore_index
        out: dict[str, Any] = {}
        for side in ('chosen', 'rejected'):
            labels = shift_labels(batch[f'{side}_labels'], ii)
            logps, model_out = sequence_logps(self.model, batch[f'{side}_input_ids'],
                                              batch.get(f'{side}_attention_mask'),
                                              batch.get(f'{side}_position_ids'), labels, ii, average=True)
            out[f'{side}_logps'] = logps
            out[f'{side}_labels'] = labels
            out[f'{side}_mean_logit'] = self._mean_logit(model_out)
        return out

    def get_logps(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Length-normalised label log-probabilities from materialised logits (``labels`` already shifted), the
        reference's helper (lms/orpo/orpo.py:61-93).  The training path does not call it: ``forward_batch`` uses the
        fused linear + log-prob op on the hidden states instead."""
        mask = labels != self.config.ignore_index
        logp = torch.log_softmax(logits.float(), dim=-1)
        picked = logp.gather(-1, labels.clamp_min(0).unsqueeze(-1)).squeeze(-1) * mask
        return picked.sum(-1) / mask.sum(-1).clamp_min(1)

    def add_suffix_to_metrics(self, metrics: dict[str, torch.Tensor], suffix: str) -> dict[str, torch.Tensor]:
        return {k + suffix: v for k, v in metrics.items()}

    def _mean_logit(self, model_out) -> torch.Tensor:
        if model_out.logits is not None:
            return model_out.logits.detach().float().mean()
        with torch.no_grad():   # mean over the vocabulary of h.W^T == h . mean(W)
            w_mean = self.model.get_output_embeddings().weight.float().mean(0)
            return (model_out.normed_hidden_states.float() @ w_mean).mean()

    def compute_loss(self, chosen_logps: torch.Tensor, chosen_labels: torch.Tensor, rejected_logps: torch.Tensor,
                     rejected_labels: torch.Tensor, chosen_mean_logit: torch.Tensor,
                     rejected_mean_logit: torch.Tensor, metrics: dict[str, torch.Tensor]) -> torch.Tensor:
        beta, ii = self.config.beta, self.config.ignore_index
        log_odds = (chosen_logps - rejected_logps) - (
            torch.log1p(-torch.exp(chosen_logps)) - torch.log1p(-torch.exp(rejected_logps)))
        ratio = F.logsigmoid(log_odds)
        or_loss = -(beta * ratio).mean()
        # token-mean CE over the whole chosen batch, from the per-sequence means
        counts = chosen_labels.ne(ii).sum(-1).float()
        ce_loss = -(chosen_logps * counts).sum() / counts.sum().clamp_min(1)
        chosen_rewards, rejected_rewards = beta * chosen_logps.detach(), beta * rejected_logps.detach()
        metrics['OR Loss'] = or_loss
        metrics['Chosen Rewards'] = chosen_rewards.mean()
        metrics['Rejected Rewards'] = rejected_rewards.mean()
        metrics['Reward Accuracy'] = (chosen_rewards > rejected_rewards).float().mean()
        metrics['Reward Margin'] = (chosen_rewards - rejected_rewards).mean()
        metrics['Chosen Log P'] = chosen_logps.detach().mean()
        metrics['Rejected Log P'] = rejected_logps.detach().mean()
        metrics['Chosen Logits'] = chosen_mean_logit
        metrics['Rejected Logits'] = rejected_mean_logit
        metrics['Log Odds Ratio'] = ratio.detach().mean()
        metrics['Log Odds Chosen'] = log_odds.detach().mean()
        metrics['CE Loss'] = ce_loss
        return or_loss + ce_loss

    def _step(self, batch: dict[str, torch.Tensor | Any]) -> tuple[torch.Tensor, dict[str, torch.Tensor]]:
        metrics: dict[str, torch.Tensor] = {}
        loss = self.compute_loss(**self.forward_batch(batch), metrics=metrics)
        metrics['Loss'] = loss
        return loss, metrics

    def training_step(self, batch: dict[str, torch.Tensor | Any], batch_idx: int) -> torch.Tensor:
        longest = max(batch['chosen_input_ids'].size(1), batch['rejected_input_ids'].size(1))
        if self.config.empty_cache_threshold is not None and longest >= self.config.empty_cache_threshold:
            gc.collect()
            if torch.cuda.is_available():
                torch.cuda.empty_cache()
        loss, metrics = self._step(batch)
        self.log('loss', loss, prog_bar=True, logger=False)
        self.log_dict({k + '/Train/Step': v for k, v in metrics.items()})
        return loss

    def validation_step(self, batch: dict[str, torch.Tensor | Any], batch_idx: int, dataloader_idx: int = 0) -> None:
        _, metrics = self._step(batch)
        self.log_dict({k + '/Val': v for k, v in metrics.items()}, sync_dist=True,
                      batch_size=batch['chosen_input_ids'].size(0))
