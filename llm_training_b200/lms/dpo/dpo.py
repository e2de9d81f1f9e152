"""Direct Preference Optimisation (reference: src/llm_training/lms/dpo/dpo.py:30-248).

Per-sequence log-probs come from ``fused_linear_logprob`` (no ``[B, T, V]`` log-softmax tensor; the
reference materialises it four times per step, dpo.py:110-114).  The frozen reference model runs under
``no_grad`` and is fed the SHIFTED labels like the policy (the reference passes unshifted labels to it,
dpo.py:139,146 — SURVEY Appendix B3; deliberate fix).
"""
import logging
from typing import Any

import torch
import torch.nn.functional as F

from ...models.base_model import BaseModel
from ...ops import fused_linear_logprob, shift_labels
from ..base_lm import BaseLightningModule
from ..utils import get_model
from .dpo_config import DPOConfig

logger = logging.getLogger(__name__)


def sequence_logps(model: BaseModel, input_ids: torch.Tensor, attention_mask: torch.Tensor | None,
                   position_ids: torch.Tensor | None, labels: torch.Tensor, ignore_index: int,
                   average: bool = False) -> tuple[torch.Tensor, Any]:
    """sum (or mean) over supervised positions of log p(label) -> [B].  ``labels`` are already shifted."""
    fused = getattr(model, 'supports_fused_loss', False)
    out = model(input_ids=input_ids, attention_mask=attention_mask, position_ids=position_ids,
                **({'skip_logits': True} if fused else {}))
    if fused:
        tp = getattr(model, 'tp_context', None)
        weight = model.get_output_embeddings().weight
        if tp is not None:
            per_token = tp.vocab_parallel_logprob(out.normed_hidden_states, weight, labels.reshape(-1), ignore_index)
        else:
            per_token = fused_linear_logprob(out.normed_hidden_states, weight, labels.reshape(-1), ignore_index)
        per_token = per_token.view(labels.shape)
    else:
        mask = labels.ne(ignore_index)
        logp = out.logits.float().log_softmax(-1)
        per_token = logp.gather(2, labels.masked_fill(~mask, 0).unsqueeze(2)).squeeze(2) * mask
    total = per_token.sum(-1)
    if average:
        total = total / labels.ne(ignore_index).sum(-1).clamp_min(1)
    return total, out


class DPO(BaseLightningModule):
    config: DPOConfig

    def __init__(self, config: DPOConfig) -> None:
        super().__init__(config)
        self.model: BaseModel | None = None
        self.ref_model: BaseModel | None = None

    @property
    def has_pre_trained_weights(self) -> bool:
        return self.model is not None and self.model.has_pre_trained_weights

    def get_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        state = {f'model.{k}': v for k, v in self.model.get_pre_trained_weights().items()}
        if self.config.ref_model is None:
            state |= {'ref_' + k: v for k, v in state.items()}
        else:
            state |= {f'ref_model.{k}': v for k, v in self.ref_model.get_pre_trained_weights().items()}
        return state

    def configure_model(self) -> None:
        self.model = get_model(self.config.model)
        self.ref_model = get_model(self.config.ref_model if self.config.ref_model is not None else self.config.model)
        self.ref_model.eval().requires_grad_(False)
        if self.global_rank == 0:
            logger.info(f'Config:\n{self.config!r}')

    def on_after_configure_model(self) -> None:
        super().on_after_configure_model()
        if not self.should_load_pre_trained_weights and self.config.ref_model is None:
            # no checkpoint to read: the reference policy starts as an exact copy of the policy
            with torch.no_grad():
                for r, p in zip(self.ref_model.parameters(), self.model.parameters()):
                    r.copy_(p)

    def on_fsdp_parallelize_model(self, **kwargs: Any) -> None:
        self.model.parallelize(**kwargs)
        self.ref_model.parallelize(**kwargs)

    def get_model(self) -> BaseModel:
        return self.model

    def forward_batch(self, batch: dict[str, torch.Tensor | Any]) -> dict[str, torch.Tensor]:
        ii = self.config.ignore_index
        out: dict[str, torch.Tensor] = {}
        for side in ('chosen', 'rejected'):
            labels = shift_labels(batch[f'{side}_labels'], ii)
            args = (batch[f'{side}_input_ids'], batch.get(f'{side}_attention_mask'),
                    batch.get(f'{side}_position_ids'), labels, ii)
            out[f'policy_{side}_logps'], _ = sequence_logps(self.model, *args)
            with torch.no_grad():
                out[f'reference_{side}_logps'], _ = sequence_logps(self.ref_model, *args)
        return out

    def get_logps(self, model: BaseModel, input_ids: torch.Tensor, attention_mask: torch.Tensor | None,
                  position_ids: torch.Tensor | None, labels: torch.Tensor) -> torch.Tensor:
        """Per-sequence SUM of label log-probabilities (``labels`` already shifted), same contract as the
        reference (lms/dpo/dpo.py:73-114) — computed by the fused linear + log-prob path, so no ``[B, T, V]``
        log-softmax is materialised (vocab-parallel under TP)."""
        return sequence_logps(model, input_ids, attention_mask, position_ids, labels, self.config.ignore_index)[0]

    def add_suffix_to_metrics(self, metrics: dict[str, torch.Tensor], suffix: str) -> dict[str, torch.Tensor]:
        return {k + suffix: v for k, v in metrics.items()}

    def compute_loss(self, policy_chosen_logps: torch.Tensor, policy_rejected_logps: torch.Tensor,
                     reference_chosen_logps: torch.Tensor, reference_rejected_logps: torch.Tensor,
                     metrics: dict[str, torch.Tensor]) -> torch.Tensor:
        beta, eps = self.config.beta, self.config.label_smoothing
        logits = (policy_chosen_logps - policy_rejected_logps) - (reference_chosen_logps - reference_rejected_logps)
        loss = (-F.logsigmoid(beta * logits) * (1 - eps) - F.logsigmoid(-beta * logits) * eps).mean()
        chosen_rewards = beta * (policy_chosen_logps - reference_chosen_logps).detach()
        rejected_rewards = beta * (policy_rejected_logps - reference_rejected_logps).detach()
        metrics['Chosen Reward'] = chosen_rewards.mean()
        metrics['Rejected Reward'] = rejected_rewards.mean()
        metrics['Reward Accuracy'] = (chosen_rewards > rejected_rewards).float().mean()
        metrics['Reward Margin'] = (chosen_rewards - rejected_rewards).mean()
        metrics['Chosen Log P'] = policy_chosen_logps.detach().mean()
        metrics['Rejected Log P'] = policy_rejected_logps.detach().mean()
        return loss

    def _step(self, batch: dict[str, torch.Tensor | Any]) -> tuple[torch.Tensor, dict[str, torch.Tensor]]:
        metrics: dict[str, torch.Tensor] = {}
        loss = self.compute_loss(**self.forward_batch(batch), metrics=metrics)
        metrics['Loss'] = loss
        return loss, metrics

    def training_step(self, batch: dict[str, torch.Tensor | Any], batch_idx: int) -> torch.Tensor:
        loss, metrics = self._step(batch)
        self.log('loss', loss, prog_bar=True, logger=False)
        self.log_dict({k + '/Train/Step': v for k, v in metrics.items()})
        return loss

    def validation_step(self, batch: dict[str, torch.Tensor | Any], batch_idx: int, dataloader_idx: int = 0) -> None:
        _, metrics = self._step(batch)
        self.log_dict({k + '/Val': v for k, v in metrics.items()}, sync_dist=True,
                      batch_size=batch['chosen_input_ids'].size(0))
