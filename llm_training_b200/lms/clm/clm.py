"""Causal-LM objective for pre-training and instruction tuning
(reference: src/llm_training/lms/clm/clm.py:25-197).

Loss path on CUDA/bf16: the model is called with ``skip_logits=True`` and the final-norm hidden states go
straight into ``fused_linear_cross_entropy`` (tcgen05 GEMM chunks + in-place CE kernel), so the fp32
``[B, T, V]`` logits of the reference (clm.py:147) never exist; under TP the vocab-parallel variant reduces
max / sum-exp / target logit across ranks.  NEFTune noise is fused into the embedding kernel instead of a
forward hook; consumed-sample / token counters stay on device and are reduced only when logged
(reference all-gathers them every step, clm.py:162-167).
"""
import logging
from typing import Any

import torch

from ...metrics import ConsumedSamples, ConsumedTokens, Perplexity
from ...models.base_model import BaseModel
from ...ops import cross_entropy, fused_linear_cross_entropy, shift_labels
from ..base_lm import BaseLightningModule
from ..utils import get_model
from .clm_config import CLMConfig

logger = logging.getLogger(__name__)


class CLM(BaseLightningModule):
    config: CLMConfig
    # one forward/backward per weight per step: wgrad GEMMs may accumulate directly into the flat grad buffers
    weights_used_once_per_step = True

    def __init__(self, config: CLMConfig) -> None:
        super().__init__(config)
        self.model: BaseModel | None = None

    @property
    def has_pre_trained_weights(self) -> bool:
        return self.model is not None and self.model.has_pre_trained_weights

    def get_pre_trained_weights(self) -> dict[str, torch.Tensor]:
        return {f'model.{k}': v for k, v in self.model.get_pre_trained_weights().items()}

    def configure_model(self) -> None:
        group = None
        if self.strategy is not None and self.strategy.dp_size > 1:
            group = self.strategy.dp_mesh.get_group()
        self.consumed_samples = ConsumedSamples(process_group=group)
        self.consumed_tokens = ConsumedTokens(ignore_index=self.config.ignore_index, process_group=group)
        if self.config.log_perplexity:
            self.train_perplexity = Perplexity(ignore_index=self.config.ignore_index, process_group=group)
            self.val_perplexity = Perplexity(ignore_index=self.config.ignore_index, process_group=group)
        self.model = get_model(self.config.model)
        if self.global_rank == 0:
            logger.info(f'Config:\n{self.config!r}')
            logger.info(f'Model:\n{self.model}')
        if self.config.neftune_alpha is not None:
            self.register_neftune()

    def register_neftune(self) -> None:
        if hasattr(self.model, 'neftune_alpha'):
            self.model.neftune_alpha = self.config.neftune_alpha     # fused into the embedding kernel
        else:                                                        # generic models: forward hook
            self.model.get_input_embeddings().register_forward_hook(self._neftune_hook)

    # names of the reference's hooks (lms/clm/clm.py:45-82)
    def register_neftune_hook(self) -> None:
        self.register_neftune()

    def neftune_forward_hook(self, module, input, output):     # noqa: A002 -- reference signature
        return self._neftune_hook(module, input, output)

    def _neftune_hook(self, module, inputs, output):
        if not module.training:
            return output
        mask = getattr(self, '_current_attention_mask', None)
        mask = torch.ones(output.shape[:2], device=output.device) if mask is None else mask.bool().to(output.dtype)
        noise = torch.empty_like(output).uniform_(-1, 1)
        scale = self.config.neftune_alpha / torch.sqrt(mask.sum(1) * output.size(-1))
        return output + (noise * mask.unsqueeze(2) * scale.view(-1, 1, 1)).detach()

    def on_fsdp_parallelize_model(self, **kwargs: Any) -> None:
        self.model.parallelize(**kwargs)

    def get_model(self) -> BaseModel:
        return self.model

    # ---------------------------------------------------------------- loss
    def compute_loss(self, logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        return cross_entropy(logits=logits, labels=labels, ignore_index=self.config.ignore_index)

    def _forward_loss(self, batch: dict[str, torch.Tensor | Any]) -> tuple[torch.Tensor, torch.Tensor]:
        labels = shift_labels(batch['labels'], self.config.ignore_index)
        self._current_attention_mask = batch.get('attention_mask')
        fused = self.config.fused_loss and getattr(self.model, 'supports_fused_loss', False)
        kwargs = {'skip_logits': True} if fused else {}
        outputs = self.model(input_ids=batch['input_ids'], attention_mask=batch.get('attention_mask'),
                             position_ids=batch.get('position_ids'), **kwargs)
        self._current_attention_mask = None
        if fused:
            tp = getattr(self.model, 'tp_context', None)
            if tp is not None:
                loss = tp.vocab_parallel_cross_entropy(outputs.normed_hidden_states,
                                                       self.model.get_output_embeddings().weight,
                                                       labels.reshape(-1), self.config.ignore_index)
            else:
                loss = fused_linear_cross_entropy(outputs.normed_hidden_states,
                                                  self.model.get_output_embeddings().weight, labels.reshape(-1),
                                                  self.config.ignore_index)
        else:
            loss = self.compute_loss(outputs.logits.float(), labels)
        return self._normalize_over_data_parallel_ranks(loss, labels), labels

    def _normalize_over_data_parallel_ranks(self, loss: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """``loss_normalization='global_tokens'``: turn the rank-local token mean into this rank's share of the mean
        over ALL valid tokens of the data-parallel step, so that the gradient average over ranks equals the gradient
        of the global token mean (ranks with few valid tokens no longer count as much as full ones)."""
        strategy = self.strategy
        if self.config.loss_normalization != 'global_tokens' or strategy is None or strategy.dp_size <= 1 \
                or not torch.is_grad_enabled():
            return loss
        import torch.distributed as dist
        n_local = labels.ne(self.config.ignore_index).sum().to(torch.float32)
        n_global = n_local.clone()
        dist.all_reduce(n_global, group=strategy.dp_mesh.get_group())
        return loss * (n_local * strategy.dp_size / n_global.clamp_min(1.0))

    def training_step(self, batch: dict[str, torch.Tensor | Any], batch_idx: int) -> torch.Tensor:
        loss, labels = self._forward_loss(batch)
        if self.config.neftune_alpha is not None:
            self.log('NEFTune Alpha', self.config.neftune_alpha)
        self.log('loss', loss, prog_bar=True, logger=False)
        self.log('Loss/Train/Step', loss)
        if self.config.log_perplexity:
            self.log('Perplexity/Train/Step', self.train_perplexity(loss))
        self.consumed_samples.update(labels)
        self.consumed_tokens.update(labels)
        self.log('Consumed Samples', self.consumed_samples)     # reduced lazily, only when a logger flushes
        self.log('Consumed Tokens', self.consumed_tokens)
        return loss

    def validation_step(self, batch: dict[str, torch.Tensor | Any], batch_idx: int, dataloader_idx: int = 0):
        batch_size = batch['input_ids'].size(0)
        loss, _ = self._forward_loss(batch)
        self.log('Loss/Val', loss, batch_size=batch_size, sync_dist=True)
        if self.config.log_perplexity:
            self.val_perplexity.update(loss)
            self.log('Perplexity/Val', self.val_perplexity)
