"""Phi-3 / 3.5 / 4 decoder (reference: src/llm_training/models/phi3/phi3_model.py:31-903).

Differences from Llama that matter here: fused ``qkv_proj`` and ``gate_up_proj`` parameters (HF layout,
phi3_model.py:508-509, :421-422), LongRoPE, optional sliding window, dropout on embeddings / residual
branches.  The fused weights already are what the packed compute path wants, so q|k|v and gate|up are
single GEMMs with no view tricks.  KV heads are NOT repeated before attention (reference does,
phi3_model.py:641-642) — the kernel handles GQA natively.
"""
from typing import Any

import torch
import torch.nn.functional as F
import torch.utils.checkpoint
from torch import nn

from ...ops import embedding, linear, neftune_noise_scale, silu_mul_fused
from ..decoder import (PackedContext, RMSNorm, RotaryEmbedding, attention_core, build_packed_context,
                       dense_core_attention, repeat_kv)  # noqa: F401  (repeat_kv: public name of the reference's model modules)
from ..hf_compat_model import HFCompatModel
from ..utils.modeling_outputs import CausalLMOutput
from .phi3_config import Phi3Config


class Phi3RMSNorm(RMSNorm):
    def __init__(self, config: Phi3Config) -> None:
        super().__init__(config.hidden_size, config.rms_norm_eps)


class Phi3RotaryEmbedding(RotaryEmbedding):
    def __init__(self, config: Phi3Config) -> None:
        super().__init__(config.hidden_size // config.num_attention_heads, config.rope_theta,
                         config.max_position_embeddings, config.rope_scaling,
                         original_max_position_embeddings=config.original_max_position_embeddings)


class Phi3MLP(nn.Module):
    def __init__(self, config: Phi3Config) -> None:
        super().__init__()
        self.gate_up_proj = nn.Linear(config.hidden_size, 2 * config.intermediate_size, bias=False)
        self.down_proj = nn.Linear(config.intermediate_size, config.hidden_size, bias=False)

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None, tp: Any = None,
                resid_pdrop: float = 0.0) -> torch.Tensor:
        gate_up = linear(x, self.gate_up_proj.weight) if tp is None else tp.column_linear(x, [self.gate_up_proj.weight])
        act = silu_mul_fused(gate_up)
        if resid_pdrop > 0.0 and self.training:
            out = tp.row_linear(act, self.down_proj.weight, None, None) if tp is not None else \
                linear(act, self.down_proj.weight)
            return residual + F.dropout(out, resid_pdrop, True)
        if tp is not None:
            return tp.row_linear(act, self.down_proj.weight, None, residual)
        return linear(act, self.down_proj.weight, residual=residual)


def fused_core_attention_forward(query_states: torch.Tensor, key_states: torch.Tensor, value_states: torch.Tensor,
                                 attention_mask: torch.Tensor | None, dropout_p: float = 0.0,
                                 training: bool = False) -> torch.Tensor:
    """Name of the reference's ``torch.jit.script`` eager attention (phi3_model.py:436-480)."""
    return dense_core_attention(query_states, key_states, value_states, attention_mask, dropout_p, training)


class Phi3Attention(nn.Module):
    def __init__(self, config: Phi3Config, layer_idx: int | None = None) -> None:
        super().__init__()
        self.config = config
        self.layer_idx = layer_idx
        self.hidden_size = config.hidden_size
        self.num_heads = config.num_attention_heads
        self.head_dim = self.hidden_size // self.num_heads
        self.num_key_value_heads = config.num_key_value_heads or self.num_heads
        self.num_key_value_groups = self.num_heads // self.num_key_value_heads
        self.attention_dropout = config.attention_dropout
        self.sliding_window = config.sliding_window
        if self.head_dim * self.num_heads != self.hidden_size:
            raise ValueError(f'hidden_size {self.hidden_size} is not divisible by num_heads {self.num_heads}')
        op_size = (self.num_heads + 2 * self.num_key_value_heads) * self.head_dim
        self.qkv_proj = nn.Linear(self.hidden_size, op_size, bias=False)
        self.o_proj = nn.Linear(self.num_heads * self.head_dim, self.hidden_size, bias=False)

    def core_attention_forward(self, query_states: torch.Tensor, key_states: torch.Tensor,
                               value_states: torch.Tensor, attention_mask: torch.Tensor | None) -> torch.Tensor:
        """Dense attention on ``[B, H, S, D]`` tensors with a 4-D additive mask (the reference's method of the same
        name); ``forward`` uses the packed kernels instead."""
        impl = 'sdpa' if self.config._attn_implementation == 'sdpa' else 'eager'
        return dense_core_attention(query_states, key_states, value_states, attention_mask,
                                    getattr(self, 'attention_dropout', 0.0), self.training, impl)

    def forward(self, hidden_states: torch.Tensor, ctx: PackedContext, residual: torch.Tensor | None = None,
                resid_pdrop: float = 0.0) -> torch.Tensor:
        tp = ctx.tp
        shards = 1 if tp is None else tp.size
        qkv = linear(hidden_states, self.qkv_proj.weight) if tp is None \
            else tp.column_linear(hidden_states, [self.qkv_proj.weight])
        attn = attention_core(
            qkv, ctx, num_heads=self.num_heads // shards, num_kv_heads=self.num_key_value_heads // shards,
            head_dim=self.head_dim, impl=self.config._attn_implementation, window=self.sliding_window,
            dropout_p=self.attention_dropout, training=self.training,
            selective_recompute=(self.config.enable_gradient_checkpointing
                                 and self.config.recompute_granularity == 'selective'))
        if resid_pdrop > 0.0 and self.training:
            out = tp.row_linear(attn, self.o_proj.weight, None, None) if tp is not None else \
                linear(attn, self.o_proj.weight)
            return residual + F.dropout(out, resid_pdrop, True)
        if tp is not None:
            return tp.row_linear(attn, self.o_proj.weight, None, residual)
        return linear(attn, self.o_proj.weight, residual=residual)


Phi3FlashAttention2 = Phi3Attention
Phi3SdpaAttention = Phi3Attention


class Phi3DecoderLayer(nn.Module):
    def __init__(self, config: Phi3Config, layer_idx: int) -> None:
        super().__init__()
        self.config = config
        self.self_attn = Phi3Attention(config, layer_idx)
        self.mlp = Phi3MLP(config)
        self.input_layernorm = Phi3RMSNorm(config)
        self.post_attention_layernorm = Phi3RMSNorm(config)
        self.resid_pdrop = config.resid_pdrop

    def forward(self, hidden_states: torch.Tensor, ctx: PackedContext) -> torch.Tensor:
        tp = ctx.tp
        normed = self.input_layernorm(hidden_states)          # [T/tp, d] token shard under tensor parallelism
        hidden_states = self.self_attn(normed, ctx, residual=hidden_states, resid_pdrop=self.resid_pdrop)
        normed = self.post_attention_layernorm(hidden_states)
        return self.mlp(normed, residual=hidden_states, tp=tp, resid_pdrop=self.resid_pdrop)


class Phi3(HFCompatModel):
    config: Phi3Config
    config_class = Phi3Config
    no_split_modules = ['Phi3DecoderLayer']
    supports_fused_loss = True

    def __init__(self, config: Phi3Config) -> None:
        super().__init__(config)
        config = self.config
        self.padding_idx = config.pad_token_id
        self.vocab_size = config.vocab_size
        self.embed_tokens = nn.Embedding(config.vocab_size, config.hidden_size, self.padding_idx)
        self.embed_dropout = nn.Dropout(config.embd_pdrop)
        self.layers = nn.ModuleList([Phi3DecoderLayer(config, i) for i in range(config.num_hidden_layers)])
        self.norm = Phi3RMSNorm(config)
        self.rotary_emb = Phi3RotaryEmbedding(config)
        self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, bias=False)
        if config.tie_word_embeddings:
            self.lm_head.weight = self.embed_tokens.weight
        self.tp_context: Any = None
        self.neftune_alpha: float | None = None
        self._neftune_seed = 0

    @classmethod
    def _hf_config_class(cls):
        from transformers import Phi3Config as HFPhi3Config
        return HFPhi3Config

    @classmethod
    def _hf_model_class(cls):
        from transformers import Phi3ForCausalLM
        return Phi3ForCausalLM

    _MERGED_FIELDS = ('vocab_size', 'hidden_size', 'intermediate_size', 'num_hidden_layers', 'num_attention_heads',
                      'num_key_value_heads', 'resid_pdrop', 'embd_pdrop', 'attention_dropout',
                      'max_position_embeddings', 'original_max_position_embeddings', 'initializer_range',
                      'rms_norm_eps', 'pad_token_id', 'bos_token_id', 'eos_token_id', 'sliding_window',
                      'tie_word_embeddings')

    def merge_hf_config(self, hf_config: Any) -> None:
        assert getattr(hf_config, 'hidden_act', 'silu') == 'silu'
        for name in self._MERGED_FIELDS:
            if hasattr(hf_config, name):
                setattr(self.config, name, getattr(hf_config, name))
        rope_params = getattr(hf_config, 'rope_parameters', None) or {}
        theta = getattr(hf_config, 'rope_theta', None) or rope_params.get('rope_theta')
        if theta is not None:
            self.config.rope_theta = float(theta)
        scaling = getattr(hf_config, 'rope_scaling', None)
        if scaling is None and rope_params.get('rope_type', 'default') != 'default':
            scaling = {k: v for k, v in rope_params.items()
                       if k in ('rope_type', 'type', 'short_factor', 'long_factor', 'factor', 'attention_factor')}
        self.config.rope_scaling = Phi3Config.validate_rope_scaling.__func__(
            Phi3Config, scaling, type('I', (), {'data': {'hidden_size': self.config.hidden_size,
                                                         'num_attention_heads': self.config.num_attention_heads}})()
        ) if scaling is not None else None

    def build_hf_config(self):
        c = self.config
        kwargs = {name: getattr(c, name) for name in self._MERGED_FIELDS}
        return self._hf_config_class()(rope_theta=c.rope_theta, rope_scaling=c.rope_scaling, hidden_act='silu',
                                       **kwargs)

    def convert_state_dict_from_hf(self, hf_state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return {k.removeprefix('model.'): v for k, v in hf_state_dict.items()}

    def convert_state_dict_to_hf(self, state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return {('' if k.startswith('lm_head') else 'model.') + k: v for k, v in state_dict.items()}

    def _init_weights_impl(self, module: nn.Module) -> None:
        std = self.config.initializer_range
        if isinstance(module, nn.Linear):
            module.weight.data.normal_(mean=0.0, std=std)
            if module.bias is not None:
                module.bias.data.zero_()
        elif isinstance(module, nn.Embedding):
            module.weight.data.normal_(mean=0.0, std=std)
            if module.padding_idx is not None:
                module.weight.data[module.padding_idx].zero_()
        elif isinstance(module, RMSNorm):
            module.weight.data.fill_(1.0)

    def get_input_embeddings(self) -> nn.Embedding:
        return self.embed_tokens

    def get_output_embeddings(self) -> nn.Linear:
        return self.lm_head

    def set_input_embeddings(self, embedding_module: nn.Embedding) -> None:
        self.embed_tokens = embedding_module

    def set_output_embeddings(self, linear_module: nn.Linear) -> None:
        self.lm_head = linear_module

    def sharding_units(self) -> list[tuple[str, nn.Module]]:
        units: list[tuple[str, nn.Module]] = [('embed_tokens', self.embed_tokens)]
        units += [(f'layers.{i}', layer) for i, layer in enumerate(self.layers)]
        units += [('norm', self.norm)]
        if not self.config.tie_word_embeddings:
            units += [('lm_head', self.lm_head)]
        return units

    def configure_tensor_parallel(self, tp_mesh: Any) -> None:
        if tp_mesh is None or tp_mesh.size() == 1:
            return
        from ...parallel.tensor_parallel import apply_decoder_tensor_parallel
        # head-aligned sharding of the fused weights (fixes SURVEY Appendix B4)
        apply_decoder_tensor_parallel(self, tp_mesh, fused_qkv=True, fused_gate_up=True)

    def configure_fully_sharded_data_parallel(self, dp_mesh: Any, **kwargs: Any) -> None:
        if dp_mesh is None or dp_mesh.size() == 1:
            return
        from ...parallel.fsdp import apply_fully_sharded
        apply_fully_sharded(self, dp_mesh, **kwargs)

    def forward_decoder_layer(self, layer: Phi3DecoderLayer, hidden_states: torch.Tensor, ctx: PackedContext
                              ) -> torch.Tensor:
        if (self.config.enable_gradient_checkpointing and self.config.recompute_granularity == 'full'
                and self.training and torch.is_grad_enabled()):
            return torch.utils.checkpoint.checkpoint(layer, hidden_states, ctx, use_reentrant=False)
        return layer(hidden_states, ctx)

    def forward(self, input_ids: torch.Tensor | None = None, attention_mask: torch.Tensor | None = None,
                position_ids: torch.Tensor | None = None, inputs_embeds: torch.Tensor | None = None,
                return_last_hidden_states: bool = False, skip_logits: bool = False) -> CausalLMOutput:
        if (input_ids is None) == (inputs_embeds is None):
            raise ValueError('specify exactly one of input_ids and inputs_embeds')
        if input_ids is not None:
            batch_size, seq_len = input_ids.shape[:2]
            w = self.embed_tokens.weight
            noise = None
            if self.neftune_alpha is not None and self.training:
                noise = neftune_noise_scale(attention_mask, batch_size, seq_len, w.shape[1], self.neftune_alpha,
                                            w.device)
                self._neftune_seed += 1
            if self.tp_context is not None:
                hidden_states = self.tp_context.vocab_parallel_embedding(input_ids, w, self.padding_idx, noise,
                                                                         self._neftune_seed)
            else:
                hidden_states = embedding(input_ids.reshape(-1), w, self.padding_idx, noise, self._neftune_seed)
        else:
            batch_size, seq_len = inputs_embeds.shape[:2]
            hidden_states = inputs_embeds.reshape(batch_size * seq_len, -1)
            if self.tp_context is not None:
                hidden_states = self.tp_context.scatter_sequence(hidden_states)
        hidden_states = self.embed_dropout(hidden_states)
        ctx = build_packed_context(self.rotary_emb, batch_size, seq_len, hidden_states.device, attention_mask,
                                   position_ids, tp=self.tp_context)
        for layer in self.layers:
            hidden_states = self.forward_decoder_layer(layer, hidden_states, ctx)
        last_hidden = hidden_states.view(batch_size, -1, hidden_states.shape[-1]) if return_last_hidden_states else None
        normed = self.norm(hidden_states)
        if skip_logits:
            return CausalLMOutput(None, last_hidden, normed)
        if self.tp_context is not None:
            logits = self.tp_context.lm_head(normed, self.lm_head.weight)
        else:
            logits = linear(normed, self.lm_head.weight)
        return CausalLMOutput(logits.view(batch_size, seq_len, -1), last_hidden, None)
