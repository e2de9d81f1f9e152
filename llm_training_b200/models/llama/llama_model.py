"""LLaMA 2 / 3 / 3.x decoder (reference: src/llm_training/models/llama/llama_model.py:32-821).

Same module tree and parameter names as the reference (and therefore as Hugging Face after the
``model.`` prefix rename), so checkpoints and ``convert_to_hf.py`` are interchangeable:

    embed_tokens.weight, layers.{i}.self_attn.{q,k,v,o}_proj.weight, layers.{i}.mlp.{gate,up,down}_proj.weight,
    layers.{i}.{input,post_attention}_layernorm.weight, norm.weight, lm_head.weight

The compute path is new: activations are packed ``[T, hidden]``; q|k|v and gate|up are each ONE tcgen05
GEMM over adjacent weight views; RoPE runs in place on the fused projection output; attention is the
packed varlen kernel driven by per-token document bounds; residual adds are fused into the o_proj /
down_proj GEMM epilogues; the LM head is normally not applied here at all (``skip_logits``) because
the loss kernels consume the final-norm hidden states directly.
"""
import logging
from typing import Any

import torch
import torch.utils.checkpoint
from torch import nn

from ...ops import (embedding, linear, multi_linear, neftune_noise_scale, norm_multi_linear, silu_mul_fused,
                    swiglu_down_linear)
from ..decoder import (PackedContext, RMSNorm, RotaryEmbedding, attention_core, build_packed_context,
                       dense_core_attention, repeat_kv)  # noqa: F401  (repeat_kv: public name of the reference's model modules)
from ..hf_compat_model import HFCompatModel
from ..utils.modeling_outputs import CausalLMOutput
from .llama_config import LlamaConfig

logger = logging.getLogger(__name__)


class LlamaRMSNorm(RMSNorm):
    def __init__(self, config: LlamaConfig) -> None:
        super().__init__(config.hidden_size, config.rms_norm_eps)


class LlamaRotaryEmbedding(RotaryEmbedding):
    def __init__(self, config: LlamaConfig) -> None:
        head_dim = config.head_dim or config.hidden_size // config.num_attention_heads
        super().__init__(head_dim, config.rope_theta, config.max_position_embeddings, config.rope_scaling)


class LlamaMLP(nn.Module):
    def __init__(self, config: LlamaConfig) -> None:
        super().__init__()
        self.hidden_size = config.hidden_size
        self.intermediate_size = config.intermediate_size
        self.gate_proj = nn.Linear(self.hidden_size, self.intermediate_size, bias=config.mlp_bias)
        self.up_proj = nn.Linear(self.hidden_size, self.intermediate_size, bias=config.mlp_bias)
        self.down_proj = nn.Linear(self.intermediate_size, self.hidden_size, bias=config.mlp_bias)

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None, tp: Any = None,
                pre_norm: nn.Module | None = None) -> torch.Tensor:
        """``pre_norm``: the RMSNorm to apply to ``x`` first, fused with the gate|up projection."""
        no_bias = self.gate_proj.bias is None and self.down_proj.bias is None
        if pre_norm is not None:
            gate_up, x_pass = norm_multi_linear(x, pre_norm.weight, pre_norm.variance_epsilon,
                                                [self.gate_proj.weight, self.up_proj.weight], return_residual=True)
            if residual is x:
                residual = x_pass       # both gradient branches of x meet in the fused block's backward
        elif tp is not None:
            gate_up = tp.column_linear(x, [self.gate_proj.weight, self.up_proj.weight],
                                       [self.gate_proj.bias, self.up_proj.bias])
        else:
            gate_up = multi_linear(x, [self.gate_proj.weight, self.up_proj.weight],
                                   [self.gate_proj.bias, self.up_proj.bias])
        if tp is None and no_bias:
            return swiglu_down_linear(gate_up, self.down_proj.weight, residual)
        act = silu_mul_fused(gate_up)
        if tp is not None:
            return tp.row_linear(act, self.down_proj.weight, self.down_proj.bias, residual)
        return linear(act, self.down_proj.weight, self.down_proj.bias, residual=residual)


class LlamaAttention(nn.Module):
    def __init__(self, config: LlamaConfig, layer_idx: int | None = None) -> None:
        super().__init__()
        self.config = config
        self.layer_idx = layer_idx
        self.hidden_size = config.hidden_size
        self.num_heads = config.num_attention_heads
        self.head_dim = config.head_dim or self.hidden_size // self.num_heads
        self.num_key_value_heads = config.num_key_value_heads
        self.num_key_value_groups = self.num_heads // self.num_key_value_heads
        self.attention_dropout = config.attention_dropout
        self.sliding_window: int | None = None
        if config.head_dim is None and self.head_dim * self.num_heads != self.hidden_size:
            raise ValueError(f'hidden_size {self.hidden_size} is not divisible by num_heads {self.num_heads}')
        bias = config.attention_bias
        self.q_proj = nn.Linear(self.hidden_size, self.num_heads * self.head_dim, bias=bias)
        self.k_proj = nn.Linear(self.hidden_size, self.num_key_value_heads * self.head_dim, bias=bias)
        self.v_proj = nn.Linear(self.hidden_size, self.num_key_value_heads * self.head_dim, bias=bias)
        self.o_proj = nn.Linear(self.num_heads * self.head_dim, self.hidden_size, bias=bias)

    def core_attention_forward(self, query_states: torch.Tensor, key_states: torch.Tensor,
                               value_states: torch.Tensor, attention_mask: torch.Tensor | None) -> torch.Tensor:
        """Dense attention on ``[B, H, S, D]`` tensors with a 4-D additive mask (the reference's method of the same
        name); ``forward`` uses the packed kernels instead."""
        impl = 'sdpa' if self.config._attn_implementation == 'sdpa' else 'eager'
        return dense_core_attention(query_states, key_states, value_states, attention_mask,
                                    getattr(self, 'attention_dropout', 0.0), self.training, impl)

    def forward(self, hidden_states: torch.Tensor, ctx: PackedContext, residual: torch.Tensor | None = None,
                pre_norm: nn.Module | None = None) -> torch.Tensor:
        tp = ctx.tp
        shards = 1 if tp is None else tp.size
        if pre_norm is not None:
            qkv, x_pass = norm_multi_linear(hidden_states, pre_norm.weight, pre_norm.variance_epsilon,
                                            [self.q_proj.weight, self.k_proj.weight, self.v_proj.weight],
                                            return_residual=True)
            if residual is hidden_states:
                residual = x_pass       # both gradient branches of the layer input meet in the fused block's backward
        elif tp is not None:
            qkv = tp.column_linear(hidden_states, [self.q_proj.weight, self.k_proj.weight, self.v_proj.weight],
                                   [self.q_proj.bias, self.k_proj.bias, self.v_proj.bias])
        else:
            qkv = multi_linear(hidden_states, [self.q_proj.weight, self.k_proj.weight, self.v_proj.weight],
                               [self.q_proj.bias, self.k_proj.bias, self.v_proj.bias])
        attn = attention_core(
            qkv, ctx, num_heads=self.num_heads // shards, num_kv_heads=self.num_key_value_heads // shards,
            head_dim=self.head_dim, impl=self.config._attn_implementation, window=self.sliding_window,
            dropout_p=self.attention_dropout, training=self.training,
            selective_recompute=(self.config.enable_gradient_checkpointing
                                 and self.config.recompute_granularity == 'selective'))
        if tp is not None:
            return tp.row_linear(attn, self.o_proj.weight, self.o_proj.bias, residual)
        return linear(attn, self.o_proj.weight, self.o_proj.bias, residual=residual)


# names kept for config / checkpoint tooling that refers to the reference's attention classes
LlamaFlashAttention2 = LlamaAttention
LlamaSdpaAttention = LlamaAttention


class LlamaDecoderLayer(nn.Module):
    def __init__(self, config: LlamaConfig, layer_idx: int) -> None:
        super().__init__()
        self.config = config
        self.hidden_size = config.hidden_size
        self.self_attn = LlamaAttention(config, layer_idx)
        self.mlp = LlamaMLP(config)
        self.input_layernorm = LlamaRMSNorm(config)
        self.post_attention_layernorm = LlamaRMSNorm(config)

    def forward(self, hidden_states: torch.Tensor, ctx: PackedContext) -> torch.Tensor:
        tp = ctx.tp
        if tp is None and not self.config.attention_bias and not self.config.mlp_bias:
            # norm + projection fused (normalised activations are recomputed in backward, not stored)
            hidden_states = self.self_attn(hidden_states, ctx, residual=hidden_states, pre_norm=self.input_layernorm)
            return self.mlp(hidden_states, residual=hidden_states, pre_norm=self.post_attention_layernorm)
        # tensor parallel: the norms run on [T/tp, d] token shards (sequence parallel); the column-parallel
        # projections gather them (fused all-gather -> GEMM on NVLink), the row-parallel ones reduce-scatter back
        normed = self.input_layernorm(hidden_states)
        hidden_states = self.self_attn(normed, ctx, residual=hidden_states)
        normed = self.post_attention_layernorm(hidden_states)
        return self.mlp(normed, residual=hidden_states, tp=tp)


class Llama(HFCompatModel):
    config: LlamaConfig
    config_class = LlamaConfig
    no_split_modules = ['LlamaDecoderLayer']
    supports_fused_loss = True

    def __init__(self, config: LlamaConfig) -> None:
        super().__init__(config)
        config = self.config
        self.padding_idx = config.pad_token_id
        self.vocab_size = config.vocab_size
        self.embed_tokens = nn.Embedding(config.vocab_size, config.hidden_size, self.padding_idx)
        self.layers = nn.ModuleList([LlamaDecoderLayer(config, i) for i in range(config.num_hidden_layers)])
        self.norm = LlamaRMSNorm(config)
        self.rotary_emb = LlamaRotaryEmbedding(config)
        self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, bias=False)
        if config.tie_word_embeddings:
            self.lm_head.weight = self.embed_tokens.weight
        self.tp_context: Any = None            # set by parallel.tensor_parallel
        self.neftune_alpha: float | None = None  # set by CLM (fused into the embedding kernel)
        self._neftune_seed = 0

    # ---------------------------------------------------------------- HF interop
    @classmethod
    def _hf_config_class(cls):
        from transformers import LlamaConfig as HFLlamaConfig
        return HFLlamaConfig

    @classmethod
    def _hf_model_class(cls):
        from transformers import LlamaForCausalLM
        return LlamaForCausalLM

    _MERGED_FIELDS = ('tie_word_embeddings', 'vocab_size', 'hidden_size', 'intermediate_size', 'num_hidden_layers',
                      'num_attention_heads', 'num_key_value_heads', 'max_position_embeddings', 'initializer_range',
                      'rms_norm_eps', 'pad_token_id', 'bos_token_id', 'eos_token_id', 'attention_dropout',
                      'attention_bias', 'mlp_bias')

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
            scaling = {k: v for k, v in rope_params.items() if k != 'rope_theta'}
        if scaling is not None and self.config.rope_scaling is None:
            self.config.rope_scaling = dict(scaling)
        if getattr(hf_config, 'head_dim', None):
            self.config.head_dim = hf_config.head_dim

    def build_hf_config(self):
        c = self.config
        kwargs = {name: getattr(c, name) for name in self._MERGED_FIELDS}
        return self._hf_config_class()(rope_theta=c.rope_theta, rope_scaling=c.rope_scaling, hidden_act='silu',
                                       **kwargs)

    def convert_state_dict_from_hf(self, hf_state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return {k.removeprefix('model.'): v for k, v in hf_state_dict.items()}

    def convert_state_dict_to_hf(self, state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return {('' if k.startswith('lm_head') else 'model.') + k: v for k, v in state_dict.items()}

    # ---------------------------------------------------------------- init / accessors
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
        """Same grouping as the reference's ``fully_shard`` calls (llama_model.py:246-268)."""
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
        apply_decoder_tensor_parallel(self, tp_mesh, fused_qkv=False, fused_gate_up=False)

    def configure_fully_sharded_data_parallel(self, dp_mesh: Any, **kwargs: Any) -> None:
        if dp_mesh is None or dp_mesh.size() == 1:
            return
        from ...parallel.fsdp import apply_fully_sharded
        apply_fully_sharded(self, dp_mesh, **kwargs)

    # ---------------------------------------------------------------- forward
    def _embed(self, input_ids: torch.Tensor, attention_mask: torch.Tensor | None) -> torch.Tensor:
        b, s = input_ids.shape
        w = self.embed_tokens.weight
        noise = None
        if self.neftune_alpha is not None and self.training:
            noise = neftune_noise_scale(attention_mask, b, s, w.shape[1], self.neftune_alpha, w.device)
            self._neftune_seed += 1
        if self.tp_context is not None:
            return self.tp_context.vocab_parallel_embedding(input_ids, w, self.padding_idx, noise, self._neftune_seed)
        return embedding(input_ids.reshape(-1), w, self.padding_idx, noise, self._neftune_seed)

    def forward_decoder_layer(self, layer: LlamaDecoderLayer, hidden_states: torch.Tensor, ctx: PackedContext
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
            hidden_states = self._embed(input_ids, attention_mask)
        else:
            batch_size, seq_len = inputs_embeds.shape[:2]
            hidden_states = inputs_embeds.reshape(batch_size * seq_len, -1)
            if self.tp_context is not None:
                hidden_states = self.tp_context.scatter_sequence(hidden_states)
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
