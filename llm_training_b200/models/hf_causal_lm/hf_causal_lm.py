"""Any Hugging Face ``AutoModelForCausalLM`` as a model (reference:
src/llm_training/models/hf_causal_lm/hf_causal_lm.py:19-131).  The HF module's own code runs the layers (no tensor
parallel, as in the reference); sharding wraps every ``_no_split_modules`` block plus the leftovers (:88-114).

Attention: with ``attn_implementation`` ``flash_attention_2`` (the auto-pick on a GPU) the HF attention layers call
the in-tree packed attention kernel through HF's ``AttentionInterface`` — O(T) memory, segment-id masks honoured as
per-token document bounds computed once per batch.  The reference gets there by monkey-patching HF's
``_get_unpad_data`` so flash-attn's varlen kernel sees the segment ids (:14-20); the dense ``[B,1,T,T]`` mask is only
built for ``sdpa`` / ``eager``."""
from typing import Any

import torch
from torch import nn

from ..hf_compat_model import HFCompatModel
from ..utils.modeling_outputs import CausalLMOutput
from .hf_causal_lm_config import HFCausalLMConfig


PACKED_ATTENTION = 'llmt_packed'


def _packed_attention_interface(module: nn.Module, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                attention_mask: torch.Tensor | None, dropout: float = 0.0,
                                scaling: float | None = None, is_causal: bool | None = None,
                                sliding_window: int | None = None, softcap: float | None = None,
                                **kwargs: Any) -> tuple[torch.Tensor, None]:
    """HF attention-interface function: q ``[B, Hq, T, D]``, k / v ``[B, Hkv, T, D]`` -> ``[B, T, Hq, D]``.
    The per-token document bounds were derived from the batch's segment ids by ``HFCausalLM.forward`` and are shared
    with every attention layer through ``module._llmt_packed_state``."""
    from ...ops.attention_op import packed_attention
    from ...ops.segments import segment_bounds
    if softcap is not None or (dropout and module.training):
        raise NotImplementedError('HFCausalLM packed attention: attention dropout / softcap are not supported; use '
                                  'attn_implementation="sdpa" for such a model')
    b, hq, t, d = query.shape
    hkv = key.shape[1]
    state = getattr(module, '_llmt_packed_state', None) or {}
    bounds = state.get('bounds')
    if bounds is None or bounds[0].numel() != b * t or bounds[0].device != query.device:
        bounds = segment_bounds(None, b, t, query.device)
    qkv = torch.cat((query.transpose(1, 2).reshape(b * t, hq * d), key.transpose(1, 2).reshape(b * t, hkv * d),
                     value.transpose(1, 2).reshape(b * t, hkv * d)), dim=-1)
    out = packed_attention(qkv, bounds[0], bounds[1], hq, hkv, d, scaling, sliding_window)
    return out.view(b, t, hq, d), None


def _register_packed_attention() -> None:
    from transformers import AttentionInterface
    from transformers.masking_utils import AttentionMaskInterface
    if PACKED_ATTENTION not in AttentionInterface._global_mapping:
        AttentionInterface.register(PACKED_ATTENTION, _packed_attention_interface)
        # no 4-D mask: documents, causality and the sliding window are all expressed by the bounds
        AttentionMaskInterface.register(PACKED_ATTENTION, lambda *args, **kwargs: None)


def patch_hf_modules_with_in_tree_kernels(model: nn.Module) -> int:
    """``enable_liger_kernel`` of the reference patches Hugging Face modules with Liger's Triton kernels
    (hf_causal_lm.py:62-86).  Here the same switch routes every RMSNorm and every SiLU-gated MLP of the wrapped model
    through the in-tree ops (sm_100a kernels on GPU, the fp32 restatements on CPU); modules that do not match the
    expected layout are left untouched.  Returns the number of patched modules."""
    import types

    from ...ops import linear, rms_norm, silu_mul

    def norm_forward(self, hidden_states):
        eps = getattr(self, 'variance_epsilon', None)
        return rms_norm(hidden_states, self.weight, eps if eps is not None else getattr(self, 'eps', 1e-6))

    def mlp_forward(self, x):
        gate = linear(x, self.gate_proj.weight, self.gate_proj.bias)
        up = linear(x, self.up_proj.weight, self.up_proj.bias)
        return linear(silu_mul(gate, up), self.down_proj.weight, self.down_proj.bias)

    patched = 0
    for module in model.modules():
        name = type(module).__name__
        if name.endswith('RMSNorm') and isinstance(getattr(module, 'weight', None), nn.Parameter) \
                and (hasattr(module, 'variance_epsilon') or hasattr(module, 'eps')) and 'Gemma' not in name:
            module.forward = types.MethodType(norm_forward, module)      # Gemma scales by (1 + w): not this formula
            patched += 1
        elif all(isinstance(getattr(module, n, None), nn.Linear) for n in ('gate_proj', 'up_proj', 'down_proj')) \
                and type(getattr(module, 'act_fn', None)).__name__ in ('SiLU', 'SiLUActivation'):
            module.forward = types.MethodType(mlp_forward, module)
            patched += 1
    return patched


class HFCausalLM(HFCompatModel):
    config: HFCausalLMConfig
    config_class = HFCausalLMConfig
    supports_fused_loss = False

    @classmethod
    def _hf_model_class(cls):
        from transformers import AutoModelForCausalLM
        return AutoModelForCausalLM

    def __init__(self, config: HFCausalLMConfig) -> None:
        super().__init__(config)
        if self.hf_config is None and self.config.hf_config is not None:
            self.hf_config = self.config.hf_config
        self.config.hf_config = self.hf_config
        self._packed_state: dict[str, Any] = {}
        self.uses_packed_attention = self.config._attn_implementation == 'flash_attention_2'
        if self.uses_packed_attention:
            _register_packed_attention()
            self.hf_model = self.construct_hf_model(attn_implementation=PACKED_ATTENTION)
            for module in self.hf_model.modules():
                if hasattr(module, 'q_proj') or hasattr(module, 'qkv_proj') or type(module).__name__.endswith('Attention'):
                    module._llmt_packed_state = self._packed_state
        else:
            self.hf_model = self.construct_hf_model()
        if self.config.enable_liger_kernel:
            self.patched_modules = patch_hf_modules_with_in_tree_kernels(self.hf_model)
        if self.config.enable_gradient_checkpointing:
            self.hf_model.gradient_checkpointing_enable({'use_reentrant': False})
        self.hf_model.tie_weights()

    def build_hf_config(self):
        raise ValueError('HFCausalLM needs `hf_path` or an explicit `hf_config`')

    def convert_state_dict_from_hf(self, hf_state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return {'hf_model.' + k: v for k, v in hf_state_dict.items()}

    def convert_state_dict_to_hf(self, state_dict: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        return {k.removeprefix('hf_model.'): v for k, v in state_dict.items()}

    def forward(self, input_ids: torch.Tensor | None = None, attention_mask: torch.Tensor | None = None,
                position_ids: torch.Tensor | None = None, inputs_embeds: torch.Tensor | None = None,
                return_last_hidden_states: bool = False, skip_logits: bool = False) -> CausalLMOutput:
        if self.uses_packed_attention:
            # segment ids -> per-token document bounds, once per batch, on device, no host sync; the attention layers
            # read them from the shared state (also during activation recompute in backward)
            from ...ops.segments import segment_bounds
            ref = input_ids if input_ids is not None else inputs_embeds
            bsz, seq = ref.shape[:2]
            self._packed_state['bounds'] = segment_bounds(
                attention_mask if (attention_mask is not None and attention_mask.dim() == 2) else None, bsz, seq,
                ref.device)
            attention_mask = None
        elif attention_mask is not None and attention_mask.dim() == 2:
            if bool((attention_mask > 1).any()):
                # packed batch (segment ids 1, 2, 3, ...): HF modules only know 0/1 masks, so hand them the
                # block-diagonal causal 4-D mask instead — documents of a pack stay isolated.  The reference gets the
                # same effect by monkey-patching HF's `_get_unpad_data` for flash-attn (hf_causal_lm.py:40-60).
                from ...ops import prepare_4d_causal_attention_mask
                ref = inputs_embeds if inputs_embeds is not None else self.get_input_embeddings().weight
                attention_mask = prepare_4d_causal_attention_mask(
                    attention_mask, tuple(attention_mask.shape), torch.empty(0, dtype=ref.dtype, device=ref.device), 0)
            else:
                attention_mask = attention_mask.clamp_max(1)
        outputs = self.hf_model(input_ids=input_ids, attention_mask=attention_mask, position_ids=position_ids,
                                inputs_embeds=inputs_embeds, output_hidden_states=return_last_hidden_states)
        last = outputs.hidden_states[-1] if return_last_hidden_states else None
        return CausalLMOutput(logits=outputs.logits, last_hidden_states=last)

    def get_input_embeddings(self) -> nn.Embedding:
        return self.hf_model.get_input_embeddings()

    def get_output_embeddings(self) -> nn.Linear:
        return self.hf_model.get_output_embeddings()

    def sharding_units(self) -> list[tuple[str, nn.Module]]:
        no_split = set(getattr(self.hf_model, '_no_split_modules', None) or [])
        units: list[tuple[str, nn.Module]] = []
        claimed: list[str] = []
        for name, module in self.named_modules():
            if type(module).__name__ in no_split:
                units.append((name, module))
                claimed.append(name)
        for name, module in self.named_modules():
            if any(c.startswith(name) or name.startswith(c) for c in claimed):
                continue
            if any(True for _ in module.parameters(recurse=False)) or any(True for _ in module.parameters()):
                units.append((name, module))
                claimed.append(name)
        return units

    def configure_fully_sharded_data_parallel(self, dp_mesh: Any, **kwargs: Any) -> None:
        if dp_mesh is None or dp_mesh.size() == 1:
            return
        from ...parallel.fsdp import apply_fully_sharded
        apply_fully_sharded(self, dp_mesh, **kwargs)
