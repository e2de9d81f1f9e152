"""Any Hugging Face ``AutoModelForCausalLM`` as a model (reference:
src/llm_training/models/hf_causal_lm/hf_causal_lm.py:19-131).  Best-effort path: it runs the HF
module's own PyTorch code (no in-tree kernels, no tensor parallel); sharding wraps every
``_no_split_modules`` block plus the leftovers, like the reference (:88-114)."""
from typing import Any

import torch
from torch import nn

from ..hf_compat_model import HFCompatModel
from ..utils.modeling_outputs import CausalLMOutput
from .hf_causal_lm_config import HFCausalLMConfig


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
        if attention_mask is not None and attention_mask.dim() == 2:
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
