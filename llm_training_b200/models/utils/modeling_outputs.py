"""Model output container (reference: src/llm_training/models/utils/modeling_outputs.py:5-14)."""
from dataclasses import dataclass

import torch


@dataclass
class CausalLMOutput:
    logits: torch.Tensor | None = None
    last_hidden_states: torch.Tensor | None = None
    # extension: final-norm output, handed to the fused linear + loss kernels instead of logits
    normed_hidden_states: torch.Tensor | None = None
