"""Model output container (reference: src/llm_training/models/utils/modeling_outputs.py:5-14)."""
from dataclasses import dataclass

import torch


@dataclass
class ModelOutput:
    """Base of the output containers (the reference's is a pydantic model; plain dataclasses here so outputs are
    cheap to build inside the step and keep working under ``torch.utils.checkpoint``)."""


@dataclass
class CausalLMOutput(ModelOutput):
    logits: torch.Tensor | None = None
    last_hidden_states: torch.Tensor | None = None
    # extension: final-norm output, handed to the fused linear + loss kernels instead of logits
    normed_hidden_states: torch.Tensor | None = None
