"""Structural type every causal LM must satisfy (reference: src/llm_training/lms/protos/clm_proto.py:9-26)."""
from typing import Protocol, runtime_checkable

import torch
from torch import nn

from ...models.utils.modeling_outputs import CausalLMOutput


@runtime_checkable
class CausalLMProto(Protocol):
    def __call__(self, input_ids: torch.Tensor | None = None, attention_mask: torch.Tensor | None = None,
                 position_ids: torch.Tensor | None = None, inputs_embeds: torch.Tensor | None = None,
                 return_last_hidden_states: bool = False) -> CausalLMOutput: ...

    def get_input_embeddings(self) -> nn.Embedding: ...

    def get_output_embeddings(self) -> nn.Linear: ...

    def set_input_embeddings(self, embed_tokens: nn.Embedding) -> None: ...

    def set_output_embeddings(self, linear_module: nn.Linear) -> None: ...
