from .modeling_outputs import CausalLMOutput, ModelOutput
from .utils import init_empty_weights, init_on_device
