from .clm_proto import CausalLMProto
