from .modules import (FlaxLLaMAAttention, FlaxLLaMABlock, FlaxLLaMABlockCollection, FlaxLLaMAForCausalLMModule,  # noqa: E402,F401
                      FlaxLLaMAMLP, FlaxLLaMAModule, LLaMAAttention, LLaMABlock, LLaMABlockCollection,
                      LLaMAForCausalLMModule, LLaMAMLP, LLaMAModule, RMSNorm)
