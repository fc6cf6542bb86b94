"""Same-named module as the reference's `jax_llama/model.py`: re-exports the model classes, the module-style
wrappers and the functional building blocks from where they live in this package."""
from .models.llama import (FlaxLLaMAForCausalLM, FlaxLLaMAModel, LLaMAForCausalLM, LLaMAModel,  # noqa: F401
                           LLaMAPreTrainedModel)
from .models.modules import (FlaxLLaMAAttention, FlaxLLaMABlock, FlaxLLaMABlockCollection,  # noqa: F401
                             FlaxLLaMAForCausalLMModule, FlaxLLaMAMLP, FlaxLLaMAModule, LLaMAAttention, LLaMABlock,
                             LLaMABlockCollection, LLaMAForCausalLMModule, LLaMAMLP, LLaMAModule, RMSNorm)
from .models.reference import apply_rotary_emb, precompute_freqs_cis, repeat_kv  # noqa: F401

FlaxLLaMAPreTrainedModel = LLaMAPreTrainedModel
