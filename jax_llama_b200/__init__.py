"""jax_llama_b200 — a B200-native LLaMA inference framework with the capabilities of LSaldyt/JAX_llama.

Public surface mirrors /root/reference/jax_llama/__init__.py:4-10 (plus torch-flavoured names).
"""
from .config import LLaMAConfig, ModelArgs, config_from_params, preset
from .convert_weights import convert_llama_weights, convert_llama_weights_for_rank, write_meta_checkpoint
from .generation import LLaMA
from .models.llama import (FlaxLLaMAForCausalLM, FlaxLLaMAModel, LLaMAForCausalLM, LLaMAModel)
from .parallel.layout import (Mesh, P, PartitionSpec, get_llama_param_partition_spec, get_partition_spec,
                              with_named_sharding_constraint, with_sharding_constraint)
from .tokenizers import ChatFormat, LLaMA2Tokenizer, LLaMA3Tokenizer

__version__ = "0.1.0"

__all__ = [
    "LLaMAConfig", "ModelArgs", "config_from_params", "preset",
    "convert_llama_weights", "convert_llama_weights_for_rank", "write_meta_checkpoint",
    "LLaMA", "LLaMAForCausalLM", "LLaMAModel", "FlaxLLaMAForCausalLM", "FlaxLLaMAModel",
    "Mesh", "P", "PartitionSpec", "get_llama_param_partition_spec", "get_partition_spec",
    "with_named_sharding_constraint", "with_sharding_constraint",
    "LLaMA2Tokenizer", "LLaMA3Tokenizer", "ChatFormat",
]
