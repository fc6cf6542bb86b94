import json

import pytest

from jax_llama_b200.config import (LLaMAConfig, ModelArgs, PRESETS, config_from_params, preset,
                                   swiglu_intermediate_size)


def test_defaults_match_reference_schema():
    c = LLaMAConfig()
    assert (c.vocab_size, c.hidden_size, c.intermediate_size, c.num_hidden_layers) == (32000, 4096, 11008, 32)
    assert c.num_key_value_heads == c.num_attention_heads == 32
    assert c.max_sequence_length == 2048 and c.rms_norm_eps == 1e-6 and c.rope_theta == 10000.0
    assert (c.pad_token_id, c.bos_token_id, c.eos_token_id) == (-1, 1, 2)
    assert c.tie_word_embeddings is False and c.use_cache is True
    assert c.to_dict()["model_type"] == "llama"


def test_json_roundtrip(tmp_path):
    c = preset("llama3-8b", some_future_key=3)
    c.save_pretrained(tmp_path)
    d = json.loads((tmp_path / "config.json").read_text())
    assert d["num_key_value_heads"] == 8 and d["some_future_key"] == 3
    assert LLaMAConfig.from_pretrained(str(tmp_path)) == c


@pytest.mark.parametrize("dim,mult,mo,expect", [
    (4096, None, 256, 11008), (5120, None, 256, 13824), (8192, None, 256, 22016),
    (4096, 1.3, 1024, 14336), (8192, 1.3, 4096, 28672), (256, None, 256, 768), (32, None, 2, 86)])
def test_swiglu_width(dim, mult, mo, expect):
    assert swiglu_intermediate_size(dim, mult, mo) == expect


def test_config_from_params_and_unknown_keys():
    args = ModelArgs.from_params_json({"dim": 4096, "n_layers": 32, "n_heads": 32, "n_kv_heads": 8,
                                       "vocab_size": 128256, "multiple_of": 1024, "ffn_dim_multiplier": 1.3,
                                       "norm_eps": 1e-5, "rope_theta": 500000.0, "use_scaled_rope": True})
    c = config_from_params(args)
    assert c.intermediate_size == 14336 and c.num_key_value_heads == 8 and c.rope_theta == 500000.0
    assert c.head_dim == 128 and c.n_rep == 4


def test_presets_are_valid_and_sized():
    for name in PRESETS:
        preset(name)
    assert abs(preset("llama2-7b").num_params() / 1e9 - 6.74) < 0.05
    assert abs(preset("llama3-8b").num_params() / 1e9 - 8.03) < 0.05
    one_b = preset("llama3.2-1b")                      # 64-wide heads, tied embeddings: 1.24 B parameters
    assert one_b.head_dim == 64 and one_b.tie_word_embeddings and abs(one_b.num_params() / 1e9 - 1.24) < 0.02
    assert abs(preset("llama3-70b").num_params() / 1e9 - 70.55) < 0.3


def test_invalid():
    with pytest.raises(ValueError):
        LLaMAConfig(hidden_size=100, num_attention_heads=3)
    with pytest.raises(ValueError):
        LLaMAConfig(num_attention_heads=32, num_key_value_heads=5)


def test_reference_module_paths_resolve():
    """`from jax_llama.<module> import <name>` keeps working with the package name swapped."""
    import importlib
    wanted = {
        "model": ["FlaxLLaMAForCausalLM", "FlaxLLaMAModel", "FlaxLLaMAPreTrainedModel", "RMSNorm", "FlaxLLaMAAttention",
                  "FlaxLLaMAMLP", "FlaxLLaMABlock", "FlaxLLaMABlockCollection", "FlaxLLaMAModule",
                  "FlaxLLaMAForCausalLMModule", "precompute_freqs_cis", "apply_rotary_emb", "repeat_kv"],
        "config": ["LLaMAConfig"],
        "convert_weights": ["ModelArgs", "config_from_params", "convert_llama_weights", "test_intermediate_size"],
        "generation": ["LLaMA"],
        "partition": ["get_partition_spec", "get_llama_param_partition_spec", "global_mesh_defined",
                      "with_sharding_constraint", "with_named_sharding_constraint", "_get_partition_rules_llama"],
        "llama2_tokenizer": ["Tokenizer"],
        "llama3_tokenizer": ["Tokenizer", "ChatFormat", "Message", "Dialog", "Role"],
    }
    for mod, names in wanted.items():
        m = importlib.import_module(f"jax_llama_b200.{mod}")
        for n in names:
            assert hasattr(m, n), f"jax_llama_b200.{mod}.{n}"
