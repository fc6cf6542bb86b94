"""Independent oracle: HF *torch* LlamaForCausalLM loaded from our canonical tree.

HF uses the half-split rotary convention, Meta/ours the interleaved one, so wq/wk rows are permuted
(the standard Meta→HF conversion).  Two independent implementations agreeing stands in for the Meta
repository the reference's own harness compared against (/root/reference/jax_test.py:316-413).
"""
import torch


def _permute(w_out_in: torch.Tensor, n_heads: int) -> torch.Tensor:
    out, inp = w_out_in.shape
    return w_out_in.view(n_heads, out // n_heads // 2, 2, inp).transpose(1, 2).reshape(out, inp)


def hf_model_from_tree(tree, config):
    from transformers import LlamaConfig, LlamaForCausalLM

    hc = LlamaConfig(vocab_size=config.vocab_size, hidden_size=config.hidden_size,
                     intermediate_size=config.intermediate_size, num_hidden_layers=config.num_hidden_layers,
                     num_attention_heads=config.num_attention_heads,
                     num_key_value_heads=config.num_key_value_heads, rms_norm_eps=config.rms_norm_eps,
                     rope_theta=config.rope_theta, max_position_embeddings=config.max_sequence_length * 2,
                     tie_word_embeddings=False, attention_bias=False, mlp_bias=False)
    hc._attn_implementation = "eager"
    m = LlamaForCausalLM(hc).float().eval()
    t = tree["transformer"]
    sd = {"model.embed_tokens.weight": t["wte"]["embedding"], "model.norm.weight": t["ln_f"]["kernel"],
          "lm_head.weight": tree["lm_head"]["kernel"].t()}
    for i in range(config.num_hidden_layers):
        l, p = t["h"][str(i)], f"model.layers.{i}."
        a, f = l["attention"], l["feed_forward"]
        sd[p + "self_attn.q_proj.weight"] = _permute(a["wq"]["kernel"].t(), config.num_attention_heads)
        sd[p + "self_attn.k_proj.weight"] = _permute(a["wk"]["kernel"].t(), config.num_key_value_heads)
        sd[p + "self_attn.v_proj.weight"] = a["wv"]["kernel"].t()
        sd[p + "self_attn.o_proj.weight"] = a["wo"]["kernel"].t()
        sd[p + "mlp.gate_proj.weight"] = f["w1"]["kernel"].t()
        sd[p + "mlp.up_proj.weight"] = f["w3"]["kernel"].t()
        sd[p + "mlp.down_proj.weight"] = f["w2"]["kernel"].t()
        sd[p + "input_layernorm.weight"] = l["attention_norm"]["kernel"]
        sd[p + "post_attention_layernorm.weight"] = l["ffn_norm"]["kernel"]
    missing, unexpected = m.load_state_dict({k: v.float().contiguous() for k, v in sd.items()}, strict=False)
    assert not [k for k in missing if "rotary" not in k], missing
    assert not unexpected, unexpected
    return m
