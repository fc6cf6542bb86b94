"""Unit parity tests mirroring /root/reference/jax_test.py:73-413 (same tiny config, random trials,
atol 1e-2 — we hold 1e-4 because both sides are fp32 PyTorch) plus the gaps SURVEY §4 lists."""
import math

import pytest
import torch

from jax_llama_b200 import LLaMAConfig, LLaMAForCausalLM, LLaMAModel
from jax_llama_b200.models import reference as ref
from tests.oracles import hf_model_from_tree

CFG = LLaMAConfig(vocab_size=256, hidden_size=32, intermediate_size=86, num_hidden_layers=4,
                  num_attention_heads=4, num_key_value_heads=2, rms_norm_eps=1e-5, max_sequence_length=64)
TRIALS = 16


def test_rmsnorm_matches_definition():
    for s in range(TRIALS):
        g = torch.Generator().manual_seed(s)
        x, w = torch.randn(1, 32, generator=g), torch.randn(32, generator=g)
        want = x / torch.sqrt((x * x).mean(-1, keepdim=True) + 1e-5) * w
        assert torch.allclose(ref.rms_norm(x, w, 1e-5), want, atol=1e-5)


def test_precompute_freqs_cis():
    f = ref.precompute_freqs_cis(8, 128, 10000.0)
    assert f.shape == (128, 4) and f.dtype == torch.complex64
    for pos in (0, 1, 77):
        for i in range(4):
            ang = pos / (10000.0 ** (2 * i / 8))
            assert abs(f[pos, i].real - math.cos(ang)) < 1e-5 and abs(f[pos, i].imag - math.sin(ang)) < 1e-5
    c, s = ref.rope_cos_sin(8, 128)
    assert torch.equal(c, f.real) and torch.equal(s, f.imag)


def test_apply_rotary_emb_interleaved():
    g = torch.Generator().manual_seed(0)
    q, k = torch.randn(1, 64, 4, 8, generator=g), torch.randn(1, 64, 2, 8, generator=g)
    f = ref.precompute_freqs_cis(8, 64)[None]
    rq, rk = ref.apply_rotary_emb(q, k, f)
    # explicit pairwise rotation
    c, s = f.real[:, :, None, :], f.imag[:, :, None, :]
    want = torch.stack([q[..., 0::2] * c - q[..., 1::2] * s, q[..., 0::2] * s + q[..., 1::2] * c], -1).flatten(-2)
    assert torch.allclose(rq, want, atol=1e-6)
    assert torch.allclose(rk.norm(dim=-1), k.norm(dim=-1), atol=1e-5)   # rotation preserves length


def test_repeat_kv():
    x = torch.arange(2 * 3 * 2 * 4.0).view(2, 3, 2, 4)
    y = ref.repeat_kv(x, 3)
    assert y.shape == (2, 3, 6, 4)
    for h in range(6):
        assert torch.equal(y[:, :, h], x[:, :, h // 3])


@pytest.fixture(scope="module")
def tree():
    t = ref.init_params(CFG, seed=3)
    g = torch.Generator().manual_seed(5)
    for i in range(CFG.num_hidden_layers):       # random norm scales, as jax_test.py:222-314 does
        for n in ("attention_norm", "ffn_norm"):
            t["transformer"]["h"][str(i)][n]["kernel"] = 1 + 0.1 * torch.randn(32, generator=g)
    t["transformer"]["ln_f"]["kernel"] = 1 + 0.1 * torch.randn(32, generator=g)
    # larger weights so that attention is not trivially uniform
    for k in ("wq", "wk"):
        for i in range(CFG.num_hidden_layers):
            t["transformer"]["h"][str(i)]["attention"][k]["kernel"] *= 20
    return t


@pytest.fixture(scope="module")
def hf(tree):
    return hf_model_from_tree(tree, CFG)


def test_transformer_matches_hf_torch(tree, hf):
    """Whole model vs an independent implementation (GQA, RoPE permutation, SwiGLU, norms)."""
    worst = 0.0
    for s in range(TRIALS):
        ids = torch.randint(0, 256, (2, 48), generator=torch.Generator().manual_seed(s))
        with torch.no_grad():
            want = hf(ids).logits
            got = ref.forward(tree, CFG, ids).logits
        worst = max(worst, (want - got).abs().max().item())
    assert worst < 1e-4, worst


def test_block_pieces_match_hf(tree, hf):
    ids = torch.randint(0, 256, (1, 24), generator=torch.Generator().manual_seed(9))
    with torch.no_grad():
        out = ref.forward(tree, CFG, ids, output_hidden_states=True, output_attentions=True)
        want = hf(ids, output_hidden_states=True, output_attentions=True)
    assert len(out.hidden_states) == CFG.num_hidden_layers + 1 and len(out.attentions) == CFG.num_hidden_layers
    for a, b in zip(out.hidden_states[:-1], want.hidden_states[:-1]):
        assert torch.allclose(a, b, atol=1e-4)
    for a, b in zip(out.attentions, want.attentions):
        assert a.shape == (1, 4, 24, 24) and torch.allclose(a, b, atol=1e-5)


def test_feed_forward_formula(tree):
    layer = tree["transformer"]["h"]["0"]
    x = torch.randn(3, 32)
    f = layer["feed_forward"]
    a, b = x @ f["w1"]["kernel"], x @ f["w3"]["kernel"]
    assert torch.allclose(ref.feed_forward(layer, x), (a * torch.sigmoid(a) * b) @ f["w2"]["kernel"], atol=1e-6)


def test_cache_equals_no_cache(tree):
    ids = torch.randint(0, 256, (2, 20), generator=torch.Generator().manual_seed(1))
    full = ref.forward(tree, CFG, ids).logits
    cache = ref.KVCache.allocate(CFG, 2, 32, torch.float32)
    pre = ref.forward(tree, CFG, ids[:, :12], cache=cache).logits
    assert torch.allclose(pre, full[:, :12], atol=1e-5) and cache.index == 12
    for t in range(12, 20):
        pos = torch.full((2, 1), t)
        step = ref.forward(tree, CFG, ids[:, t:t + 1], position_ids=pos, cache=cache).logits
        assert torch.allclose(step[:, 0], full[:, t], atol=1e-5), t
    assert cache.index == 20
    with pytest.raises(ValueError):
        ref.forward(tree, CFG, ids[:, :1], cache=cache)       # positions required once the cache is warm


def test_left_padding_is_invisible(tree):
    """A left-padded row must produce the logits of the un-padded prompt (generation.py:55-60)."""
    ids = torch.randint(3, 256, (1, 9), generator=torch.Generator().manual_seed(2))
    want = ref.forward(tree, CFG, ids).logits[0, -1]
    pad = 4
    padded = torch.cat([torch.full((1, pad), 2), ids], 1)
    mask = torch.cat([torch.zeros(1, pad, dtype=torch.int32), torch.ones(1, 9, dtype=torch.int32)], 1)
    pos = ref.default_position_ids(mask, 1, pad + 9, "cpu")
    assert pos[0].tolist() == [0] * pad + list(range(9))
    got = ref.forward(tree, CFG, padded, attention_mask=mask, position_ids=pos).logits[0, -1]
    assert torch.allclose(got, want, atol=1e-5)


def test_wrapper_api(tree):
    m = LLaMAForCausalLM(CFG, _do_init=False)
    with pytest.raises(ValueError):
        _ = m.params
    ids = torch.randint(0, 256, (2, 6))
    out = m(ids, params=tree, output_hidden_states=True)
    assert out.logits.shape == (2, 6, 256) and out.past_key_values is None
    tup = m(ids, params=tree, return_dict=False)
    assert torch.equal(tup[0], out.logits)
    cache = m.init_cache(2, 16, tree)
    assert cache.k.shape == (4, 2, 2, 16, 8)
    o1 = m(ids, params=tree, past_key_values=cache)
    assert o1.past_key_values.index == 6
    with pytest.raises(ValueError, match="position_ids"):
        m(ids[:, :1], params=tree, past_key_values=cache)
    base = LLaMAModel(CFG, _do_init=False)(ids, params=tree)
    assert base.logits is None and base.last_hidden_state.shape == (2, 6, 32)
    with pytest.raises(NotImplementedError):
        m(ids, params=tree, train=True)


def test_tied_embeddings():
    cfg = LLaMAConfig(vocab_size=64, hidden_size=16, intermediate_size=32, num_hidden_layers=1,
                      num_attention_heads=2, tie_word_embeddings=True)
    m = LLaMAForCausalLM(cfg)
    assert "lm_head" not in m.params
    ids = torch.randint(0, 64, (1, 5))
    out = m(ids)
    h = LLaMAModel(cfg, _do_init=False)(ids, params=m.params).last_hidden_state
    assert torch.allclose(out.logits, h @ m.params["transformer"]["wte"]["embedding"].t(), atol=1e-6)


def test_save_and_from_pretrained(tmp_path, tree):
    m = LLaMAForCausalLM(CFG, _do_init=False)
    m.save_pretrained(str(tmp_path), params=tree)
    m2 = LLaMAForCausalLM.from_pretrained(str(tmp_path))
    ids = torch.randint(0, 256, (1, 5))
    assert torch.equal(m2(ids).logits, m(ids, params=tree).logits) and m2.config == CFG


def test_module_style_wrappers(tree, hf):
    """The reference's Flax module names, driven with `.apply({"params": subtree}, ...)` like jax_test.py does."""
    from jax_llama_b200.models import (FlaxLLaMAAttention, FlaxLLaMABlock, FlaxLLaMABlockCollection,
                                       FlaxLLaMAForCausalLMModule, FlaxLLaMAMLP, FlaxLLaMAModule, RMSNorm)
    layer = tree["transformer"]["h"]["0"]
    x = torch.randn(2, 10, 32, generator=torch.Generator().manual_seed(4))
    f = ref._freqs_table(CFG, "cpu")[torch.arange(10)[None].expand(2, 10)]
    norm = RMSNorm(32, eps=CFG.rms_norm_eps)
    assert torch.equal(norm.apply({"params": layer["attention_norm"]}, x),
                       ref.rms_norm(x, layer["attention_norm"]["kernel"], CFG.rms_norm_eps))
    assert norm.init()["params"]["kernel"].shape == (32,)
    got = FlaxLLaMAAttention(CFG).apply({"params": layer["attention"]}, x)[0]
    assert torch.equal(got, ref.attention(layer, CFG, x, f, None, None, 0)[0])
    assert torch.equal(FlaxLLaMAMLP(CFG).apply({"params": layer["feed_forward"]}, x), ref.feed_forward(layer, x))
    out, probs = FlaxLLaMABlock(CFG).apply({"params": layer}, x, output_attentions=True)
    want, wprobs = ref.block(layer, CFG, x, f, None, None, 0, True)
    assert torch.equal(out, want) and torch.equal(probs, wprobs)
    ids = torch.randint(0, 256, (2, 12), generator=torch.Generator().manual_seed(6))
    full = ref.forward(tree, CFG, ids, output_hidden_states=True)
    h0 = tree["transformer"]["wte"]["embedding"][ids]
    stack = FlaxLLaMABlockCollection(CFG).apply({"params": tree["transformer"]["h"]}, h0, output_hidden_states=True)
    assert torch.allclose(ref.rms_norm(stack[0], tree["transformer"]["ln_f"]["kernel"], CFG.rms_norm_eps),
                          full.last_hidden_state, atol=1e-6) and len(stack[1]) == CFG.num_hidden_layers
    back = FlaxLLaMAModule(CFG).apply({"params": tree["transformer"]}, ids)
    assert back.logits is None and torch.equal(back.last_hidden_state, full.last_hidden_state)
    lm = FlaxLLaMAForCausalLMModule(CFG).apply({"params": tree}, ids)
    assert torch.equal(lm.logits, full.logits)
    with torch.no_grad():
        assert (lm.logits - hf(ids).logits).abs().max() < 1e-4
    assert LLaMAForCausalLM.module_class is FlaxLLaMAForCausalLMModule and LLaMAModel.module_class is FlaxLLaMAModule
