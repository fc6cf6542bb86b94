import pytest
import torch

from jax_llama_b200 import LLaMA, LLaMA2Tokenizer, LLaMAConfig, LLaMAForCausalLM, preset
from jax_llama_b200.ops.sampling import filter_logits, sample


@pytest.fixture(scope="module")
def model():
    cfg = LLaMAConfig(vocab_size=300, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_sequence_length=128,
                      initializer_range=0.2)
    return LLaMAForCausalLM(cfg, seed=0)


def test_greedy_matches_full_recompute(model):
    ids = torch.randint(3, 300, (2, 7), generator=torch.Generator().manual_seed(0))
    seqs = model.generate(ids, max_new_tokens=6, eos_token_id=-5).sequences
    assert seqs.shape == (2, 13) and torch.equal(seqs[:, :7], ids)
    for t in range(7, 13):
        assert torch.equal(model(seqs[:, :t]).logits[:, -1].argmax(-1), seqs[:, t])


def test_eos_stops_row_and_pads(model):
    ids = torch.randint(3, 300, (2, 5), generator=torch.Generator().manual_seed(1))
    free = model.generate(ids, max_new_tokens=8, eos_token_id=-5).sequences
    eos = int(free[0, 6])                           # second generated token of row 0 becomes "EOS"
    seqs = model.generate(ids, max_new_tokens=8, eos_token_id=eos, pad_token_id=299).sequences
    row = seqs[0].tolist()
    assert row[6] == eos and all(t == 299 for t in row[7:])
    if eos not in free[1, 5:].tolist():
        assert torch.equal(seqs[1], free[1])


def test_tiny_preset_greedy_on_cpu():
    """BASELINE.json config #1: dim 256, 2 layers, 4 heads, vocab 32000, greedy, 1 process."""
    m = LLaMAForCausalLM(preset("tiny"))
    ids = torch.randint(0, 32000, (1, 8), generator=torch.Generator().manual_seed(0))
    seqs = m.generate(ids, max_new_tokens=8).sequences
    assert seqs.shape == (1, 16)
    assert torch.equal(m(seqs[:, :-1]).logits[0, 7:].argmax(-1), seqs[0, 8:])


def test_filter_top_k_top_p():
    logits = torch.log(torch.tensor([[0.5, 0.25, 0.15, 0.07, 0.03]]))
    assert torch.isfinite(filter_logits(logits, top_k=2)).sum() == 2
    assert torch.isfinite(filter_logits(logits, top_p=0.7)).sum() == 2      # 0.5, then 0.25 crosses 0.7
    assert torch.isfinite(filter_logits(logits, top_p=0.76)).sum() == 3
    assert torch.isfinite(filter_logits(logits, top_p=0.01)).sum() == 1     # min_tokens_to_keep
    assert torch.isfinite(filter_logits(logits, top_k=0, top_p=1.0)).sum() == 5


def test_sample_distribution_and_repeatability():
    logits = torch.log(torch.tensor([[0.6, 0.3, 0.1]])).repeat(20000, 1)
    g = torch.Generator().manual_seed(0)
    draws = sample(logits, temperature=1.0, top_k=None, top_p=1.0, generator=g)
    freq = torch.bincount(draws, minlength=3).float() / 20000
    assert torch.allclose(freq, torch.tensor([0.6, 0.3, 0.1]), atol=0.015)
    cold = sample(logits, temperature=0.5, generator=torch.Generator().manual_seed(0))
    f2 = torch.bincount(cold, minlength=3).float() / 20000
    want = torch.tensor([0.36, 0.09, 0.01]) / 0.46
    assert torch.allclose(f2, want, atol=0.015)
    assert torch.equal(sample(logits[:4], temperature=0.0), torch.zeros(4, dtype=torch.long))


def test_sampling_is_repeatable_per_seed(model):
    ids = torch.randint(3, 300, (2, 5))
    a = model.generate(ids, max_new_tokens=6, do_sample=True, temperature=1.0, top_p=0.9, seed=3, eos_token_id=-5).sequences
    b = model.generate(ids, max_new_tokens=6, do_sample=True, temperature=1.0, top_p=0.9, seed=3, eos_token_id=-5).sequences
    assert torch.equal(a, b)


def test_generate_from_str_left_pad_equals_single(spm_model):
    tok = LLaMA2Tokenizer(spm_model)
    cfg = LLaMAConfig(vocab_size=len(tok), hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                      num_attention_heads=4, max_sequence_length=128, initializer_range=0.2,
                      bos_token_id=tok.bos_id, eos_token_id=tok.eos_id)
    model = LLaMAForCausalLM(cfg, seed=1)
    gen = LLaMA(model.params, model, tok)
    prompts = ["The capital of germany is the city of", "Hello world"]
    both = gen.generate_from_str(prompts, max_gen_len=8, temperature=0.0, top_p=1.0)
    singles = [gen.generate_from_str([p], max_gen_len=8, temperature=0.0, top_p=1.0)[0] for p in prompts]
    assert both == singles                                  # left padding must not change the result
    for text, p in zip(both, prompts):
        assert text.startswith(p)                            # output includes the prompt (Q12)
    tokens, mask, rows = gen.encode_prompts(prompts)
    assert tokens[1, 0] == tok.eos_id and mask[1, 0] == 0 and mask[0].all()
    out = gen.generate(tokens, mask, 4, temperature=0.0)
    assert out.shape == (2, tokens.shape[1] + 4)


def test_llama_replace_returns_modified_copy():
    from jax_llama_b200 import LLaMA
    a = LLaMA(params={"x": 1}, model=None, tokenizer="tok")
    b = a.replace(tokenizer="other")
    assert b.tokenizer == "other" and a.tokenizer == "tok" and b.params is a.params
