import pytest

from jax_llama_b200 import ChatFormat, LLaMA2Tokenizer, LLaMA3Tokenizer
from jax_llama_b200.tokenizers.llama3 import split_long_runs

TEXTS = ["The capital of Germany is the city of", "hello world this is the city", ""]


def test_llama2_roundtrip(spm_model):
    tok = LLaMA2Tokenizer(spm_model)
    assert len(tok) == tok.n_words == 300 and (tok.bos_id, tok.eos_id, tok.pad_id) == (1, 2, -1)
    for s in TEXTS:
        ids = tok.encode(s, bos=True, eos=True)
        assert ids[0] == tok.bos_id and ids[-1] == tok.eos_id
        assert tok.encode(s, bos=False, eos=False) == ids[1:-1]
        assert tok.decode(ids) == s
    # parity with the raw sentencepiece processor (the reference's oracle is Meta's identical wrapper)
    assert tok.encode(TEXTS[0], False, False) == tok.sp_model.encode(TEXTS[0])


def test_llama3_specials_and_roundtrip(tiktoken_model):
    tok = LLaMA3Tokenizer(tiktoken_model)
    base = 256 + 20
    assert len(tok) == base + 256
    assert tok.bos_id == base and tok.eos_id == base + 1 and tok.pad_id == -1
    sp = tok.special_tokens
    assert sp["<|reserved_special_token_0|>"] == base + 2 and sp["<|reserved_special_token_3|>"] == base + 5
    assert sp["<|start_header_id|>"] == base + 6 and sp["<|end_header_id|>"] == base + 7
    assert sp["<|reserved_special_token_4|>"] == base + 8 and sp["<|eot_id|>"] == base + 9
    assert sp["<|reserved_special_token_5|>"] == base + 10 and sp["<|reserved_special_token_250|>"] == base + 255
    assert tok.stop_tokens == {tok.eos_id, sp["<|eot_id|>"]}
    for s in TEXTS + ["<|begin_of_text|> stays text", "héllo wörld ✓"]:
        ids = tok.encode(s, bos=True, eos=True)
        assert ids[0] == tok.bos_id and ids[-1] == tok.eos_id and tok.bos_id not in ids[1:]
        assert tok.decode(ids[1:-1]) == s
    ids = tok.encode("<|eot_id|>", bos=False, eos=False, allowed_special="all")
    assert ids == [sp["<|eot_id|>"]]
    assert tok.encode("hello world", bos=False, eos=False) == tok.model.encode("hello world")


def test_llama3_long_run_splitting(tiktoken_model):
    s = "a" * 60_000 + " " * 30_000 + "b"
    pieces = list(split_long_runs(s, 25_000))
    assert "".join(pieces) == s and max(len(p) for p in pieces) <= 25_000 + 1 or len(pieces) > 2
    tok = LLaMA3Tokenizer(tiktoken_model)
    assert tok.decode(tok.encode(s, bos=False, eos=False)) == s
    assert list(split_long_runs("", 5)) == [""]


def test_chat_format(tiktoken_model):
    tok = LLaMA3Tokenizer(tiktoken_model)
    fmt = ChatFormat(tok)
    sp = tok.special_tokens
    msg = {"role": "user", "content": "  hello  "}
    head = fmt.encode_header(msg)
    assert head[0] == sp["<|start_header_id|>"] and sp["<|end_header_id|>"] in head
    assert tok.decode(head[1:head.index(sp["<|end_header_id|>"])]) == "user"
    m = fmt.encode_message(msg)
    assert m[-1] == sp["<|eot_id|>"] and tok.decode(m[len(head):-1]) == "hello"
    d = fmt.encode_dialog_prompt([{"role": "system", "content": "s"}, msg])
    assert d[0] == tok.bos_id and d.count(sp["<|eot_id|>"]) == 2
    tail = fmt.encode_header({"role": "assistant", "content": ""})
    assert d[-len(tail):] == tail


def test_missing_file():
    with pytest.raises(FileNotFoundError):
        LLaMA2Tokenizer("/nonexistent.model")
    with pytest.raises(FileNotFoundError):
        LLaMA3Tokenizer("/nonexistent.model")
