"""Converter round-trips on synthetic Meta-format shards."""
import json

import pytest
import torch

from jax_llama_b200 import (LLaMAConfig, Mesh, convert_llama_weights, convert_llama_weights_for_rank,
                            write_meta_checkpoint)
from jax_llama_b200.models.reference import init_params
from jax_llama_b200.parallel.layout import flatten_tree, shard_tree

CFG = LLaMAConfig(vocab_size=64, hidden_size=32, intermediate_size=96, num_hidden_layers=2,
                  num_attention_heads=8, num_key_value_heads=4, rms_norm_eps=1e-5, rope_theta=500000.0)


class FakeTok:
    bos_id, eos_id = 1, 2

    def __len__(self):
        return 64


@pytest.fixture(scope="module")
def tree():
    t = init_params(CFG, seed=7)
    t["transformer"]["ln_f"]["kernel"] = torch.rand(32)
    return t


@pytest.mark.parametrize("mp", [1, 2, 4])
@pytest.mark.parametrize("llama3", [False, True])
def test_roundtrip(tmp_path, tree, mp, llama3):
    write_meta_checkpoint(tree, CFG, str(tmp_path), mp=mp, dtype=torch.float32, llama3_embedding=llama3,
                          multiple_of=32)
    assert len(list(tmp_path.glob("consolidated.*.pth"))) == mp
    got, cfg = convert_llama_weights(str(tmp_path), FakeTok(), max_seq_len=128)
    assert cfg.vocab_size == 64 and cfg.max_sequence_length == 128 and cfg.intermediate_size == 96
    assert cfg.num_key_value_heads == 4 and cfg.rope_theta == 500000.0 and cfg.rms_norm_eps == 1e-5
    flat = flatten_tree(got)
    for k, v in flatten_tree(tree).items():
        assert flat[k].dtype == torch.float32 and torch.equal(flat[k], v), k


def test_shard_shapes_follow_meta_layout(tmp_path, tree):
    write_meta_checkpoint(tree, CFG, str(tmp_path), mp=2, dtype=torch.float32, llama3_embedding=False,
                          multiple_of=32)
    sd = torch.load(tmp_path / "consolidated.01.pth")
    assert sd["tok_embeddings.weight"].shape == (64, 16)            # hidden-sharded (LLaMA-1/2)
    assert sd["layers.0.attention.wq.weight"].shape == (16, 32)     # column parallel: out/mp
    assert sd["layers.0.attention.wk.weight"].shape == (8, 32)
    assert sd["layers.0.attention.wo.weight"].shape == (32, 16)     # row parallel: in/mp
    assert sd["layers.0.feed_forward.w2.weight"].shape == (32, 48)
    assert sd["output.weight"].shape == (32, 32)
    assert json.loads((tmp_path / "params.json").read_text())["n_kv_heads"] == 4


@pytest.mark.parametrize("mp,tp", [(2, 2), (4, 2), (1, 2), (2, 4), (1, 1)])
@pytest.mark.parametrize("llama3", [False, True])
def test_per_rank_loader_equals_sharding_the_full_tree(tmp_path, tree, mp, tp, llama3):
    write_meta_checkpoint(tree, CFG, str(tmp_path), mp=mp, dtype=torch.float32, llama3_embedding=llama3,
                          multiple_of=32)
    mesh = Mesh({"dp": 1, "mp": tp})
    for r in range(tp):
        got, cfg = convert_llama_weights_for_rank(str(tmp_path), FakeTok(), tp, r, dtype=torch.float32)
        want = flatten_tree(shard_tree(tree, mesh, r))
        assert cfg.vocab_size == 64
        for k, v in flatten_tree(got).items():
            assert torch.equal(v, want[k]), (k, r)


def test_unknown_params_key_is_tolerated(tmp_path, tree):
    write_meta_checkpoint(tree, CFG, str(tmp_path), mp=1, dtype=torch.float32, multiple_of=32)
    p = json.loads((tmp_path / "params.json").read_text())
    p["use_scaled_rope"] = True
    (tmp_path / "params.json").write_text(json.dumps(p))
    _, cfg = convert_llama_weights(str(tmp_path), None)
    assert cfg.vocab_size == 64


def test_no_shards(tmp_path):
    with pytest.raises(FileNotFoundError):
        convert_llama_weights(str(tmp_path), None)
