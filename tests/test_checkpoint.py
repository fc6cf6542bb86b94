import torch

from jax_llama_b200 import LLaMAConfig, Mesh, write_meta_checkpoint
from jax_llama_b200.checkpoint import (export_canonical, load_rank, read_layout, save_sharded,
                                       save_sharded_from_meta)
from jax_llama_b200.models.reference import init_params
from jax_llama_b200.parallel.layout import flatten_tree, shard_tree

CFG = LLaMAConfig(vocab_size=64, hidden_size=32, intermediate_size=96, num_hidden_layers=2,
                  num_attention_heads=8, num_key_value_heads=4, rms_norm_eps=1e-5)


def test_sharded_cache_roundtrip(tmp_path):
    tree = init_params(CFG, seed=3)
    save_sharded(str(tmp_path), CFG, tree, tp=4, dtype=torch.float32)
    assert read_layout(str(tmp_path))["tp"] == 4
    mesh = Mesh({"dp": 1, "mp": 4})
    for r in range(4):
        got, cfg = load_rank(str(tmp_path), r, tp=4)
        want = flatten_tree(shard_tree(tree, mesh, r))
        assert cfg == CFG
        for k, v in flatten_tree(got).items():
            assert torch.equal(v, want[k]), k
    full, cfg = export_canonical(str(tmp_path))
    for k, v in flatten_tree(tree).items():
        assert torch.equal(flatten_tree(full)[k], v), k


def test_cache_from_meta_directory(tmp_path):
    tree = init_params(CFG, seed=4)
    meta, cache = tmp_path / "meta", tmp_path / "cache"
    write_meta_checkpoint(tree, CFG, str(meta), mp=2, dtype=torch.float32, multiple_of=32)
    cfg = save_sharded_from_meta(str(cache), str(meta), None, tp=2, dtype=torch.float32)
    assert cfg.vocab_size == 64
    full, _ = export_canonical(str(cache))
    for k, v in flatten_tree(tree).items():
        assert torch.equal(flatten_tree(full)[k], v), k
