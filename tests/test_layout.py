import pytest
import torch

from jax_llama_b200 import (Mesh, P, get_llama_param_partition_spec, get_partition_spec, preset,
                            with_named_sharding_constraint, with_sharding_constraint)
from jax_llama_b200.models.reference import init_params
from jax_llama_b200.parallel.layout import (check_tp_divisibility, flatten_tree, shard_tree, unshard_tree)


@pytest.fixture(scope="module")
def tree():
    return init_params(preset("tiny-gqa", num_hidden_layers=2), seed=1)


def test_tp_rules(tree):
    spec = flatten_tree(get_llama_param_partition_spec(tree))
    assert spec[("transformer", "wte", "embedding")] == P("mp", None)
    for w in ("wq", "wk", "wv"):
        assert spec[("transformer", "h", "0", "attention", w, "kernel")] == P(None, "mp")
    assert spec[("transformer", "h", "1", "attention", "wo", "kernel")] == P("mp", None)
    assert spec[("transformer", "h", "1", "feed_forward", "w1", "kernel")] == P(None, "mp")
    assert spec[("transformer", "h", "1", "feed_forward", "w3", "kernel")] == P(None, "mp")
    assert spec[("transformer", "h", "1", "feed_forward", "w2", "kernel")] == P("mp", None)
    assert spec[("transformer", "h", "0", "attention_norm", "kernel")] == P(None)
    assert spec[("transformer", "h", "0", "ffn_norm", "kernel")] == P(None)
    assert spec[("transformer", "ln_f", "kernel")] == P(None)
    assert spec[("lm_head", "kernel")] == P(None, "mp")
    assert set(spec) == set(flatten_tree(tree))


def test_fsdp_rules(tree):
    spec = flatten_tree(get_llama_param_partition_spec(tree, fsdp=True))
    assert spec[("transformer", "wte", "embedding")] == P("mp", "dp")
    assert spec[("transformer", "h", "0", "attention", "wq", "kernel")] == P("dp", "mp")
    assert spec[("transformer", "h", "0", "attention", "wo", "kernel")] == P("mp", "dp")
    assert spec[("transformer", "h", "0", "feed_forward", "w2", "kernel")] == P("mp", "dp")
    assert spec[("lm_head", "kernel")] == P("dp", "mp")


def test_unmatched_leaf_is_an_error():
    with pytest.raises(ValueError, match="Incomplete partition spec"):
        get_partition_spec({"foo": {"bar": torch.zeros(1)}}, [(("baz",), P(None))])


def test_window_matching_is_full_match():
    rules = [(("attention", "w", "kernel"), P("mp")), ((".*",), P(None))]
    spec = flatten_tree(get_partition_spec({"attention": {"wq": {"kernel": 0}, "w": {"kernel": 0}}}, rules))
    assert spec[("attention", "wq", "kernel")] == P(None) and spec[("attention", "w", "kernel")] == P("mp")


@pytest.mark.parametrize("tp", [1, 2])
def test_shard_unshard_roundtrip_tp(tree, tp):
    mesh = Mesh({"dp": 1, "mp": tp})
    shards = [shard_tree(tree, mesh, r) for r in range(tp)]
    q = shards[0]["transformer"]["h"]["0"]["attention"]["wq"]["kernel"]
    assert q.shape == (512, 512 // tp)
    assert shards[0]["transformer"]["h"]["0"]["attention"]["wo"]["kernel"].shape == (512 // tp, 512)
    assert shards[0]["transformer"]["wte"]["embedding"].shape == (1024 // tp, 512)
    back = flatten_tree(unshard_tree(shards, mesh))
    for k, v in flatten_tree(tree).items():
        assert torch.equal(back[k], v), k


def test_shard_unshard_roundtrip_fsdp(tree):
    mesh = Mesh({"dp": 2, "mp": 2})
    shards = [shard_tree(tree, mesh, r, fsdp=True) for r in range(4)]
    assert shards[3]["transformer"]["h"]["0"]["attention"]["wq"]["kernel"].shape == (256, 256)
    back = flatten_tree(unshard_tree(shards, mesh, fsdp=True))
    for k, v in flatten_tree(tree).items():
        assert torch.equal(back[k], v), k


def test_mesh_coords():
    m = Mesh({"dp": 2, "mp": 4})
    assert m.size() == 8 and m.coords(5) == {"mp": 1, "dp": 1} and m.coords(3) == {"mp": 3, "dp": 0}


def test_divisibility():
    check_tp_divisibility(preset("llama3-70b"), 8)
    with pytest.raises(ValueError, match="num_key_value_heads"):
        check_tp_divisibility(preset("llama3-8b"), 16)


def test_constraint_wrappers():
    x = torch.zeros(4, 3)
    assert with_named_sharding_constraint(x, None, P("dp", None)) is x
    assert with_named_sharding_constraint(x, Mesh({"dp": 2, "mp": 1}), P("dp", None)) is x
    with pytest.raises(ValueError):
        with_named_sharding_constraint(x, Mesh({"dp": 3, "mp": 1}), P("dp", None))
    assert with_sharding_constraint(x, P("dp", None)) is x


def test_reference_named_helpers():
    """Names a user of the reference may import: ambient-mesh query, private rule table, width formula."""
    from jax_llama_b200.convert_weights import test_intermediate_size
    from jax_llama_b200.parallel.layout import (Mesh, _get_partition_rules_llama, current_mesh, global_mesh_defined,
                                                partition_rules)
    assert not global_mesh_defined() and current_mesh() is None
    with Mesh({"dp": 1, "mp": 4}) as m:
        assert global_mesh_defined() and current_mesh() is m
        with Mesh({"dp": 2, "mp": 2}) as inner:
            assert current_mesh() is inner
        assert current_mesh() is m
    assert not global_mesh_defined()
    assert _get_partition_rules_llama(fsdp=True) == partition_rules(fsdp=True)
    assert test_intermediate_size(4096, 1.3, 1024) == 14336          # LLaMA-3 8B
    assert test_intermediate_size(8192, 1.3, 4096) == 28672          # LLaMA-3 70B
    assert test_intermediate_size(4096, None, 256) == 11008          # LLaMA-2 7B
