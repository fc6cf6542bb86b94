"""Tensor-parallel layout executed with real collectives on CPU (gloo, 2 and 4 processes)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from jax_llama_b200 import LLaMAConfig, Mesh
from jax_llama_b200.models import reference as ref
from jax_llama_b200.parallel.layout import shard_tree
from jax_llama_b200.parallel.tp_reference import tp_forward

CFG = LLaMAConfig(vocab_size=64, hidden_size=32, intermediate_size=96, num_hidden_layers=2,
                  num_attention_heads=8, num_key_value_heads=4, rms_norm_eps=1e-5, max_sequence_length=32)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    full = ref.init_params(CFG, seed=5)
    ids = torch.randint(0, CFG.vocab_size, (2, 11), generator=torch.Generator().manual_seed(1))
    want = ref.forward(full, CFG, ids).logits
    mine = shard_tree(full, Mesh({"dp": 1, "mp": world}), rank)
    got = tp_forward(mine, CFG, ids, world, rank)
    out[rank] = float((got - want).abs().max())
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_tp_forward_matches_unsharded_reference(world):
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert len(out) == world and max(out.values()) < 1e-4, dict(out)


def test_tp_forward_single_process_is_the_reference():
    full = ref.init_params(CFG, seed=6)
    ids = torch.randint(0, CFG.vocab_size, (1, 9))
    assert torch.allclose(tp_forward(full, CFG, ids, 1, 0), ref.forward(full, CFG, ids).logits, atol=1e-5)
