"""Sharded checkpoint cache: convert once, reload each tensor-parallel rank's slice directly.

The reference only knows how to load Meta ``consolidated.*.pth`` files — every start pays the
"load all shards, concatenate, cast to fp32" cost, ~3× the model in host RAM
(/root/reference/jax_llama/convert_weights.py:55-89; its ``convert.sh --out_dir`` is stale and writes
nothing).  This module adds the missing persistence layer:

    cache_dir/
      config.json                      LLaMAConfig (same keys as the reference's HF config)
      layout.json                      {"tp": N, "format": "canonical-in-out", "dtype": "bfloat16"}
      rank{r}-of-{N}.safetensors       rank r's slice of the canonical tree (reference names, [in, out])

`save_sharded` writes it from a full canonical tree or straight from a Meta directory (one rank at a
time, so 70B never needs more than one slice in RAM); `load_rank` memory-maps one file.  `export_canonical`
re-assembles the full reference-named tree for interchange with the reference's own format.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional, Tuple

import torch

from .config import LLaMAConfig
from .convert_weights import convert_llama_weights_for_rank
from .models.llama import _flatten, _unflatten
from .parallel.layout import Mesh, shard_tree, unshard_tree

Tree = Dict[str, Any]


def _rank_file(cache_dir: str, rank: int, tp: int) -> str:
    return os.path.join(cache_dir, f"rank{rank}-of-{tp}.safetensors")


def _write_rank(cache_dir: str, tree: Tree, rank: int, tp: int) -> None:
    from safetensors.torch import save_file

    flat = {k: v.detach().cpu().contiguous() for k, v in _flatten(tree).items()}
    save_file(flat, _rank_file(cache_dir, rank, tp), metadata={"rank": str(rank), "tp": str(tp)})


def _write_meta(cache_dir: str, config: LLaMAConfig, tp: int, dtype: torch.dtype) -> None:
    os.makedirs(cache_dir, exist_ok=True)
    config.save_pretrained(cache_dir)
    with open(os.path.join(cache_dir, "layout.json"), "w") as f:
        json.dump({"tp": tp, "format": "canonical-in-out", "dtype": str(dtype).replace("torch.", "")}, f)


def save_sharded(cache_dir: str, config: LLaMAConfig, tree: Tree, tp: int,
                 dtype: torch.dtype = torch.bfloat16) -> None:
    """Cut a full canonical tree `tp` ways with the partition rules and write one file per rank."""
    _write_meta(cache_dir, config, tp, dtype)
    mesh = Mesh({"dp": 1, "mp": tp})
    for r in range(tp):
        shard = shard_tree(tree, mesh, r)
        _write_rank(cache_dir, {k: v for k, v in _unflatten({k: t.to(dtype) for k, t in _flatten(shard).items()}).items()}, r, tp)


def save_sharded_from_meta(cache_dir: str, ckpt_dir: str, tokenizer: Optional[Any], tp: int,
                           max_seq_len: int = 2048, dtype: torch.dtype = torch.bfloat16) -> LLaMAConfig:
    """Meta directory → sharded cache, one rank at a time (peak RAM ≈ one rank's slice)."""
    config = None
    for r in range(tp):
        shard, config = convert_llama_weights_for_rank(ckpt_dir, tokenizer, tp, r, max_seq_len=max_seq_len, dtype=dtype)
        if r == 0:
            _write_meta(cache_dir, config, tp, dtype)
        _write_rank(cache_dir, shard, r, tp)
        del shard
    assert config is not None
    return config


def read_layout(cache_dir: str) -> Dict[str, Any]:
    with open(os.path.join(cache_dir, "layout.json")) as f:
        return json.load(f)


def load_rank(cache_dir: str, rank: int, tp: Optional[int] = None, device: Any = "cpu") -> Tuple[Tree, LLaMAConfig]:
    """One rank's slice (canonical names) + the config.  `tp` must match how the cache was written."""
    from safetensors.torch import load_file

    layout = read_layout(cache_dir)
    if tp is not None and tp != layout["tp"]:
        raise ValueError(f"cache was written for tp={layout['tp']}, requested tp={tp}: re-shard with save_sharded")
    flat = load_file(_rank_file(cache_dir, rank, layout["tp"]), device=str(device))
    return _unflatten(flat), LLaMAConfig.from_pretrained(cache_dir)


def export_canonical(cache_dir: str) -> Tuple[Tree, LLaMAConfig]:
    """Re-assemble the full tree in the reference's schema from all rank files."""
    layout = read_layout(cache_dir)
    tp = layout["tp"]
    shards, config = [], None
    for r in range(tp):
        s, config = load_rank(cache_dir, r)
        shards.append(s)
    return unshard_tree(shards, Mesh({"dp": 1, "mp": tp})), config
