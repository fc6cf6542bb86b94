"""Meta checkpoint (``consolidated.NN.pth`` + ``params.json``) → canonical parameter tree.

Same contract as /root/reference/jax_llama/convert_weights.py:52-92: the result is the tree of
SURVEY §2.2 (``transformer/h/{i}/attention/wq/kernel`` …) with linear kernels in ``[in, out]``
layout, plus a `LLaMAConfig` whose ``vocab_size`` comes from the tokenizer.  Differences, on purpose:

* the embedding is re-assembled along whichever axis the shards were actually cut on (the reference
  always uses the hidden axis, which mis-assembles multi-shard LLaMA-3);
* unknown ``params.json`` keys are ignored with a warning instead of raising;
* `convert_llama_weights_for_rank` builds only one tensor-parallel rank's slice, reading just the
  Meta shards that overlap it, so a 70B load never holds 3× the model in host RAM;
* `write_meta_checkpoint` is the inverse (canonical tree → Meta shards), used for interchange and to
  synthesise test fixtures.
"""
from __future__ import annotations

import json
import logging
import re
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

from .config import LLaMAConfig, ModelArgs, config_from_params
from .parallel.layout import Mesh, shard_tree

logger = logging.getLogger(__name__)

Tree = Dict[str, Any]

# Meta key → (canonical path under the layer, axis of the Meta tensor that model-parallelism cuts)
_COLUMN_PARALLEL = ("wq", "wk", "wv", "w1", "w3")   # Meta weight [out/mp, in]  → cut on dim 0
_ROW_PARALLEL = ("wo", "w2")                        # Meta weight [out, in/mp]  → cut on dim 1


def _shard_files(ckpt_dir: str) -> List[Path]:
    found = {}
    for p in Path(ckpt_dir).glob("*.pth"):
        m = re.match(r"consolidated\.(\d+)\.pth$", p.name)
        if m:
            found[int(m.group(1))] = p
    if not found:
        raise FileNotFoundError(f"no consolidated.NN.pth shards in {ckpt_dir}")
    if sorted(found) != list(range(len(found))):
        raise ValueError(f"non-contiguous shard numbering in {ckpt_dir}: {sorted(found)}")
    return [found[i] for i in range(len(found))]


def _load_shard(path: Path) -> Dict[str, torch.Tensor]:
    try:
        return torch.load(path, map_location="cpu", mmap=True, weights_only=True)
    except (RuntimeError, TypeError, ValueError):  # legacy (non-zip) serialisation cannot be mmapped
        return torch.load(path, map_location="cpu", weights_only=True)


def _read_params(ckpt_dir: str) -> Dict[str, Any]:
    with open(Path(ckpt_dir) / "params.json") as f:
        return json.load(f)


def _embedding_axis(shard: torch.Tensor, dim: int, n_shards: int) -> int:
    """LLaMA-1/2 cut the embedding on the hidden axis, LLaMA-3 on the vocab axis: decide from shape."""
    if n_shards == 1:
        return 0
    return 0 if shard.shape[1] == dim else 1


def _cat(parts: Sequence[torch.Tensor], axis: int, dtype: torch.dtype) -> torch.Tensor:
    parts = [p.to(dtype) for p in parts]
    return parts[0] if len(parts) == 1 else torch.cat(parts, dim=axis)


def _assemble(ckpts: Sequence[Dict[str, torch.Tensor]], params: Dict[str, Any],
              dtype: torch.dtype) -> Tree:
    n = len(ckpts)
    dim = params["dim"]

    def col(name: str) -> torch.Tensor:       # [out, in] shards stacked on out → kernel [in, out]
        return _cat([c[name] for c in ckpts], 0, dtype).t().contiguous()

    def row(name: str) -> torch.Tensor:       # [out, in] shards stacked on in → kernel [in, out]
        return _cat([c[name] for c in ckpts], 1, dtype).t().contiguous()

    emb_axis = _embedding_axis(ckpts[0]["tok_embeddings.weight"], dim, n)
    layers: Tree = {}
    for i in range(params["n_layers"]):
        pre = f"layers.{i}."
        layers[str(i)] = {
            "attention": {
                **{w: {"kernel": col(f"{pre}attention.{w}.weight")} for w in ("wq", "wk", "wv")},
                "wo": {"kernel": row(f"{pre}attention.wo.weight")},
            },
            "feed_forward": {
                "w1": {"kernel": col(f"{pre}feed_forward.w1.weight")},
                "w2": {"kernel": row(f"{pre}feed_forward.w2.weight")},
                "w3": {"kernel": col(f"{pre}feed_forward.w3.weight")},
            },
            "attention_norm": {"kernel": ckpts[0][f"{pre}attention_norm.weight"].to(dtype).clone()},
            "ffn_norm": {"kernel": ckpts[0][f"{pre}ffn_norm.weight"].to(dtype).clone()},
        }
    return {
        "transformer": {
            "wte": {"embedding": _cat([c["tok_embeddings.weight"] for c in ckpts], emb_axis, dtype).contiguous()},
            "ln_f": {"kernel": ckpts[0]["norm.weight"].to(dtype).clone()},
            "h": layers,
        },
        "lm_head": {"kernel": col("output.weight")},
    }


def _config_for(params: Dict[str, Any], tokenizer: Optional[Any], max_seq_len: int,
                vocab_rows: int) -> LLaMAConfig:
    params = dict(params)
    params["vocab_size"] = len(tokenizer) if tokenizer is not None else vocab_rows
    params["max_seq_len"] = max_seq_len
    cfg = config_from_params(ModelArgs.from_params_json(params))
    if tokenizer is not None:
        cfg.bos_token_id = tokenizer.bos_id
        cfg.eos_token_id = tokenizer.eos_id
    return cfg


def convert_llama_weights(ckpt_dir: str, tokenizer: Optional[Any] = None, max_seq_len: int = 2048,
                          verbose: bool = False, dtype: torch.dtype = torch.float32
                          ) -> Tuple[Tree, LLaMAConfig]:
    """Load every Meta shard, undo model-parallel sharding, return (canonical tree, config)."""
    files = _shard_files(ckpt_dir)
    ckpts = []
    for i, f in enumerate(files):
        (logger.info if verbose else logger.debug)("loading checkpoint shard %d of %d: %s", i + 1, len(files), f)
        ckpts.append(_load_shard(f))
    params = _read_params(ckpt_dir)
    tree = _assemble(ckpts, params, dtype)
    cfg = _config_for(params, tokenizer, max_seq_len, tree["transformer"]["wte"]["embedding"].shape[0])
    if tree["transformer"]["wte"]["embedding"].shape[0] != cfg.vocab_size:
        logger.warning("tokenizer has %d entries but the embedding has %d rows",
                       cfg.vocab_size, tree["transformer"]["wte"]["embedding"].shape[0])
    return tree, cfg


def convert_llama_weights_for_rank(ckpt_dir: str, tokenizer: Optional[Any], tp: int, rank: int,
                                   max_seq_len: int = 2048, dtype: torch.dtype = torch.bfloat16
                                   ) -> Tuple[Tree, LLaMAConfig]:
    """Canonical-layout slice for one TP rank, touching only the Meta shards that overlap it.

    When the checkpoint was written with MP == tp this reads exactly one file per rank.  Other ratios
    (MP a multiple of tp, or tp a multiple of MP) read the covering shards and slice.
    """
    files = _shard_files(ckpt_dir)
    params = _read_params(ckpt_dir)
    mp = len(files)
    mesh = Mesh({"dp": 1, "mp": tp})
    if mp % tp == 0:
        group = mp // tp
        ckpts = [_load_shard(f) for f in files[rank * group:(rank + 1) * group]]
        # the covering shards ARE this rank's slice: assemble them as if they were a whole model
        sub = _assemble(ckpts, params, dtype)
        emb = sub["transformer"]["wte"]["embedding"]
        first = _load_shard(files[0])["tok_embeddings.weight"] if mp > 1 else None
        if mp > 1 and _embedding_axis(first, params["dim"], mp) == 1:
            # LLaMA-1/2 hidden-sharded embedding: our layout is vocab-parallel, so rebuild it
            full = _cat([_load_shard(f)["tok_embeddings.weight"] for f in files], 1, dtype)
            step = full.shape[0] // tp
            emb = full[rank * step:(rank + 1) * step].contiguous()
        sub["transformer"]["wte"]["embedding"] = emb
        vocab_rows = emb.shape[0] * tp
        tree = sub
    elif tp % mp == 0:
        inner = tp // mp
        ckpt = _load_shard(files[rank // inner])
        whole = _assemble([ckpt], params, dtype)
        # the one shard is 1/mp of the model; cut it `inner` more ways, except replicated leaves
        tree = shard_tree(whole, Mesh({"dp": 1, "mp": inner}), rank % inner)
        if mp > 1 and _embedding_axis(ckpt["tok_embeddings.weight"], params["dim"], mp) == 1:
            full = _cat([_load_shard(f)["tok_embeddings.weight"] for f in files], 1, dtype)
            step = full.shape[0] // tp
            tree["transformer"]["wte"]["embedding"] = full[rank * step:(rank + 1) * step].contiguous()
        vocab_rows = tree["transformer"]["wte"]["embedding"].shape[0] * tp
    else:
        full, _ = convert_llama_weights(ckpt_dir, tokenizer, max_seq_len, dtype=dtype)
        tree = shard_tree(full, mesh, rank)
        vocab_rows = full["transformer"]["wte"]["embedding"].shape[0]
    cfg = _config_for(params, tokenizer, max_seq_len, vocab_rows)
    return tree, cfg


def write_meta_checkpoint(tree: Tree, config: LLaMAConfig, out_dir: str, mp: int = 1,
                          dtype: torch.dtype = torch.bfloat16, llama3_embedding: Optional[bool] = None,
                          multiple_of: int = 256, ffn_dim_multiplier: Optional[float] = None) -> None:
    """Canonical tree → Meta-format directory with `mp` shards (inverse of `convert_llama_weights`).

    `llama3_embedding` selects the embedding cut (vocab axis for LLaMA-3, hidden axis for LLaMA-1/2);
    by default it follows the vocabulary size.  `multiple_of`/`ffn_dim_multiplier` are recorded in
    ``params.json`` and must reproduce ``config.intermediate_size`` through Meta's width formula.
    """
    from .config import swiglu_intermediate_size

    if swiglu_intermediate_size(config.hidden_size, ffn_dim_multiplier, multiple_of) != config.intermediate_size:
        raise ValueError("multiple_of / ffn_dim_multiplier do not reproduce config.intermediate_size")
    if llama3_embedding is None:
        llama3_embedding = config.vocab_size > 100_000
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    t = tree["transformer"]

    def cut(x: torch.Tensor, axis: int, r: int) -> torch.Tensor:
        step = x.shape[axis] // mp
        if x.shape[axis] % mp:
            raise ValueError(f"cannot cut dimension of size {x.shape[axis]} into {mp} shards")
        return x.narrow(axis, r * step, step).to(dtype).contiguous().clone()

    for r in range(mp):
        sd: Dict[str, torch.Tensor] = {}
        sd["tok_embeddings.weight"] = cut(t["wte"]["embedding"], 0 if llama3_embedding else 1, r)
        sd["norm.weight"] = t["ln_f"]["kernel"].to(dtype).clone()
        sd["output.weight"] = cut(tree["lm_head"]["kernel"].t(), 0, r)
        for i in range(config.num_hidden_layers):
            layer = t["h"][str(i)]
            pre = f"layers.{i}."
            for w in ("wq", "wk", "wv"):
                sd[f"{pre}attention.{w}.weight"] = cut(layer["attention"][w]["kernel"].t(), 0, r)
            sd[f"{pre}attention.wo.weight"] = cut(layer["attention"]["wo"]["kernel"].t(), 1, r)
            sd[f"{pre}feed_forward.w1.weight"] = cut(layer["feed_forward"]["w1"]["kernel"].t(), 0, r)
            sd[f"{pre}feed_forward.w3.weight"] = cut(layer["feed_forward"]["w3"]["kernel"].t(), 0, r)
            sd[f"{pre}feed_forward.w2.weight"] = cut(layer["feed_forward"]["w2"]["kernel"].t(), 1, r)
            sd[f"{pre}attention_norm.weight"] = layer["attention_norm"]["kernel"].to(dtype).clone()
            sd[f"{pre}ffn_norm.weight"] = layer["ffn_norm"]["kernel"].to(dtype).clone()
        sd["rope.freqs"] = torch.zeros(config.head_dim // 2, dtype=dtype)  # present in Meta files; ignored
        torch.save(sd, out / f"consolidated.{r:02d}.pth")
    params = {
        "dim": config.hidden_size, "n_layers": config.num_hidden_layers,
        "n_heads": config.num_attention_heads, "n_kv_heads": config.num_key_value_heads,
        "multiple_of": multiple_of, "norm_eps": config.rms_norm_eps, "rope_theta": config.rope_theta,
        "vocab_size": -1,
    }
    if ffn_dim_multiplier is not None:
        params["ffn_dim_multiplier"] = ffn_dim_multiplier
    with open(out / "params.json", "w") as f:
        json.dump(params, f)


def test_intermediate_size(dim: int, ffn_dim_multiplier: Optional[float], multiple_of: int) -> int:
    """The reference's name for the SwiGLU width formula (convert_weights.py:28-33); same as
    `config.swiglu_intermediate_size`."""
    from .config import swiglu_intermediate_size
    return swiglu_intermediate_size(dim, ffn_dim_multiplier, multiple_of)


test_intermediate_size.__test__ = False                  # a helper with a reference-given name, not a pytest test


def main(argv: Optional[Sequence[str]] = None) -> int:
    """``python -m jax_llama_b200.convert_weights --ckpt_dir D --out_dir O [--tp N]`` — the command the
    reference's convert.sh calls but never shipped (/root/reference/convert.sh:10): Meta shards → the sharded
    safetensors cache of `checkpoint.py`, one rank slice at a time."""
    import argparse

    from .checkpoint import save_sharded_from_meta

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--ckpt_dir", required=True)
    ap.add_argument("--out_dir", required=True)
    ap.add_argument("--tokenizer_path", default=None, help="sets vocab_size / bos / eos ids in config.json")
    ap.add_argument("--is_llama3", type=lambda s: str(s).lower() in ("1", "true", "yes"), default=False)
    ap.add_argument("--tp", type=int, default=1, help="number of tensor-parallel rank files to write")
    ap.add_argument("--max_seq_len", type=int, default=2048)
    ap.add_argument("--dtype", default="bfloat16", choices=["bfloat16", "float16", "float32"])
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO if a.verbose else logging.WARNING)
    tok = None
    if a.tokenizer_path:
        from .tokenizers import LLaMA2Tokenizer, LLaMA3Tokenizer
        tok = (LLaMA3Tokenizer if a.is_llama3 else LLaMA2Tokenizer)(a.tokenizer_path)
    cfg = save_sharded_from_meta(a.out_dir, a.ckpt_dir, tok, a.tp, max_seq_len=a.max_seq_len,
                                 dtype=getattr(torch, a.dtype))
    logger.info("wrote %d rank file(s) to %s (%d layers, hidden %d)", a.tp, a.out_dir, cfg.num_hidden_layers,
                cfg.hidden_size)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
