"""Tensor-parallel forward pass in plain PyTorch + torch.distributed (gloo on CPU, NCCL on GPU).

Executes the Megatron layout that the partition rules describe (column-parallel wq/wk/wv/w1/w3, row-parallel
wo/w2 followed by an all-reduce, vocab-parallel embedding and LM head — /root/reference/jax_llama/
partition.py:62-78) with explicit collectives.  It is the host-side oracle for the fused NVLink kernels and
the way the layout is tested without GPUs (2-process gloo runs in tests/test_tp_cpu.py).
"""
from __future__ import annotations

import math
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist

from ..config import LLaMAConfig
from ..models import reference as ref

Tree = Dict[str, Any]


def tp_forward(shard: Tree, config: LLaMAConfig, input_ids: torch.Tensor, tp: int, rank: int,
               group: Optional[dist.ProcessGroup] = None) -> torch.Tensor:
    """Full-vocabulary logits [B, S, V] from one rank's parameter slice (no KV cache: oracle use)."""
    b, s = input_ids.shape
    nh, nkv, hd = config.num_attention_heads // tp, config.num_key_value_heads // tp, config.head_dim
    t = shard["transformer"]
    # vocab-parallel embedding: masked local lookup + all-reduce (K1 in SURVEY §2.5)
    emb = t["wte"]["embedding"]
    v_local = emb.shape[0]
    local_ids = input_ids - rank * v_local
    mine = (local_ids >= 0) & (local_ids < v_local)
    h = emb[local_ids.clamp(0, v_local - 1)] * mine[..., None].to(emb.dtype)
    if tp > 1:
        dist.all_reduce(h, group=group)
    freqs = ref.precompute_freqs_cis(hd, s, config.rope_theta)[None].expand(b, s, hd // 2)
    causal = torch.ones(s, s, dtype=torch.bool).tril()[None, None]
    for i in range(config.num_hidden_layers):
        l = t["h"][str(i)]
        a = l["attention"]
        x = ref.rms_norm(h, l["attention_norm"]["kernel"], config.rms_norm_eps)
        q = (x @ a["wq"]["kernel"]).view(b, s, nh, hd)
        k = (x @ a["wk"]["kernel"]).view(b, s, nkv, hd)
        v = (x @ a["wv"]["kernel"]).view(b, s, nkv, hd)
        q, k = ref.apply_rotary_emb(q, k, freqs)
        k, v = ref.repeat_kv(k, nh // nkv), ref.repeat_kv(v, nh // nkv)
        sc = torch.einsum("bqhd,bkhd->bhqk", q.float(), k.float()) / math.sqrt(hd)
        sc = sc.masked_fill(~causal, torch.finfo(torch.float32).min)
        o = torch.einsum("bhqk,bkhd->bqhd", torch.softmax(sc, -1), v.float()).to(h.dtype).reshape(b, s, nh * hd)
        part = o @ a["wo"]["kernel"]                       # row-parallel: partial sums (K2)
        if tp > 1:
            dist.all_reduce(part, group=group)
        h = h + part
        x = ref.rms_norm(h, l["ffn_norm"]["kernel"], config.rms_norm_eps)
        f = l["feed_forward"]
        part = (torch.nn.functional.silu(x @ f["w1"]["kernel"]) * (x @ f["w3"]["kernel"])) @ f["w2"]["kernel"]   # K3
        if tp > 1:
            dist.all_reduce(part, group=group)
        h = h + part
    h = ref.rms_norm(h, t["ln_f"]["kernel"], config.rms_norm_eps)
    logits = h @ shard["lm_head"]["kernel"]               # vocab-parallel head: [B, S, V/tp]
    if tp > 1:
        parts = [torch.empty_like(logits) for _ in range(tp)]
        dist.all_gather(parts, logits.contiguous(), group=group)    # K4
        logits = torch.cat(parts, -1)
    return logits
