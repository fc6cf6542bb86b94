"""Logit processing and token selection: temperature → top-k → top-p → categorical / argmax.

The reference gets this pipeline from HF's Flax generation mixin, configured at
/root/reference/jax_llama/generation.py:32-40 (`do_sample = temperature != 0`, nucleus `top_p`,
HF-4.x default `top_k=50`, `min_tokens_to_keep=1`).  This module is the fp32 PyTorch statement of that
pipeline (CPU path and oracle); the CUDA path (`csrc/sampling.cu`) implements the same selection on
candidates gathered across tensor-parallel ranks.
"""
from __future__ import annotations

from typing import Optional

import torch


def filter_logits(logits: torch.Tensor, temperature: float = 1.0, top_k: Optional[int] = None,
                  top_p: float = 1.0, min_tokens_to_keep: int = 1) -> torch.Tensor:
    """Return fp32 logits with everything outside the top-k / nucleus set at -inf."""
    x = logits.float()
    if temperature != 1.0:
        x = x / temperature
    v = x.shape[-1]
    if top_k is not None and 0 < top_k < v:
        kth = torch.topk(x, max(top_k, min_tokens_to_keep), dim=-1).values[..., -1:]
        x = x.masked_fill(x < kth, float("-inf"))
    if top_p < 1.0:
        sx, si = torch.sort(x, dim=-1, descending=True)
        cum = torch.softmax(sx, dim=-1).cumsum(-1)
        # keep tokens until the cumulative mass first reaches top_p (the crossing token is kept)
        remove = (cum - torch.softmax(sx, dim=-1)) >= top_p
        remove[..., :min_tokens_to_keep] = False
        sx = sx.masked_fill(remove, float("-inf"))
        x = torch.full_like(x, float("-inf")).scatter(-1, si, sx)
    return x


def sample(logits: torch.Tensor, temperature: float = 1.0, top_k: Optional[int] = None,
           top_p: float = 1.0, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """[B, V] logits → [B] token ids.  temperature == 0 selects greedy argmax (generation.py:34)."""
    if temperature == 0.0:
        return logits.float().argmax(dim=-1)
    probs = torch.softmax(filter_logits(logits, temperature, top_k, top_p), dim=-1)
    return torch.multinomial(probs, 1, generator=generator).squeeze(-1)
