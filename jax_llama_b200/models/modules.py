"""Module-style handles on the building blocks, for code written against the reference's Flax modules.

The reference exposes its layers as `flax.linen` modules (`RMSNorm`, `FlaxLLaMAAttention`, `FlaxLLaMAMLP`,
`FlaxLLaMABlock`, `FlaxLLaMABlockCollection`, `FlaxLLaMAModule`, `FlaxLLaMAForCausalLMModule`;
/root/reference/jax_llama/model.py:28-48, 105-300, 302-340, 342-400, 548-600, 602-677, 691-741) and its own
parity harness drives them with ``module.apply({"params": subtree}, ...)`` (/root/reference/jax_test.py:73-413).
Here the math lives in plain functions (`models/reference.py`, the fp32 oracle) and in CUDA kernels
(`models/engine.py`); these classes give the same call shape over the same parameter sub-trees so that such
code keeps working.  They are stateless: all weights come in through `variables["params"]`.
"""
from __future__ import annotations

from typing import Any, Dict, Mapping, Optional, Tuple

import torch

from ..config import LLaMAConfig
from . import reference as ref

Variables = Mapping[str, Any]


def _params(variables: Variables) -> Mapping[str, Any]:
    return variables["params"] if "params" in variables else variables


class RMSNorm:
    """y = x · rsqrt(mean(x²) + eps) · kernel (model.py:28-48).  params: {"kernel": [dim]}."""

    def __init__(self, dim: int, eps: float = 1e-6, dtype: Any = None, param_dtype: Any = None):
        self.dim, self.eps = dim, eps

    def init(self, rng: Any = None, *_: Any) -> Dict[str, Any]:
        return {"params": {"kernel": torch.ones(self.dim)}}

    def apply(self, variables: Variables, x: torch.Tensor) -> torch.Tensor:
        return ref.rms_norm(x, _params(variables)["kernel"], self.eps)

    __call__ = apply


class _ConfigModule:
    def __init__(self, config: LLaMAConfig, dtype: Any = None, param_dtype: Any = None, precision: Any = None):
        self.config = config

    def _freqs(self, position_ids: Optional[torch.Tensor], batch: int, seq: int, device: Any) -> torch.Tensor:
        table = ref._freqs_table(self.config, device)
        if position_ids is None:
            position_ids = torch.arange(seq, device=device)[None].expand(batch, seq)
        return table[position_ids.clamp(min=0)]


class LLaMAAttention(_ConfigModule):
    """Rotary GQA attention with optional KV cache (model.py:105-300).
    params: {"wq","wk","wv","wo"} each {"kernel": [in, out]}."""

    def apply(self, variables: Variables, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
              position_ids: Optional[torch.Tensor] = None, cache: Optional[ref.KVCache] = None, layer_idx: int = 0,
              output_attentions: bool = False) -> Tuple[torch.Tensor, ...]:
        b, s, _ = hidden_states.shape
        freqs = self._freqs(position_ids, b, s, hidden_states.device)
        out, probs = ref.attention({"attention": _params(variables)}, self.config, hidden_states, freqs, attention_mask,
                                   cache, layer_idx, output_attentions)
        return (out, probs) if output_attentions else (out,)

    __call__ = apply


class LLaMAMLP(_ConfigModule):
    """SwiGLU feed-forward w2(silu(w1 x) · w3 x) (model.py:302-340).  params: {"w1","w2","w3"}."""

    def apply(self, variables: Variables, x: torch.Tensor, deterministic: bool = True) -> torch.Tensor:
        return ref.feed_forward({"feed_forward": _params(variables)}, x)

    __call__ = apply


class LLaMABlock(_ConfigModule):
    """Pre-norm transformer block (model.py:342-400).
    params: {"attention", "feed_forward", "attention_norm", "ffn_norm"}."""

    def apply(self, variables: Variables, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
              position_ids: Optional[torch.Tensor] = None, cache: Optional[ref.KVCache] = None, layer_idx: int = 0,
              output_attentions: bool = False) -> Tuple[torch.Tensor, ...]:
        b, s, _ = hidden_states.shape
        freqs = self._freqs(position_ids, b, s, hidden_states.device)
        h, probs = ref.block(_params(variables), self.config, hidden_states, freqs, attention_mask, cache, layer_idx,
                             output_attentions)
        return (h, probs) if output_attentions else (h,)

    __call__ = apply


class LLaMABlockCollection(_ConfigModule):
    """The stack of blocks (model.py:548-600).  params: {"0": block, "1": block, ...}."""

    def apply(self, variables: Variables, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
              position_ids: Optional[torch.Tensor] = None, cache: Optional[ref.KVCache] = None,
              output_attentions: bool = False, output_hidden_states: bool = False) -> Tuple[Any, ...]:
        p = _params(variables)
        b, s, _ = hidden_states.shape
        freqs = self._freqs(position_ids, b, s, hidden_states.device)
        all_h, all_p = [], []
        for i in range(self.config.num_hidden_layers):
            if output_hidden_states:
                all_h.append(hidden_states)
            hidden_states, probs = ref.block(p[str(i)], self.config, hidden_states, freqs, attention_mask, cache, i,
                                             output_attentions)
            if output_attentions:
                all_p.append(probs)
        if cache is not None:
            cache.index += s
        return (hidden_states, tuple(all_h) if output_hidden_states else None,
                tuple(all_p) if output_attentions else None)

    __call__ = apply


class LLaMAModule(_ConfigModule):
    """Backbone: embedding → blocks → final norm (model.py:602-677).  params: {"wte", "h", "ln_f"}."""

    def apply(self, variables: Variables, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
              position_ids: Optional[torch.Tensor] = None, cache: Optional[ref.KVCache] = None,
              output_attentions: bool = False, output_hidden_states: bool = False) -> ref.ModelOutput:
        return ref.forward({"transformer": _params(variables)}, self.config, input_ids, attention_mask, position_ids,
                           cache, output_attentions, output_hidden_states, with_head=False)

    __call__ = apply


class LLaMAForCausalLMModule(_ConfigModule):
    """Backbone + LM head over every position (model.py:691-741).  params: {"transformer", "lm_head"}."""

    def apply(self, variables: Variables, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
              position_ids: Optional[torch.Tensor] = None, cache: Optional[ref.KVCache] = None,
              output_attentions: bool = False, output_hidden_states: bool = False) -> ref.ModelOutput:
        return ref.forward(_params(variables), self.config, input_ids, attention_mask, position_ids, cache,
                           output_attentions, output_hidden_states, with_head=True)

    __call__ = apply


# the reference's names
FlaxLLaMAAttention = LLaMAAttention
FlaxLLaMAMLP = LLaMAMLP
FlaxLLaMABlock = LLaMABlock
FlaxLLaMABlockCollection = LLaMABlockCollection
FlaxLLaMAModule = LLaMAModule
FlaxLLaMAForCausalLMModule = LLaMAForCausalLMModule

__all__ = ["RMSNorm", "LLaMAAttention", "LLaMAMLP", "LLaMABlock", "LLaMABlockCollection", "LLaMAModule",
           "LLaMAForCausalLMModule", "FlaxLLaMAAttention", "FlaxLLaMAMLP", "FlaxLLaMABlock",
           "FlaxLLaMABlockCollection", "FlaxLLaMAModule", "FlaxLLaMAForCausalLMModule"]
