"""Plain-PyTorch LLaMA forward pass: the numerical oracle and the CPU execution path.

This is a deliberately un-fused restatement of the model math the reference implements in Flax
(/root/reference/jax_llama/model.py:28-300 RMSNorm / RoPE / GQA attention with KV cache,
:302-400 SwiGLU MLP and pre-norm block, :602-741 backbone and LM head).  It works on the canonical
parameter tree (kernels ``[in, out]``), supports arbitrary key-padding masks, ``output_attentions`` /
``output_hidden_states``, and is what every CUDA kernel in ``csrc/`` is tested against.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple

import torch

from ..config import LLaMAConfig

Tree = Dict[str, Any]


# --------------------------------------------------------------------------- building blocks
def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """x · rsqrt(mean(x²) + eps) · w, statistics in fp32 (model.py:42-48)."""
    xf = x.float()
    y = xf * torch.rsqrt(xf.pow(2).mean(dim=-1, keepdim=True) + eps)
    return (y * weight.float()).to(x.dtype)


def precompute_freqs_cis(dim: int, end: int, theta: float = 10000.0) -> torch.Tensor:
    """Complex rotation table [end, dim/2] (model.py:50-56), built in float64 then cast to complex64."""
    inv = 1.0 / (theta ** (torch.arange(0, dim, 2, dtype=torch.float64)[: dim // 2] / dim))
    ang = torch.outer(torch.arange(end, dtype=torch.float64), inv)
    return torch.polar(torch.ones_like(ang), ang).to(torch.complex64)


def rope_cos_sin(dim: int, end: int, theta: float = 10000.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """Real-valued view of the same table, the layout the CUDA kernels consume: two [end, dim/2] fp32."""
    f = precompute_freqs_cis(dim, end, theta)
    return f.real.contiguous().float(), f.imag.contiguous().float()


def apply_rotary_emb(xq: torch.Tensor, xk: torch.Tensor, freqs_cis: torch.Tensor
                     ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Interleaved-pair (Meta) rotary embedding (model.py:66-92).

    xq/xk: [B, S, heads, hd]; freqs_cis: [B, S, hd/2] complex rows already gathered per position.
    """
    def rot(x: torch.Tensor) -> torch.Tensor:
        xc = torch.view_as_complex(x.float().reshape(*x.shape[:-1], -1, 2))
        out = torch.view_as_real(xc * freqs_cis[:, :, None, :]).flatten(-2)
        return out.to(x.dtype)

    return rot(xq), rot(xk)


def repeat_kv(x: torch.Tensor, n_rep: int) -> torch.Tensor:
    """[B, T, nkv, hd] → [B, T, nkv·n_rep, hd] (model.py:94-103)."""
    if n_rep == 1:
        return x
    b, t, h, d = x.shape
    return x[:, :, :, None, :].expand(b, t, h, n_rep, d).reshape(b, t, h * n_rep, d)


@dataclass
class KVCache:
    """Static KV cache: k/v ``[L, B, nkv, T, hd]`` (head-major so decode streams contiguous rows).

    `index` is the number of filled slots and is shared by the whole batch, as in the reference
    (model.py:177-199); unequal prompts are handled by left padding (`pad_left[b]` leading slots of
    row b hold padding and are never attended).
    """

    k: torch.Tensor
    v: torch.Tensor
    index: int = 0
    pad_left: Optional[torch.Tensor] = None

    @property
    def capacity(self) -> int:
        return self.k.shape[3]

    @classmethod
    def allocate(cls, config: LLaMAConfig, batch: int, max_length: int, dtype: torch.dtype,
                 device: Any = "cpu", n_kv_heads: Optional[int] = None) -> "KVCache":
        nkv = n_kv_heads if n_kv_heads is not None else config.num_key_value_heads
        shape = (config.num_hidden_layers, batch, nkv, max_length, config.head_dim)
        return cls(torch.zeros(shape, dtype=dtype, device=device),
                   torch.zeros(shape, dtype=dtype, device=device), 0,
                   torch.zeros(batch, dtype=torch.int32, device=device))


@dataclass
class ModelOutput:
    logits: Optional[torch.Tensor] = None
    last_hidden_state: Optional[torch.Tensor] = None
    hidden_states: Optional[Tuple[torch.Tensor, ...]] = None
    attentions: Optional[Tuple[torch.Tensor, ...]] = None
    past_key_values: Optional[KVCache] = None

    def to_tuple(self) -> Tuple[Any, ...]:
        return tuple(v for v in (self.logits if self.logits is not None else self.last_hidden_state,
                                 self.hidden_states, self.attentions, self.past_key_values)
                     if v is not None)

    def __getitem__(self, i: int) -> Any:
        return self.to_tuple()[i]


def attention(layer: Tree, config: LLaMAConfig, x: torch.Tensor, freqs: torch.Tensor,
              key_mask: Optional[torch.Tensor], cache: Optional[KVCache], layer_idx: int,
              output_attentions: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Dense masked attention with optional cache (model.py:201-300).

    key_mask: [B, T_visible] bool/int over key slots (T_visible ≥ filled length), or None.
    """
    b, s, _ = x.shape
    nh, nkv, hd = config.num_attention_heads, config.num_key_value_heads, config.head_dim
    a = layer["attention"]
    q = (x @ a["wq"]["kernel"]).view(b, s, nh, hd)
    k = (x @ a["wk"]["kernel"]).view(b, s, nkv, hd)
    v = (x @ a["wv"]["kernel"]).view(b, s, nkv, hd)
    q, k = apply_rotary_emb(q, k, freqs)

    start = 0
    if cache is not None:
        start = cache.index
        if start + s > cache.capacity:
            raise ValueError(f"KV cache overflow: {start}+{s} > capacity {cache.capacity}")
        cache.k[layer_idx, :, :, start:start + s] = k.transpose(1, 2).to(cache.k.dtype)
        cache.v[layer_idx, :, :, start:start + s] = v.transpose(1, 2).to(cache.v.dtype)
        k = cache.k[layer_idx, :, :, :start + s].transpose(1, 2).to(x.dtype)   # only the filled prefix
        v = cache.v[layer_idx, :, :, :start + s].transpose(1, 2).to(x.dtype)
    t = k.shape[1]
    k = repeat_kv(k, nh // nkv)
    v = repeat_kv(v, nh // nkv)

    scores = torch.einsum("bqhd,bkhd->bhqk", q.float(), k.float()) / math.sqrt(hd)
    qpos = torch.arange(start, start + s, device=x.device)[:, None]
    kpos = torch.arange(t, device=x.device)[None, :]
    allowed = (kpos <= qpos)[None, None]                                       # causal
    if key_mask is not None:
        allowed = allowed & key_mask[:, None, None, :t].bool()
    scores = scores.masked_fill(~allowed, torch.finfo(torch.float32).min)      # finite, like model.py:263-267
    probs = torch.softmax(scores, dim=-1)
    out = torch.einsum("bhqk,bkhd->bqhd", probs, v.float()).to(x.dtype).reshape(b, s, nh * hd)
    out = out @ a["wo"]["kernel"]
    return out, (probs if output_attentions else None)


def feed_forward(layer: Tree, x: torch.Tensor) -> torch.Tensor:
    f = layer["feed_forward"]
    return (torch.nn.functional.silu(x @ f["w1"]["kernel"]) * (x @ f["w3"]["kernel"])) @ f["w2"]["kernel"]


def block(layer: Tree, config: LLaMAConfig, h: torch.Tensor, freqs: torch.Tensor,
          key_mask: Optional[torch.Tensor], cache: Optional[KVCache], layer_idx: int,
          output_attentions: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    a, p = attention(layer, config, rms_norm(h, layer["attention_norm"]["kernel"], config.rms_norm_eps),
                     freqs, key_mask, cache, layer_idx, output_attentions)
    h = h + a
    h = h + feed_forward(layer, rms_norm(h, layer["ffn_norm"]["kernel"], config.rms_norm_eps))
    return h, p


_FREQS_CACHE: Dict[Tuple[int, int, float, str], torch.Tensor] = {}


def _freqs_table(config: LLaMAConfig, device: Any) -> torch.Tensor:
    key = (config.head_dim, config.max_sequence_length * 2, config.rope_theta, str(device))
    if key not in _FREQS_CACHE:
        _FREQS_CACHE[key] = precompute_freqs_cis(key[0], key[1], key[2]).to(device)
    return _FREQS_CACHE[key]


def default_position_ids(attention_mask: Optional[torch.Tensor], batch: int, seq: int,
                         device: Any) -> torch.Tensor:
    """cumsum(mask)−1 for left-padded prompts (model.py:758), clamped at 0; arange otherwise."""
    if attention_mask is None:
        return torch.arange(seq, device=device)[None].expand(batch, seq)
    return (attention_mask[:, :seq].long().cumsum(-1) - 1).clamp_(min=0)


def forward(params: Tree, config: LLaMAConfig, input_ids: torch.Tensor,
            attention_mask: Optional[torch.Tensor] = None, position_ids: Optional[torch.Tensor] = None,
            cache: Optional[KVCache] = None, output_attentions: bool = False,
            output_hidden_states: bool = False, with_head: bool = True,
            last_only: bool = False) -> ModelOutput:
    """Whole-model forward.  `attention_mask` covers key slots [0, T) (T ≥ filled cache + S)."""
    b, s = input_ids.shape
    dev = input_ids.device
    t = params["transformer"]
    if position_ids is None:
        if cache is not None and cache.index > 0:
            raise ValueError("position_ids must be given when continuing from a KV cache")
        position_ids = torch.arange(s, device=dev)[None].expand(b, s)
    table = _freqs_table(config, dev)
    if int(position_ids.max()) >= table.shape[0]:
        raise ValueError(f"position {int(position_ids.max())} exceeds the rotary table ({table.shape[0]})")
    freqs = table[position_ids.clamp(min=0)]
    h = t["wte"]["embedding"][input_ids]
    all_h: List[torch.Tensor] = []
    all_p: List[torch.Tensor] = []
    for i in range(config.num_hidden_layers):
        if output_hidden_states:
            all_h.append(h)
        h, p = block(t["h"][str(i)], config, h, freqs, attention_mask, cache, i, output_attentions)
        if output_attentions:
            all_p.append(p)
    if cache is not None:
        cache.index += s
    h = rms_norm(h, t["ln_f"]["kernel"], config.rms_norm_eps)
    if output_hidden_states:
        all_h.append(h)
    out = ModelOutput(last_hidden_state=h, past_key_values=cache,
                      hidden_states=tuple(all_h) if output_hidden_states else None,
                      attentions=tuple(all_p) if output_attentions else None)
    if with_head:
        hh = h[:, -1:] if last_only else h
        w = t["wte"]["embedding"].t() if config.tie_word_embeddings else params["lm_head"]["kernel"]
        out.logits = hh @ w
    return out


def init_params(config: LLaMAConfig, seed: int = 0, dtype: torch.dtype = torch.float32,
                device: Any = "cpu") -> Tree:
    """Random canonical tree: N(0, initializer_range) matrices, unit norm scales (model.py:424-457)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    std = config.initializer_range

    def mat(i: int, o: int) -> torch.Tensor:
        return (torch.randn(i, o, generator=g, dtype=torch.float32) * std).to(dtype).to(device)

    h, f, v = config.hidden_size, config.intermediate_size, config.vocab_size
    kv = config.num_key_value_heads * config.head_dim
    qd = config.num_attention_heads * config.head_dim          # == h unless the config sets head_dim explicitly
    ones = lambda: torch.ones(h, dtype=dtype, device=device)  # noqa: E731
    layers = {}
    for i in range(config.num_hidden_layers):
        layers[str(i)] = {
            "attention": {"wq": {"kernel": mat(h, qd)}, "wk": {"kernel": mat(h, kv)},
                          "wv": {"kernel": mat(h, kv)}, "wo": {"kernel": mat(qd, h)}},
            "feed_forward": {"w1": {"kernel": mat(h, f)}, "w2": {"kernel": mat(f, h)},
                             "w3": {"kernel": mat(h, f)}},
            "attention_norm": {"kernel": ones()}, "ffn_norm": {"kernel": ones()},
        }
    tree: Tree = {"transformer": {"wte": {"embedding": mat(v, h)}, "ln_f": {"kernel": ones()}, "h": layers}}
    if not config.tie_word_embeddings:
        tree["lm_head"] = {"kernel": mat(h, v)}
    return tree
