"""Model configuration.

`LLaMAConfig` keeps the field names, defaults and JSON schema of the reference's HF-style config
(/root/reference/jax_llama/config.py:72-116) but is a free-standing class: it does not depend on
`transformers`, so it loads in every process of a torchrun job without importing HF.

`ModelArgs` / `config_from_params` mirror Meta's ``params.json`` schema and the SwiGLU width formula
(/root/reference/jax_llama/convert_weights.py:13-50).  Unlike the reference, unknown ``params.json``
keys (e.g. LLaMA-3.1's ``use_scaled_rope``) are tolerated and reported instead of raising.
"""
from __future__ import annotations

import copy
import dataclasses
import json
import logging
import os
from dataclasses import dataclass
from typing import Any, Dict, Optional

logger = logging.getLogger(__name__)

CONFIG_NAME = "config.json"


class LLaMAConfig:
    """LLaMA-1/2/3 architecture description (same keys as the reference's config.json)."""

    model_type = "llama"

    def __init__(
        self,
        vocab_size: int = 32000,
        hidden_size: int = 4096,
        intermediate_size: int = 11008,
        num_hidden_layers: int = 32,
        num_attention_heads: int = 32,
        num_key_value_heads: Optional[int] = None,
        max_sequence_length: int = 2048,
        rms_norm_eps: float = 1e-6,
        initializer_range: float = 0.02,
        use_cache: bool = True,
        pad_token_id: int = -1,
        bos_token_id: int = 1,
        eos_token_id: int = 2,
        resid_pdrop: float = 0.0,
        embd_pdrop: float = 0.0,
        attn_pdrop: float = 0.0,
        tie_word_embeddings: bool = False,
        gradient_checkpointing: bool = False,
        rope_theta: float = 10000.0,
        **kwargs: Any,
    ):
        self.vocab_size = int(vocab_size)
        self.hidden_size = int(hidden_size)
        self.intermediate_size = int(intermediate_size)
        self.num_hidden_layers = int(num_hidden_layers)
        self.num_attention_heads = int(num_attention_heads)
        if num_key_value_heads is None:
            num_key_value_heads = num_attention_heads
        self.num_key_value_heads = int(num_key_value_heads)
        self.max_sequence_length = int(max_sequence_length)
        self.rms_norm_eps = float(rms_norm_eps)
        self.initializer_range = float(initializer_range)
        self.use_cache = bool(use_cache)
        self.pad_token_id = pad_token_id
        self.bos_token_id = bos_token_id
        self.eos_token_id = eos_token_id
        self.resid_pdrop = float(resid_pdrop)
        self.embd_pdrop = float(embd_pdrop)
        self.attn_pdrop = float(attn_pdrop)
        self.tie_word_embeddings = bool(tie_word_embeddings)
        # accepted for schema parity; inference-only framework ignores it (SURVEY §2.4 "remat")
        self.gradient_checkpointing = bool(gradient_checkpointing)
        self.rope_theta = float(rope_theta)
        kwargs.pop("model_type", None)
        # unknown keys are preserved so a config.json written by HF round-trips
        self.extra: Dict[str, Any] = dict(kwargs)
        self.validate()

    # ------------------------------------------------------------------ derived
    @property
    def head_dim(self) -> int:
        """hidden_size // num_attention_heads as in the reference (model.py:115), unless the config carries an
        explicit ``head_dim`` key (newer HF Llama configs do): then q/o project to ``num_attention_heads * head_dim``
        columns, which need not equal hidden_size — e.g. the per-rank slice of a tensor-parallel model as a model."""
        hd = self.extra.get("head_dim") if hasattr(self, "extra") else None
        return int(hd) if hd else self.hidden_size // self.num_attention_heads

    @property
    def n_rep(self) -> int:
        return self.num_attention_heads // self.num_key_value_heads

    def validate(self) -> None:
        if not self.extra.get("head_dim") and self.hidden_size % self.num_attention_heads:
            raise ValueError("hidden_size must be divisible by num_attention_heads")
        if self.num_attention_heads % self.num_key_value_heads:
            raise ValueError("num_attention_heads must be divisible by num_key_value_heads")
        if self.head_dim % 2:
            raise ValueError("head_dim must be even for rotary embeddings")

    def num_params(self) -> int:
        h, f, v, l = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        kv = self.num_key_value_heads * self.head_dim
        per_layer = h * self.num_attention_heads * self.head_dim * 2 + 2 * h * kv + 3 * h * f + 2 * h
        head = 0 if self.tie_word_embeddings else v * h
        return v * h + l * per_layer + h + head

    # ------------------------------------------------------------------ (de)serialisation
    def to_dict(self) -> Dict[str, Any]:
        d = {k: copy.deepcopy(v) for k, v in self.__dict__.items() if k != "extra"}
        d.update(copy.deepcopy(self.extra))
        d["model_type"] = self.model_type
        return d

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "LLaMAConfig":
        return cls(**dict(d))

    def to_json_string(self) -> str:
        return json.dumps(self.to_dict(), indent=2, sort_keys=True) + "\n"

    def to_json_file(self, path: str) -> None:
        with open(path, "w") as f:
            f.write(self.to_json_string())

    @classmethod
    def from_json_file(cls, path: str) -> "LLaMAConfig":
        with open(path) as f:
            return cls.from_dict(json.load(f))

    def save_pretrained(self, save_directory: str) -> None:
        os.makedirs(save_directory, exist_ok=True)
        self.to_json_file(os.path.join(save_directory, CONFIG_NAME))

    @classmethod
    def from_pretrained(cls, path: str, **overrides: Any) -> "LLaMAConfig":
        if os.path.isdir(path):
            path = os.path.join(path, CONFIG_NAME)
        cfg = cls.from_json_file(path)
        for k, v in overrides.items():
            setattr(cfg, k, v)
        cfg.validate()
        return cfg

    def __eq__(self, other: object) -> bool:
        return isinstance(other, LLaMAConfig) and self.to_dict() == other.to_dict()

    def __repr__(self) -> str:
        return f"LLaMAConfig {self.to_json_string()}"


@dataclass
class ModelArgs:
    """Meta ``params.json`` schema (/root/reference/jax_llama/convert_weights.py:13-26)."""

    dim: int = 512
    n_layers: int = 8
    n_heads: int = 8
    n_kv_heads: Optional[int] = None
    vocab_size: int = -1
    multiple_of: int = 256
    ffn_dim_multiplier: Optional[float] = None
    norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    max_batch_size: int = 32
    max_seq_len: int = 2048

    @classmethod
    def from_params_json(cls, params: Dict[str, Any]) -> "ModelArgs":
        known = {f.name for f in dataclasses.fields(cls)}
        unknown = sorted(set(params) - known)
        if unknown:
            logger.warning("params.json keys ignored (not part of the LLaMA-1/2/3 schema): %s", unknown)
        return cls(**{k: v for k, v in params.items() if k in known})


def swiglu_intermediate_size(dim: int, ffn_dim_multiplier: Optional[float], multiple_of: int) -> int:
    """Meta's FFN width rule: 2/3·4·dim, optional multiplier, rounded up to `multiple_of`."""
    width = int(2 * (dim * 4) / 3)
    if ffn_dim_multiplier is not None:
        width = int(ffn_dim_multiplier * width)
    return multiple_of * ((width + multiple_of - 1) // multiple_of)


def config_from_params(args: ModelArgs) -> LLaMAConfig:
    return LLaMAConfig(
        vocab_size=args.vocab_size,
        hidden_size=args.dim,
        intermediate_size=swiglu_intermediate_size(args.dim, args.ffn_dim_multiplier, args.multiple_of),
        num_hidden_layers=args.n_layers,
        num_attention_heads=args.n_heads,
        num_key_value_heads=args.n_kv_heads,
        max_sequence_length=args.max_seq_len,
        rms_norm_eps=args.norm_eps,
        rope_theta=args.rope_theta,
    )


# Architecture presets for the configurations BASELINE.json names (random-init benchmarking).
PRESETS: Dict[str, Dict[str, Any]] = {
    "tiny": dict(vocab_size=32000, hidden_size=256, intermediate_size=768, num_hidden_layers=2,
                 num_attention_heads=4, num_key_value_heads=4, rms_norm_eps=1e-5),
    "tiny-gqa": dict(vocab_size=1024, hidden_size=512, intermediate_size=1536, num_hidden_layers=2,
                     num_attention_heads=8, num_key_value_heads=2, rms_norm_eps=1e-5),
    "llama1-7b": dict(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=32,
                      num_attention_heads=32, rms_norm_eps=1e-6),
    "llama2-7b": dict(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=32,
                      num_attention_heads=32, rms_norm_eps=1e-5, max_sequence_length=4096),
    "llama2-13b": dict(vocab_size=32000, hidden_size=5120, intermediate_size=13824, num_hidden_layers=40,
                       num_attention_heads=40, rms_norm_eps=1e-5, max_sequence_length=4096),
    "llama1-33b": dict(vocab_size=32000, hidden_size=6656, intermediate_size=17920, num_hidden_layers=60,
                       num_attention_heads=52, rms_norm_eps=1e-6),
    "llama1-65b": dict(vocab_size=32000, hidden_size=8192, intermediate_size=22016, num_hidden_layers=80,
                       num_attention_heads=64, rms_norm_eps=1e-5),
    "llama2-70b": dict(vocab_size=32000, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                       num_attention_heads=64, num_key_value_heads=8, rms_norm_eps=1e-5,
                       max_sequence_length=4096),
    "llama3-8b": dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                      num_attention_heads=32, num_key_value_heads=8, rms_norm_eps=1e-5,
                      rope_theta=500000.0, max_sequence_length=8192, bos_token_id=128000,
                      eos_token_id=128001),
    # 1B-class LLaMA-architecture shape with 64-wide heads and tied embeddings (the architecture of Llama-3.2-1B; its
    # "llama3" RoPE frequency scaling is not applied — the reference has no RoPE scaling either)
    "llama3.2-1b": dict(vocab_size=128256, hidden_size=2048, intermediate_size=8192, num_hidden_layers=16,
                        num_attention_heads=32, num_key_value_heads=8, rms_norm_eps=1e-5, rope_theta=500000.0,
                        max_sequence_length=8192, bos_token_id=128000, eos_token_id=128001,
                        tie_word_embeddings=True),
    "llama3-70b": dict(vocab_size=128256, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                       num_attention_heads=64, num_key_value_heads=8, rms_norm_eps=1e-5,
                       rope_theta=500000.0, max_sequence_length=8192, bos_token_id=128000,
                       eos_token_id=128001),
}


def preset(name: str, **overrides: Any) -> LLaMAConfig:
    if name not in PRESETS:
        raise KeyError(f"unknown preset {name!r}; choose from {sorted(PRESETS)}")
    kw = dict(PRESETS[name])
    kw.update(overrides)
    return LLaMAConfig(**kw)
