"""HF-style model classes: `LLaMAModel`, `LLaMAForCausalLM` (+ `FlaxLLaMA*` aliases).

API surface of /root/reference/jax_llama/model.py:402-546,680-772: ctor ``(config, input_shape, seed,
dtype, _do_init)``, ``__call__(input_ids, attention_mask, position_ids, params, past_key_values, …,
output_attentions, output_hidden_states, return_dict)``, ``init_cache``, the two generation hooks,
``generate`` (inherited from HF in the reference; implemented here), ``save_pretrained`` /
``from_pretrained``.  There is no tracing compiler underneath: the call dispatches to

* the *reference backend* (`models/reference.py`, plain PyTorch; CPU, fp32 oracle, attention maps), or
* the *B200 engine* (`models/engine.py`: hand-written sm_100a kernels, CUDA-graph decode) when the
  parameters live on a CUDA device — missing native code is an error there, never a silent fallback.
"""
from __future__ import annotations

import logging
import os
from dataclasses import dataclass
import weakref
from typing import Any, Dict, Optional, Tuple, Union

import torch

from ..config import LLaMAConfig
from ..ops.sampling import sample
from . import reference as ref
from .modules import LLaMAForCausalLMModule, LLaMAModule
from .reference import KVCache, ModelOutput

logger = logging.getLogger(__name__)

Tree = Dict[str, Any]
WEIGHTS_NAME = "model.safetensors"


@dataclass
class GenerateOutput:
    sequences: torch.Tensor


def _flatten(tree: Tree, prefix: str = "") -> Dict[str, torch.Tensor]:
    out: Dict[str, torch.Tensor] = {}
    for k, v in tree.items():
        key = f"{prefix}/{k}" if prefix else str(k)
        if isinstance(v, dict):
            out.update(_flatten(v, key))
        else:
            out[key] = v
    return out


def _unflatten(flat: Dict[str, torch.Tensor]) -> Tree:
    tree: Tree = {}
    for key, v in flat.items():
        node = tree
        parts = key.split("/")
        for p in parts[:-1]:
            node = node.setdefault(p, {})
        node[parts[-1]] = v
    return tree


def tree_to(tree: Tree, *, device: Any = None, dtype: Optional[torch.dtype] = None) -> Tree:
    return _unflatten({k: v.to(device=device, dtype=dtype) for k, v in _flatten(tree).items()})


def _tree_device(tree: Tree) -> torch.device:
    return next(iter(_flatten(tree).values())).device


class LLaMAPreTrainedModel:
    config_class = LLaMAConfig
    base_model_prefix = "transformer"
    with_head = True

    def __init__(self, config: LLaMAConfig, input_shape: Tuple[int, int] = (1, 1), seed: int = 0,
                 dtype: torch.dtype = torch.float32, _do_init: bool = True, device: Any = "cpu",
                 **kwargs: Any):
        self.config = config
        self.dtype = dtype
        self.seed = seed
        self.input_shape = input_shape
        self._params: Optional[Tree] = None
        self._engines: Dict[int, Tuple[Any, Any, Any]] = {}      # id(tree) → (weakref(leaf), version stamp, engine)
        self.runtime_kwargs = kwargs            # tp / max_batch / kernel toggles, forwarded to the engine
        if _do_init:
            self._params = self.init_weights(seed, input_shape, device=device)

    # ------------------------------------------------------------------ parameters
    def init_weights(self, rng: int = 0, input_shape: Tuple[int, int] = (1, 1), params: Optional[Tree] = None,
                     device: Any = "cpu") -> Tree:
        fresh = ref.init_params(self.config, seed=int(rng), dtype=self.dtype, device=device)
        if params is not None:                  # fill in only what is missing, like the HF wrapper
            have = _flatten(params)
            fresh = _unflatten({k: have.get(k, v) for k, v in _flatten(fresh).items()})
        return fresh

    @property
    def params(self) -> Tree:
        if self._params is None:
            raise ValueError("model was built with _do_init=False: pass `params=` explicitly")
        return self._params

    @params.setter
    def params(self, value: Tree) -> None:
        self._params = value

    # ------------------------------------------------------------------ backends
    def _engine_for(self, params: Tree) -> Any:
        """B200 engine bound to `params` (built once per parameter tree).

        The cache entry is validated against the tree it was built from: a weak reference to the tree's
        embedding tensor plus the in-place version counters of all leaves.  A tree that was freed (its id may be
        reused by a new dict) or modified in place therefore gets a fresh engine, and entries of dead trees are
        dropped.  A dict of the form ``{"engine": B200Engine}`` is a handle to an engine built elsewhere
        (``B200Engine.random_init`` in the benchmarks)."""
        pre = params.get("engine") if isinstance(params, dict) else None
        if pre is not None and not isinstance(pre, (dict, torch.Tensor)):
            return pre
        leaves = list(_flatten(params).values())
        stamp = (len(leaves), tuple(int(t._version) for t in leaves))
        for k in [k for k, (ref_, _, _) in self._engines.items() if ref_() is None]:
            del self._engines[k]                                   # trees that no longer exist
        key = id(params)
        hit = self._engines.get(key)
        if hit is not None and hit[0]() is leaves[0] and hit[1] == stamp:
            return hit[2]
        from .engine import B200Engine
        eng = B200Engine.from_tree(self.config, params, **self.runtime_kwargs)
        self._engines[key] = (weakref.ref(leaves[0]), stamp, eng)
        return eng

    def _use_engine(self, params: Tree, output_attentions: bool, output_hidden_states: bool) -> bool:
        if _tree_device(params).type != "cuda":
            return False
        if output_attentions or output_hidden_states:
            # flash kernels never materialise [B, nh, Q, K]; introspection requests take the unfused path
            return False
        return True

    # ------------------------------------------------------------------ cache + forward
    def init_cache(self, batch_size: int, max_length: int, params: Optional[Tree] = None) -> KVCache:
        """Allocate an empty KV cache directly (no throw-away model init as in model.py:459-476)."""
        params = params if params is not None else self._params
        if params is not None and _tree_device(params).type == "cuda":
            return self._engine_for(params).new_cache(batch_size, max_length)
        dtype = next(iter(_flatten(params).values())).dtype if params is not None else self.dtype
        dev = _tree_device(params) if params is not None else "cpu"
        return KVCache.allocate(self.config, batch_size, max_length, dtype, dev)

    def __call__(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                 position_ids: Optional[torch.Tensor] = None, params: Optional[Tree] = None,
                 past_key_values: Optional[KVCache] = None, dropout_rng: Any = None, train: bool = False,
                 output_attentions: Optional[bool] = None, output_hidden_states: Optional[bool] = None,
                 return_dict: Optional[bool] = True, last_only: bool = False
                 ) -> Union[ModelOutput, Tuple[Any, ...]]:
        if train:
            raise NotImplementedError("inference-only framework: train=True is not supported")
        params = params if params is not None else self.params
        output_attentions = bool(output_attentions)
        output_hidden_states = bool(output_hidden_states)
        input_ids = torch.as_tensor(input_ids)
        if input_ids.ndim != 2:
            raise ValueError("input_ids must be [batch, sequence]")
        if past_key_values is not None and past_key_values.index > 0 and position_ids is None:
            raise ValueError("Make sure to provide `position_ids` when passing `past_key_values`.")
        dev = _tree_device(params)
        input_ids = input_ids.to(dev)
        if attention_mask is not None:
            attention_mask = torch.as_tensor(attention_mask).to(dev)
        if position_ids is not None:
            position_ids = torch.as_tensor(position_ids).to(dev)

        if self._use_engine(params, output_attentions, output_hidden_states):
            out = self._engine_for(params).forward(
                input_ids, attention_mask=attention_mask, position_ids=position_ids,
                cache=past_key_values, with_head=self.with_head, last_only=last_only)
        else:
            with torch.no_grad():
                out = ref.forward(params, self.config, input_ids, attention_mask, position_ids,
                                  past_key_values, output_attentions, output_hidden_states,
                                  with_head=self.with_head, last_only=last_only)
        if not self.with_head:
            out.logits = None
        return out if return_dict else out.to_tuple()

    # ------------------------------------------------------------------ checkpoint I/O
    def save_pretrained(self, save_directory: str, params: Optional[Tree] = None) -> None:
        from safetensors.torch import save_file

        params = params if params is not None else self.params
        os.makedirs(save_directory, exist_ok=True)
        self.config.save_pretrained(save_directory)
        flat = {k: v.detach().cpu().contiguous() for k, v in _flatten(params).items()}
        save_file(flat, os.path.join(save_directory, WEIGHTS_NAME), metadata={"format": "pt", "layout": "canonical-in-out"})

    @classmethod
    def from_pretrained(cls, path: str, dtype: Optional[torch.dtype] = None, device: Any = "cpu",
                        _do_init: bool = True, **kwargs: Any) -> Any:
        from safetensors.torch import load_file

        config = LLaMAConfig.from_pretrained(path)
        flat = load_file(os.path.join(path, WEIGHTS_NAME), device=str(device))
        if dtype is not None:
            flat = {k: v.to(dtype) for k, v in flat.items()}
        params = _unflatten(flat)
        any_leaf = next(iter(flat.values()))
        model = cls(config, dtype=any_leaf.dtype, _do_init=False, **kwargs)
        if _do_init:
            model.params = params
            return model
        return model, params


class LLaMAModel(LLaMAPreTrainedModel):
    """Backbone without the LM head: returns `last_hidden_state` (model.py:680-681)."""

    with_head = False
    module_class = LLaMAModule            # the reference's `module_class = FlaxLLaMAModule`


class LLaMAForCausalLM(LLaMAPreTrainedModel):
    """Backbone + LM head + generation (model.py:745-772)."""

    with_head = True
    module_class = LLaMAForCausalLMModule

    # ------------------------------------------------------------------ generation hooks
    def prepare_inputs_for_generation(self, input_ids: torch.Tensor, max_length: int,
                                      attention_mask: Optional[torch.Tensor] = None,
                                      params: Optional[Tree] = None) -> Dict[str, Any]:
        """Cache for `max_length`, static key mask [B, max_length] (prompt mask then ones), positions."""
        b, s = input_ids.shape
        cache = self.init_cache(b, max_length, params)
        dev = input_ids.device
        ext = torch.ones((b, max_length), dtype=torch.int32, device=dev)
        if attention_mask is not None:
            position_ids = ref.default_position_ids(attention_mask, b, s, dev)
            ext[:, :s] = attention_mask.to(torch.int32)
        else:
            position_ids = torch.arange(s, device=dev)[None].expand(b, s)
        return {"past_key_values": cache, "attention_mask": ext, "position_ids": position_ids}

    def update_inputs_for_generation(self, model_outputs: ModelOutput, model_kwargs: Dict[str, Any]
                                     ) -> Dict[str, Any]:
        model_kwargs["past_key_values"] = model_outputs.past_key_values
        model_kwargs["position_ids"] = model_kwargs["position_ids"][:, -1:] + 1
        return model_kwargs

    # ------------------------------------------------------------------ decode loop
    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                 params: Optional[Tree] = None, max_length: Optional[int] = None,
                 max_new_tokens: Optional[int] = None, do_sample: bool = False, temperature: float = 1.0,
                 top_k: Optional[int] = 50, top_p: float = 1.0, pad_token_id: Optional[int] = None,
                 eos_token_id: Optional[int] = None, seed: int = 0, generation_config: Any = None
                 ) -> GenerateOutput:
        """Greedy / sampling decode with a KV cache.

        Semantics of HF's Flax ``_greedy_search`` / ``_sample`` that the reference inherits
        (SURVEY §3.2): the prompt is prefilled in one call; each step takes ``logits[:, -1]``; rows that
        already produced EOS keep emitting ``pad``; the loop ends at ``max_length`` or when every row has
        finished; `sequences` is ``[B, max_length]`` padded with ``pad``.  A fixed `seed` makes sampling
        repeatable per call (the reference always used PRNGKey(0)).
        """
        if generation_config is not None:
            g = generation_config
            max_length = getattr(g, "max_length", max_length)
            do_sample = getattr(g, "do_sample", do_sample)
            temperature = getattr(g, "temperature", temperature) or 1.0
            top_p = getattr(g, "top_p", top_p) or 1.0
            top_k = getattr(g, "top_k", top_k)
            pad_token_id = getattr(g, "pad_token_id", pad_token_id)
            eos_token_id = getattr(g, "eos_token_id", eos_token_id)
        params = params if params is not None else self.params
        dev = _tree_device(params)
        input_ids = torch.as_tensor(input_ids).to(dev)
        b, s = input_ids.shape
        if max_length is None:
            if max_new_tokens is None:
                raise ValueError("give max_length or max_new_tokens")
            max_length = s + max_new_tokens
        if max_length < s:
            raise ValueError("max_length is shorter than the prompt")
        eos = self.config.eos_token_id if eos_token_id is None else eos_token_id
        pad = eos if pad_token_id is None else pad_token_id
        if attention_mask is not None:
            attention_mask = torch.as_tensor(attention_mask).to(dev)

        if dev.type == "cuda":
            seqs = self._engine_for(params).generate(
                input_ids, attention_mask, max_length=max_length, do_sample=do_sample,
                temperature=temperature, top_k=top_k, top_p=top_p, pad_token_id=pad,
                eos_token_id=eos, seed=seed)
            return GenerateOutput(sequences=seqs)

        gen = torch.Generator(device="cpu").manual_seed(seed)
        sequences = torch.full((b, max_length), pad, dtype=torch.long, device=dev)
        sequences[:, :s] = input_ids
        finished = torch.zeros(b, dtype=torch.bool, device=dev)
        kwargs = self.prepare_inputs_for_generation(input_ids, max_length, attention_mask, params)
        running = input_ids
        cur = s
        while cur < max_length and not bool(finished.all()):
            out = self(running, params=params, last_only=True, **kwargs)
            logits = out.logits[:, -1].float()
            nxt = sample(logits, temperature if do_sample else 0.0, top_k, top_p, gen)
            nxt = torch.where(finished, torch.full_like(nxt, pad), nxt)
            finished = finished | (nxt == eos)
            sequences[:, cur] = nxt
            running = nxt[:, None]
            kwargs = self.update_inputs_for_generation(out, kwargs)
            cur += 1
        return GenerateOutput(sequences=sequences)


# Names the reference exports (jax_llama/__init__.py:4): kept so existing user code keeps importing.
FlaxLLaMAForCausalLM = LLaMAForCausalLM
FlaxLLaMAModel = LLaMAModel
FlaxLLaMAPreTrainedModel = LLaMAPreTrainedModel
