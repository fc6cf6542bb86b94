"""Generation front-end: `LLaMA(params, model, tokenizer, mesh).generate / generate_from_str`.

Same call signatures and post-processing as /root/reference/jax_llama/generation.py:15-79: prompts
are encoded with BOS, left-padded with ``eos_id`` to the longest prompt, the attention mask is
``tokens != eos_id``, sampling is on iff ``temperature != 0``, and the returned strings include the
prompt, start at the first BOS and stop before the first EOS.  Added knobs (`top_k`, `seed`) default
to the behaviour the reference had through its HF dependency.
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence

import torch

from .models.llama import LLaMAForCausalLM
from .parallel.layout import Mesh, P, with_named_sharding_constraint

logger = logging.getLogger(__name__)


def _leaves(tree: Dict[str, Any]):
    for v in tree.values():
        if isinstance(v, dict):
            yield from _leaves(v)
        else:
            yield v


@dataclass
class LLaMA:
    params: Dict[str, Any]
    model: LLaMAForCausalLM
    tokenizer: Any
    mesh: Optional[Mesh] = None

    def replace(self, **updates: Any) -> "LLaMA":
        """A copy with some fields swapped (what `flax.struct.PyTreeNode.replace` gave the reference's LLaMA)."""
        import dataclasses
        return dataclasses.replace(self, **updates)

    def generate(self, tokens: torch.Tensor, attention_mask: torch.Tensor, max_gen_len: int,
                 temperature: float = 0.8, top_p: float = 0.95, top_k: Optional[int] = 50,
                 seed: int = 0) -> torch.Tensor:
        """[B, P] prompt ids (+ mask) → [B, P + max_gen_len] ids, padded with ``eos_id``."""
        tokens = with_named_sharding_constraint(torch.as_tensor(tokens), self.mesh, P("dp", None))
        attention_mask = with_named_sharding_constraint(torch.as_tensor(attention_mask), self.mesh, P("dp", None))
        logger.debug("generate: tokens %s", tuple(tokens.shape))
        out = self.model.generate(
            input_ids=tokens, attention_mask=attention_mask, params=self.params,
            max_length=max_gen_len + tokens.shape[1], do_sample=temperature != 0.0,
            temperature=temperature if temperature != 0.0 else 1.0, top_k=top_k, top_p=top_p,
            pad_token_id=self.tokenizer.eos_id, eos_token_id=self.tokenizer.eos_id, seed=seed)
        return with_named_sharding_constraint(out.sequences, self.mesh, P("dp", None))

    def last_generation_stats(self) -> Optional[Any]:
        """`GenerationStats` of the most recent `generate` call on the B200 engine (prefill ms, ms per token, tokens/s,
        fraction of the HBM roofline when a peak is set); None on the reference backend."""
        dev = next(iter(_leaves(self.params))).device if isinstance(self.params, dict) else None
        if dev is None or getattr(dev, "type", "cpu") != "cuda":
            return None
        return getattr(self.model._engine_for(self.params), "last_stats", None)

    def stream(self, tokens: torch.Tensor, attention_mask: torch.Tensor, max_gen_len: int,
               temperature: float = 0.8, top_p: float = 0.95, top_k: Optional[int] = 50, seed: int = 0) -> Any:
        """Token-at-a-time variant of `generate` for interactive serving (CUDA parameters only).

        Returns a handle whose ``step(feed=None)`` produces one token per row and hands it back in pinned
        host memory; ``last_tokens()`` is the first generated token right after the prefill."""
        dev = next(iter(_leaves(self.params))).device
        if dev.type != "cuda":
            raise NotImplementedError("LLaMA.stream needs parameters on a CUDA device")
        eng = self.model._engine_for(self.params)
        return eng.begin_stream(torch.as_tensor(tokens).to(dev), torch.as_tensor(attention_mask).to(dev),
                                max_length=max_gen_len + tokens.shape[1], do_sample=temperature != 0.0,
                                temperature=temperature if temperature != 0.0 else 1.0, top_k=top_k, top_p=top_p,
                                pad_token_id=self.tokenizer.eos_id, eos_token_id=self.tokenizer.eos_id, seed=seed)

    def encode_prompts(self, prompts: Sequence[str]) -> tuple:
        """Tokenise with BOS and left-pad with eos_id; returns (tokens [B,P] int64, mask, per-row ids)."""
        rows = [self.tokenizer.encode(p, bos=True, eos=False) for p in prompts]
        width = max(len(r) for r in rows)
        tokens = torch.full((len(rows), width), self.tokenizer.eos_id, dtype=torch.long)
        for i, r in enumerate(rows):
            tokens[i, width - len(r):] = torch.tensor(r, dtype=torch.long)
        mask = (tokens != self.tokenizer.eos_id).to(torch.int32)
        return tokens, mask, rows

    def generate_from_str(self, prompts: List[str], max_gen_len: int, temperature: float = 0.8,
                          top_p: float = 0.95, top_k: Optional[int] = 50, seed: int = 0) -> List[str]:
        tokens, mask, rows = self.encode_prompts(prompts)
        out = self.generate(tokens, mask, max_gen_len, temperature, top_p, top_k=top_k, seed=seed)
        texts = []
        for i, ids in enumerate(out.tolist()):
            ids = ids[ids.index(self.tokenizer.bos_id):]          # drop the left padding
            ids = ids[:len(rows[i]) + max_gen_len]
            if self.tokenizer.eos_id in ids:
                ids = ids[:ids.index(self.tokenizer.eos_id)]
            texts.append(self.tokenizer.decode(ids))
        return texts
