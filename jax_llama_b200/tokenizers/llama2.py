"""LLaMA-1/2 tokenizer: a thin shell over the sentencepiece C++ library.

Capability parity with /root/reference/jax_llama/llama2_tokenizer.py:14-71 (`encode(s, bos, eos)`,
`decode`, `len()`, `n_words/bos_id/eos_id/pad_id`).  Tokenisation is host-side and off the measured
path, so the native dependency is kept.
"""
from __future__ import annotations

import logging
import os
from typing import List, Sequence

from sentencepiece import SentencePieceProcessor

logger = logging.getLogger(__name__)


class Tokenizer:
    def __init__(self, model_path: str):
        if not os.path.isfile(model_path):
            raise FileNotFoundError(model_path)
        self.sp_model = SentencePieceProcessor(model_file=model_path)
        self.n_words: int = self.sp_model.vocab_size()
        self.bos_id: int = self.sp_model.bos_id()
        self.eos_id: int = self.sp_model.eos_id()
        self.pad_id: int = self.sp_model.pad_id()
        if self.n_words != self.sp_model.get_piece_size():
            raise ValueError("sentencepiece model reports inconsistent vocabulary sizes")
        logger.info("sentencepiece model %s: %d words, bos=%d eos=%d pad=%d",
                    model_path, self.n_words, self.bos_id, self.eos_id, self.pad_id)

    def encode(self, s: str, bos: bool, eos: bool) -> List[int]:
        if not isinstance(s, str):
            raise TypeError("encode() expects a str")
        ids: List[int] = self.sp_model.encode(s)
        if bos:
            ids = [self.bos_id] + ids
        if eos:
            ids = ids + [self.eos_id]
        return ids

    def decode(self, t: Sequence[int]) -> str:
        return self.sp_model.decode(list(t))

    def __len__(self) -> int:
        return self.n_words
