"""LLaMA-3 tokenizer (tiktoken BPE + 256 reserved specials) and chat prompt format.

Capability parity with /root/reference/jax_llama/llama3_tokenizer.py:38-232.  The special-token
table, its order and the split regex are part of Meta's published LLaMA-3 vocabulary definition (an
interchange format), so they are reproduced as data; the code around them is ours.
"""
from __future__ import annotations

import logging
import os
from pathlib import Path
from typing import AbstractSet, Collection, Dict, Iterator, List, Literal, Sequence, TypedDict, Union

import tiktoken
from tiktoken.load import load_tiktoken_bpe

logger = logging.getLogger(__name__)

Role = Literal["system", "user", "assistant"]


class Message(TypedDict):
    role: Role
    content: str


Dialog = Sequence[Message]

NUM_RESERVED_SPECIAL_TOKENS = 256
# tiktoken hard limits that the encoder must respect on pathological inputs
_MAX_ENCODE_CHARS = 400_000
_MAX_SAME_CLASS_RUN = 25_000

_SPLIT_REGEX = (
    r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*"
    r"|\s*[\r\n]+|\s+(?!\S)|\s+"
)


def _special_token_names() -> List[str]:
    named = {0: "<|begin_of_text|>", 1: "<|end_of_text|>", 6: "<|start_header_id|>",
             7: "<|end_header_id|>", 9: "<|eot_id|>"}
    names, reserved = [], 0
    for slot in range(NUM_RESERVED_SPECIAL_TOKENS):
        if slot in named:
            names.append(named[slot])
        else:
            names.append(f"<|reserved_special_token_{reserved}|>")
            reserved += 1
    return names


def split_long_runs(s: str, max_run: int) -> Iterator[str]:
    """Yield pieces of `s` in which no run of all-space or all-non-space characters exceeds `max_run`."""
    if not s:
        yield s
        return
    start, run, run_is_space = 0, 0, s[0].isspace()
    for i, ch in enumerate(s):
        is_space = ch.isspace()
        if is_space != run_is_space:
            run, run_is_space = 1, is_space
            continue
        run += 1
        if run > max_run:
            yield s[start:i]
            start, run = i, 1
    yield s[start:]


class Tokenizer:
    special_tokens: Dict[str, int]
    num_reserved_special_tokens = NUM_RESERVED_SPECIAL_TOKENS
    pat_str = _SPLIT_REGEX

    def __init__(self, model_path: str):
        if not os.path.isfile(model_path):
            raise FileNotFoundError(model_path)
        ranks = load_tiktoken_bpe(model_path)
        base = len(ranks)
        self.special_tokens = {name: base + i for i, name in enumerate(_special_token_names())}
        self.model = tiktoken.Encoding(name=Path(model_path).name, pat_str=self.pat_str,
                                       mergeable_ranks=ranks, special_tokens=self.special_tokens)
        self.n_words: int = self.model.n_vocab
        self.bos_id: int = self.special_tokens["<|begin_of_text|>"]
        self.eos_id: int = self.special_tokens["<|end_of_text|>"]
        self.pad_id: int = -1
        self.stop_tokens = {self.eos_id, self.special_tokens["<|eot_id|>"]}
        logger.info("tiktoken model %s: %d words, bos=%d eos=%d", model_path, self.n_words,
                    self.bos_id, self.eos_id)

    def encode(self, s: str, *, bos: bool, eos: bool,
               allowed_special: Union[Literal["all"], AbstractSet[str]] = frozenset(),
               disallowed_special: Union[Literal["all"], Collection[str]] = ()) -> List[int]:
        """Special-token text is encoded as ordinary text unless listed in `allowed_special`."""
        if not isinstance(s, str):
            raise TypeError("encode() expects a str")
        ids: List[int] = []
        for lo in range(0, len(s), _MAX_ENCODE_CHARS):
            for piece in split_long_runs(s[lo:lo + _MAX_ENCODE_CHARS], _MAX_SAME_CLASS_RUN):
                ids.extend(self.model.encode(piece, allowed_special=allowed_special,
                                             disallowed_special=disallowed_special))
        if bos:
            ids.insert(0, self.bos_id)
        if eos:
            ids.append(self.eos_id)
        return ids

    def decode(self, t: Sequence[int]) -> str:
        return self.model.decode(list(t))

    def __len__(self) -> int:
        return self.n_words

    # kept for callers that used the reference's private helper name
    _split_whitespaces_or_nonwhitespaces = staticmethod(split_long_runs)


class ChatFormat:
    """LLaMA-3 instruct prompt layout: header / message / dialog-with-open-assistant-turn."""

    def __init__(self, tokenizer: Tokenizer):
        self.tokenizer = tokenizer

    def _sp(self, name: str) -> int:
        return self.tokenizer.special_tokens[name]

    def encode_header(self, message: Message) -> List[int]:
        enc = self.tokenizer.encode
        return ([self._sp("<|start_header_id|>")] + enc(message["role"], bos=False, eos=False)
                + [self._sp("<|end_header_id|>")] + enc("\n\n", bos=False, eos=False))

    def encode_message(self, message: Message) -> List[int]:
        body = self.tokenizer.encode(message["content"].strip(), bos=False, eos=False)
        return self.encode_header(message) + body + [self._sp("<|eot_id|>")]

    def encode_dialog_prompt(self, dialog: Dialog) -> List[int]:
        tokens = [self._sp("<|begin_of_text|>")]
        for message in dialog:
            tokens.extend(self.encode_message(message))
        # leave an open assistant header for the model to complete
        tokens.extend(self.encode_header({"role": "assistant", "content": ""}))
        return tokens
