"""Host-side tokenizers (LLaMA-1/2 sentencepiece, LLaMA-3 tiktoken + chat format)."""
from .llama2 import Tokenizer as LLaMA2Tokenizer
from .llama3 import ChatFormat, Dialog, Message, Role
from .llama3 import Tokenizer as LLaMA3Tokenizer

__all__ = ["LLaMA2Tokenizer", "LLaMA3Tokenizer", "ChatFormat", "Dialog", "Message", "Role"]
