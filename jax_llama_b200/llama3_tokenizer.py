"""Same-named module as the reference's `jax_llama/llama3_tokenizer.py` (see `tokenizers/llama3.py`)."""
from .tokenizers.llama3 import ChatFormat, Dialog, Message, Role, Tokenizer  # noqa: F401
