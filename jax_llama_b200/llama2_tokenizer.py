"""Same-named module as the reference's `jax_llama/llama2_tokenizer.py` (see `tokenizers/llama2.py`)."""
from .tokenizers.llama2 import Tokenizer  # noqa: F401
