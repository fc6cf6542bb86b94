#!/usr/bin/env python
"""Entry point with the reference's file name (its `jax_example.py`): runs `examples/example.py` with the same
flags (--ckpt_dir, --tokenizer_path, --is_llama3, --max_gen_len, --temperature, --top_p)."""
import runpy
import sys
from pathlib import Path

if __name__ == "__main__":
    sys.argv[0] = str(Path(__file__).with_name("examples") / "example.py")
    runpy.run_path(sys.argv[0], run_name="__main__")
