#!/usr/bin/env python
"""Entry point with the reference's file name (its `jax_test.py`): runs the parity harness `parity_test.py`
(unit parity x N trials, tokenizer, model logits, greedy generations)."""
import runpy
import sys
from pathlib import Path

if __name__ == "__main__":
    sys.argv[0] = str(Path(__file__).with_name("parity_test.py"))
    runpy.run_path(sys.argv[0], run_name="__main__")
