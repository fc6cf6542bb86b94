#!/usr/bin/env python
"""Small eager workloads for ncu (`--mode decode|prefill|gemm`): every kernel is an ordinary launch
(no CUDA graph) so the profiler lists them one by one."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from jax_llama_b200 import preset  # noqa: E402
from jax_llama_b200.models.engine import B200Engine  # noqa: E402
from jax_llama_b200.ops import kernels as K  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="decode")
ap.add_argument("--model", default="llama3-8b")
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--prompt", type=int, default=2048)
ap.add_argument("--steps", type=int, default=2)
a = ap.parse_args()
torch.cuda.set_device(0)
cfg = preset(a.model, num_hidden_layers=a.layers)
if a.mode == "gemm":
    for (m, n, k) in [(2048, 6144, 4096), (2048, 4096, 4096), (2048, 28672, 4096), (2048, 4096, 14336),
                      (64, 28672, 4096), (64, 4096, 14336), (16, 128256, 4096)]:
        x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
        w = torch.randn(n, k, device="cuda", dtype=torch.bfloat16)
        for _ in range(2):
            K.linear(x, w, force="gemm")
    torch.cuda.synchronize()
    sys.exit(0)
eng = B200Engine.random_init(cfg, "cuda:0", use_cuda_graph=False)
ids = torch.randint(0, cfg.vocab_size, (a.batch, a.prompt), device="cuda")
if a.mode == "prefill":
    st = eng._decode_state(a.batch, a.prompt + 8)
    for _ in range(a.steps):
        eng.prepare(st, ids, None, 0)
else:
    seqs = eng.generate(ids, None, max_length=a.prompt + 1 + a.steps, eos_token_id=-1)
torch.cuda.synchronize()
print("done")
