"""Times jl_flash_prefill (CUDA events, warm, L2 flushed between iterations) on LLaMA-3 8B / 70B-per-rank shapes.
JLLAMA_FLASH_V1=1 selects the first-generation kernel for an A/B on the same box."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jax_llama_b200.ops import kernels as K  # noqa: E402

dev = torch.device("cuda")
from jax_llama_b200.ops import _lib  # noqa: E402
print("resident CTAs/SM:", {n: _lib.lib().jl_flash_prefill_occupancy(n) for n in (-1, 0, 32768, 65536, 81920, 100352)})
pr = torch.cuda.get_device_properties(0)
print("smem/SM", pr.shared_memory_per_multiprocessor, "regs/SM", pr.regs_per_multiprocessor, "max threads/SM", pr.max_threads_per_multi_processor)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
SHAPES = [(1, 2048, 32, 8), (1, 8192, 32, 8), (4, 2048, 8, 1), (1, 512, 32, 8)]
if os.environ.get("FLASH_SHAPES"):          # e.g. FLASH_SHAPES="1,8192,32,8;1,2048,32,8"
    SHAPES = [tuple(int(v) for v in t.split(",")) for t in os.environ["FLASH_SHAPES"].split(";")]
ITERS = int(os.environ.get("FLASH_ITERS", "10"))
for (B, S, nh, nkv) in SHAPES:
    g = torch.Generator(device="cpu").manual_seed(0)
    q = torch.randn(B, S, nh, 128, generator=g).to(torch.bfloat16).to(dev)
    kc = torch.randn(B, nkv, S, 128, generator=g).to(torch.bfloat16).to(dev)
    vc = torch.randn(B, nkv, S, 128, generator=g).to(torch.bfloat16).to(dev)
    pad = torch.zeros(B, dtype=torch.int32, device=dev)
    for _ in range(3):
        K.prefill_attn(q, kc, vc, 0, S, pad)
    ts = []
    for _ in range(ITERS):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        K.prefill_attn(q, kc, vc, 0, S, pad)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    ms = ts[len(ts) // 2]
    flops = 4.0 * B * nh * 128 * (S * (S + 1) / 2)
    print(f"flash B={B} S={S} nh={nh} nkv={nkv}: {ms * 1e3:8.1f} us  {flops / ms / 1e9:7.1f} TFLOP/s (causal)"
          f"  v1={os.environ.get('JLLAMA_FLASH_V1', '0')} psmem={os.environ.get('JLLAMA_FLASH_PSMEM', '0')}")
    # the warp-level tensor-core kernel (the path of 64-wide heads) on the same shape, and at head_dim 64
    for hd in (128, 64):
        q2, k2, v2 = (t[..., :hd].contiguous() for t in (q, kc, vc))
        for _ in range(2):
            K.prefill_attn(q2, k2, v2, 0, S, pad, force="mma")
        ts = []
        for _ in range(max(ITERS // 2, 3)):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            K.prefill_attn(q2, k2, v2, 0, S, pad, force="mma")
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        ms2 = ts[len(ts) // 2]
        print(f"   warp-mma head_dim {hd}: {ms2 * 1e3:8.1f} us  {flops * hd / 128 / ms2 / 1e9:7.1f} TFLOP/s")
