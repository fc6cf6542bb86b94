"""CUDA-event timing of the prefill GEMM shapes: CTA-pair kernel (cta_group::2) vs the single-CTA kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_llama_b200.ops import kernels as K

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, reps=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sorted(ts)[len(ts) // 2]

M = int(os.environ.get("M", "2048"))
for name, n, k, mode in [("wqkv", 6144, 4096, K.PLAIN), ("wo", 4096, 4096, K.PLAIN), ("w13", 28672, 4096, K.SWIGLU),
                         ("w2", 4096, 14336, K.PLAIN)]:
    x = torch.randn(M, k, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(n, k, device="cuda") * 0.02).to(torch.bfloat16)
    fl = 2.0 * M * n * k
    for path in ("gemm_2cta", "gemm"):
        if path == "gemm":
            os.environ["JLLAMA_GEMM_2CTA"] = "0"
        best, med = timeit(lambda: K.linear(x, w, mode=mode, force=path))
        print(f"{name:5s} M={M} N={n} K={k} {path:10s} best {best*1e3:8.1f} us  median {med*1e3:8.1f} us  {fl/best/1e9:7.1f} TFLOP/s (best)")
from jax_llama_b200.ops import _lib
print("resident 2-CTA clusters:", _lib.lib().jl_gemm_2cta_pairs())
