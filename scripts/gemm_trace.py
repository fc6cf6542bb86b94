"""Timeline of the swap-AB (decode) GEMM: where the microseconds of a 20 us kernel go.

Every CTA stamps %globaltimer at ten points (csrc/gemm_tcgen05.cu `gemm_stamp`); this prints, per shape, the median /
max over CTAs of each point relative to the first CTA's entry, once for an isolated launch and once for the second
of two back-to-back launches (programmatic dependent launch: its set-up and weight prefetch overlap the first)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_llama_b200.ops import _lib, kernels as K

NAMES = ["entry", "setup done", "producer past wait", "first stage landed", "last MMA issued", "accumulator complete",
         "partial published", "all splits arrived", "share emitted", "exit"]
dev = torch.device("cuda")
L = _lib.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
M = int(os.environ.get("M", "32"))


def show(tr: torch.Tensor, label: str, t0=None) -> None:
    t = tr.cpu().numpy().astype("int64")
    used = t[:, 0] > 0
    t = t[used]
    base = t[:, 0].min() if t0 is None else t0
    print(f"  {label}: {len(t)} CTAs; microseconds after {'the first entry' if t0 is None else 'the FIRST kernel entry'}")
    for i, n in enumerate(NAMES):
        col = t[:, i]
        col = col[col > 0]
        if len(col) == 0:
            continue
        rel = (col - base) / 1e3
        print(f"    {n:22s} min {rel.min():7.2f}  p50 {float(sorted(rel)[len(rel) // 2]):7.2f}  max {rel.max():7.2f}")


for name, n, k, mode, res in [("wqkv", 6144, 4096, K.PLAIN, False), ("wo+res", 4096, 4096, K.RESIDUAL, True),
                              ("w13", 28672, 4096, K.SWIGLU, False), ("w2+res", 4096, 14336, K.RESIDUAL, True)]:
    x = torch.randn(M, k, device=dev).bfloat16()
    w = (torch.randn(n, k, device=dev) * 0.02).bfloat16()
    r = torch.randn(M, n, device=dev).bfloat16() if res else None
    x2 = torch.randn(M, 4096, device=dev).bfloat16()
    w2 = (torch.randn(4096, 4096, device=dev) * 0.02).bfloat16()
    fn = lambda: K.linear(x, w, residual=r, mode=mode)
    fn(); torch.cuda.synchronize()
    print(f"--- {name}: M={M} N={n} K={k}")
    tr = torch.zeros(148, 16, dtype=torch.int64, device=dev)
    flush.max(); torch.cuda.synchronize()
    L.jl_gemm_set_trace(tr.data_ptr())
    fn()
    L.jl_gemm_set_trace(None)
    torch.cuda.synchronize()
    show(tr, "isolated")
    # second of two dependent launches (a plain 4096x4096 GEMM first, then this one), replayed from a CUDA graph so
    # that the host launch path is out of the picture
    tr1 = torch.zeros(148, 16, dtype=torch.int64, device=dev)
    tr2 = torch.zeros(148, 16, dtype=torch.int64, device=dev)
    out1 = torch.empty(M, 4096, device=dev, dtype=torch.bfloat16)
    out2 = torch.empty(M, n // 2 if mode == K.SWIGLU else n, device=dev, dtype=torch.bfloat16)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        K.linear(x2, w2, out=out1); K.linear(x, w, residual=r, mode=mode, out=out2)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        L.jl_gemm_set_trace(tr1.data_ptr())
        K.linear(x2, w2, out=out1)
        L.jl_gemm_set_trace(tr2.data_ptr())
        K.linear(x, w, residual=r, mode=mode, out=out2)
        L.jl_gemm_set_trace(None)
    flush.max(); torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    t1 = tr1.cpu().numpy().astype("int64")
    first_entry = t1[t1[:, 0] > 0][:, 0].min()
    print(f"  (preceding GEMM: entry 0.00, last share emitted {(t1[:, 8].max() - first_entry) / 1e3:.2f})")
    show(tr2, "second kernel of a graph-replayed pair", t0=first_entry)
