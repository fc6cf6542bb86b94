"""CUDA-event timing of every kernel of one per-op decode layer (batch 9..128 path of the engine) at LLaMA-3 8B
shapes: where the step time of the batched decode goes, each op against its HBM floor (bytes it must read / the
measured copy bandwidth).  Before every timed launch L2 is flushed by READING a 256 MB buffer (a written buffer
would leave 126 MB of dirty lines whose write-back is then charged to the timed kernel: +20-25 us)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from jax_llama_b200.ops import kernels as K

dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
HBM = 6447.2e9
try:
    HBM = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                      "MEASURED_PEAKS.json")))["hbm_gbs"] * 1e9
except Exception:
    pass


def timeit(fn, reps=8):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.max()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


H, F, V, nh, nkv, hd = 4096, 14336, 128256, 32, 8, 128
T = int(os.environ.get("CTX", "2048"))
batches = [int(b) for b in os.environ.get("BATCHES", "16,32,64").split(",")]
wqkv = (torch.randn((nh + 2 * nkv) * hd, H, device=dev) * 0.02).bfloat16()
wo = (torch.randn(H, nh * hd, device=dev) * 0.02).bfloat16()
w13 = (torch.randn(2 * F, H, device=dev) * 0.02).bfloat16()
w2 = (torch.randn(H, F, device=dev) * 0.02).bfloat16()
lm = (torch.randn(V, H, device=dev) * 0.02).bfloat16()
g = torch.ones(H, device=dev, dtype=torch.bfloat16)
half = hd // 2
ang = torch.arange(T + 64, device=dev, dtype=torch.float32)[:, None] * (
    500000.0 ** (-torch.arange(half, device=dev, dtype=torch.float32) / half))[None]
cos, sin = ang.cos().contiguous(), ang.sin().contiguous()

if os.environ.get("ATTN_SWEEP") == "1":
    # attention alone, both kernels, over (batch, kv heads per rank, group) shapes of the benchmark configs
    for B, kv, G in [(16, 1, 8), (64, 1, 8), (9, 8, 4), (16, 8, 4), (32, 8, 4), (8, 32, 1), (32, 32, 1), (2, 8, 4), (4, 8, 4)]:
        q = torch.randn(B, (kv * G + 2 * kv) * hd, device=dev).bfloat16()
        kc = torch.randn(B, kv, T + 64, hd, device=dev).bfloat16()
        vc = torch.randn(B, kv, T + 64, hd, device=dev).bfloat16()
        clen = torch.full((1,), T, device=dev, dtype=torch.int32)
        pad = torch.zeros(B, device=dev, dtype=torch.int32)
        splits, scratch = K.decode_attn_fused_plan(B, kv * G, kv, hd, T + 64, dev)
        out = torch.empty(B, kv * G * hd, device=dev, dtype=torch.bfloat16)
        fn = lambda: K.decode_attn_fused(q, kc, vc, cos, sin, clen, pad, kv * G, kv, hd, splits, scratch, out=out)
        ts = []
        for v in (1, 2):
            K.set_decode_attn_variant(v)
            ts.append(timeit(fn) * 1e3)
        K.set_decode_attn_variant(0)
        by = 2 * B * kv * T * hd * 2
        print(f"attn B={B:3d} nkv={kv:2d} G={G}: simt {ts[0]:7.1f} us  mma {ts[1]:7.1f} us  floor {by / HBM * 1e6:6.1f} us  "
              f"(plan splits {splits})")
    sys.exit(0)

for B in batches:
    x = torch.randn(B, H, device=dev).bfloat16()
    kc = torch.randn(B, nkv, T + 64, hd, device=dev).bfloat16()
    vc = torch.randn(B, nkv, T + 64, hd, device=dev).bfloat16()
    clen = torch.full((1,), T, device=dev, dtype=torch.int32)
    pad = torch.zeros(B, device=dev, dtype=torch.int32)
    splits, scratch = K.decode_attn_fused_plan(B, nh, nkv, hd, T + 64, dev)
    qkv = K.linear(x, wqkv)
    attn = torch.empty(B, nh * hd, device=dev, dtype=torch.bfloat16)
    hmid = torch.randn(B, F, device=dev).bfloat16()
    rows = []
    rows.append(("rmsnorm", timeit(lambda: K.rmsnorm(x, g, 1e-5)), 2 * B * H * 2))
    rows.append(("wqkv", timeit(lambda: K.linear(x, wqkv)), wqkv.numel() * 2))
    attn_fn = lambda: K.decode_attn_fused(qkv, kc, vc, cos, sin, clen, pad, nh, nkv, hd, splits, scratch, out=attn)
    K.set_decode_attn_variant(1)
    t_simt = timeit(attn_fn)
    K.set_decode_attn_variant(0)
    rows.append(("attn_fused", timeit(attn_fn), 2 * B * nkv * T * hd * 2))
    print(f"  (attention, SIMT kernel pinned: {t_simt * 1e3:.1f} us)")
    rows.append(("wo+res", timeit(lambda: K.linear(attn, wo, residual=x)), wo.numel() * 2))
    rows.append(("w13 swiglu", timeit(lambda: K.linear(x, w13, mode=K.SWIGLU)), w13.numel() * 2))
    rows.append(("w2+res", timeit(lambda: K.linear(hmid, w2, residual=x)), w2.numel() * 2))
    rows.append(("lm_head f32", timeit(lambda: K.linear(x, lm, out_f32=True)), lm.numel() * 2))
    layer = sum(t for n, t, _ in rows[:-1]) + rows[0][1]
    print(f"--- batch {B}, context {T}, splits {splits}: layer sum {layer * 1e3:.1f} us -> x32 = {layer * 32:.2f} ms "
          f"(+ lm_head {rows[-1][1] * 1e3:.0f} us)")
    for n, t, by in rows:
        print(f"  {n:12s} {t * 1e3:8.1f} us   floor {by / HBM * 1e6:7.1f} us   {by / t / 1e6:7.1f} GB/s "
              f"({by / t * 1e3 / HBM * 100:.0f} % of measured copy)")
