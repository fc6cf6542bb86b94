#!/usr/bin/env python
"""Per-op timeline of the persistent decode kernel (globaltimer stamps from every block)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from jax_llama_b200 import preset  # noqa: E402
from jax_llama_b200.models.engine import B200Engine  # noqa: E402
from jax_llama_b200.ops import _lib  # noqa: E402

from jax_llama_b200.parallel.comm import init_distributed  # noqa: E402

EV = 32
rank, world, local = init_distributed()
torch.cuda.set_device(local)
cfg = preset(os.environ.get("MODEL", "llama3-8b"))
kw = dict(tp=world, rank=rank) if world > 1 else {}
eng = B200Engine.random_init(cfg, f"cuda:{local}", use_cuda_graph=False, **kw)
L = cfg.num_hidden_layers
B = int(os.environ.get("BATCH", "1"))
ids = torch.randint(0, cfg.vocab_size, (B, 2048), device="cuda")
st = eng._decode_state(B, 2048 + 32)
cache = eng.prepare(st, ids, None, 0)
samp = (False, 1.0, 0, 1.0, 0)
eng._select_and_advance_after_prefill(st, cache, samp, -1, 0)
for _ in range(3):
    eng._decode_step(st, cache, samp, -1, 0)
NB = torch.cuda.get_device_properties(local).multi_processor_count
trace = torch.zeros(NB * L * EV, dtype=torch.int64, device="cuda")
lib = _lib.lib()
lib.jl_decode_mega_set_trace.argtypes = [C.c_void_p]
lib.jl_decode_mega_set_trace(trace.data_ptr())
eng._decode_step(st, cache, samp, -1, 0)
torch.cuda.synchronize()
lib.jl_decode_mega_set_trace(None)
tall = trace.view(NB, L, EV).cpu().double()
t = tall[[0, NB - 1]]
if rank != 0:
    sys.exit(0)
names = ["stage_qkv(poll x)", "gemv_qkv", "attn(poll qkv)", "merge(poll parts)", "stage_wo(poll attn)", "gemv_wo(+AR)",
         "stage_w13(poll x)", "gemv_w13", "stage_w2(poll h)", "gemv_w2(+AR)"]
lo, hi = 4, L - 4
for blk in range(2):
    d = (t[blk, lo:hi, 1:11] - t[blk, lo:hi, 0:10]) / 1e3          # µs
    print(f"block {'0' if blk == 0 else 'last'}: " + "  ".join(f"{n}={v:.1f}" for n, v in zip(names, d.mean(0).tolist())))
    print(f"   layer total {((t[blk, lo:hi, 10] - t[blk, lo:hi, 0]) / 1e3).mean():.1f} us; layer-to-layer "
          f"{((t[blk, lo + 1:hi, 0] - t[blk, lo:hi - 1, 0]) / 1e3).mean():.1f} us")
dall = (tall[:, lo:hi, 1:11] - tall[:, lo:hi, 0:10]) / 1e3         # [blocks, layers, phases]
print("all blocks, mean over layers, then over blocks (min / p50 / max):")
for i, n in enumerate(names):
    m = dall[:, :, i].mean(1)
    print(f"   {n:22s} {m.min():6.2f} {m.median():6.2f} {m.max():6.2f}")
print(f"   layer total p50 {((tall[:, lo:hi, 10] - tall[:, lo:hi, 0]) / 1e3).mean(1).median():.1f} us")
a = (t[0, lo:hi, 16:21] - t[0, lo:hi, 2:3]) / 1e3
print("block 0 attention internals (µs after the phase starts): " + "  ".join(
    f"{n}={v:.1f}" for n, v in zip(["setup", "q+rope", "kv_loop", "new_tok", "fold+write"], a.mean(0).tolist())))

# data-flow waits: per layer, how long after the LAST block finished producing does the FIRST / median block finish
# consuming (stamps: 6 = wo GEMV + all-reduce done, 7 = x staged for w13; 10 = w2 + all-reduce done, next layer's 1)
def tail(prod, cons, shift=0):
    tp_ = tall[:, lo:hi, prod]                                  # [blocks, layers]
    tc_ = tall[:, lo + shift:hi + shift, cons]
    last = tp_.max(0).values
    return (((last - tp_.median(0).values) / 1e3).mean().item(), ((tc_.min(0).values - last) / 1e3).mean().item(),
            ((tc_.median(0).values - last) / 1e3).mean().item())
for name, pr, co, sh in (("qkv→attn(q+rope)", 2, 17, 0), ("attn→merge", 3, 4, 0), ("merge→stage_wo", 4, 5, 0),
                         ("wo(+AR)→stage_w13", 6, 7, 0), ("w13→stage_w2", 8, 9, 0), ("w2(+AR)→stage_qkv", 10, 1, 1)):
    a_, b_, c_ = tail(pr, co, sh)
    print(f"{name:22s} producer tail (last - median) {a_:5.2f} us | first consumer done {b_:5.2f} us after the last producer, median {c_:5.2f}")
for name, ev, start in (("poll x (qkv)", 21, 0), ("poll x (w13)", 27, 6), ("poll attn", 24, 4), ("poll h", 25, 8)):
    d_ = (tall[:, lo:hi, ev] - tall[:, lo:hi, start]) / 1e3
    print(f"{name:14s} poll finished after phase start: min {d_.mean(1).min():5.2f} p50 {d_.mean(1).median():5.2f} max {d_.mean(1).max():5.2f}")

# spread across blocks: when does each block finish each phase, relative to the earliest block entering it
ends = {"gemv_qkv": 2, "attn": 3, "gemv_wo": 6, "gemv_w13": 8, "gemv_w2": 10}
starts = {"gemv_qkv": 1, "attn": 2, "gemv_wo": 5, "gemv_w13": 7, "gemv_w2": 9}
smid = tall[:, lo, 31].long()
for n in ends:
    begin = tall[:, lo:hi, starts[n]].min(0).values                      # [layers]
    fin = (tall[:, lo:hi, ends[n]] - begin[None]) / 1e3                  # [blocks, layers]
    own = (tall[:, lo:hi, ends[n]] - tall[:, lo:hi, starts[n]]) / 1e3
    m = fin.mean(1)
    order = m.argsort()
    print(f"{n:9s} finish after phase start: min {m.min():5.1f}  p50 {m.median():5.1f}  p90 {m.quantile(0.9):5.1f}  max {m.max():5.1f} us"
          f" | own duration p50 {own.mean(1).median():5.1f} max {own.mean(1).max():5.1f}"
          f" | slowest blocks {order[-6:].tolist()} (sm {smid[order[-6:]].tolist()}) fastest {order[:4].tolist()}")
