"""Eager (no CUDA graph) batched decode of a random-init model, meant to run under
`ncu --metrics gpu__time_duration.sum --csv` so that every kernel of a decode step gets its own duration:

    python scripts/decode_kernel_times.py                       # plain run (must exit 0 first)
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/kt.csv \
        python scripts/decode_kernel_times.py
    python scripts/decode_kernel_times.py --summarise gpurun_out/kt.csv

The prompt is short (the prefill is not the subject); the cache length is set with --context so that the attention
kernel reads as much K/V as in the benchmark.  ncu serialises the kernels: the sum is an upper bound of the step."""
import argparse
import collections
import csv
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def summarise(path: str, steps: int) -> None:
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    ui = hdr.index("Metric Unit")
    seq = []
    for r in rows[rows.index(hdr) + 1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        v = v / 1000.0 if r[ui] in ("ns", "nsecond") else v
        seq.append((r[ki].split("(")[0][-60:], v))
    # the last `steps` decode steps: find the period from the position of the sampler kernel
    idx = [i for i, (k, _) in enumerate(seq) if "sample_full" in k]
    if len(idx) < 2:
        print("no decode steps found"); return
    lo, hi = idx[-2] + 1, idx[-1] + 1
    step = seq[lo:hi]
    tot = collections.OrderedDict()
    for k, v in step:
        c, s = tot.get(k, (0, 0.0))
        tot[k] = (c + 1, s + v)
    total = sum(v for _, v in step)
    print(f"one decode step: {len(step)} kernels, {total / 1000:.3f} ms summed (serialised by ncu)")
    for k, (c, s) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"  {s:9.1f} us  {c:4d} x {s / c:7.2f} us  {k}")


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3-8b")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--context", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--summarise", default=None)
    a = ap.parse_args()
    if a.summarise:
        summarise(a.summarise, a.steps)
        return
    import torch
    from jax_llama_b200 import preset
    from jax_llama_b200.models.engine import B200Engine
    cfg = preset(a.model)
    dev = torch.device("cuda", 0)
    eng = B200Engine.random_init(cfg, dev, seed=0, use_cuda_graph=False)
    max_len = a.context + a.steps + 8
    cfg.max_sequence_length = max(cfg.max_sequence_length, max_len)
    ids = torch.randint(0, cfg.vocab_size, (a.batch, 16), device=dev)
    st = eng._decode_state(a.batch, max_len)
    cache = eng.prepare(st, ids, None, 0)
    sampling = (False, 1.0, 0, 1.0, 0)
    eng._select_and_advance_after_prefill(st, cache, sampling, -1, 0)
    # pretend the cache is `context` tokens long (its contents are whatever the allocation held: timing only)
    cache.k.normal_(); cache.v.normal_()
    cache.len_dev.fill_(a.context)
    st.cur_len.fill_(a.context)
    for _ in range(a.steps):
        eng._decode_step(st, cache, sampling, -1, 0)
    torch.cuda.synchronize()
    print("ok", int(cache.len_dev.item()))


if __name__ == "__main__":
    main()
