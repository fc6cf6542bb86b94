#!/usr/bin/env python
"""Headline benchmark: LLaMA-3 8B bf16 decode tokens/s (KV cache, 2k-token prompt) at TP = --gpus.

Contract (see the task statement): ``python bench.py --gpus N --steps K --warmup W`` prints ONE JSON line
from rank 0.  A *step* is one decode step (one new token for every row of the batch) of the full model:
embedding → 32 layers → final norm → LM head → greedy selection → sequence bookkeeping, i.e. exactly the
body of the generate loop, replayed from its CUDA graph.  Weights are random-init bf16 of the named
architecture, prompts are synthetic token ids (no network, no checkpoints).

* ``value``           total generated tokens per second over all GPUs (device time, CUDA events, max over ranks)
* ``e2e``             the same metric through the public API (`LLaMA.stream(...).step(feed)`): every step copies
                      that step's input tokens from pinned host memory and reads the new tokens back.
* ``prefill``         prompt tokens per second of the 2k-token prefill (same engine), reported alongside.
* ``gpu_launches``    kernels of libjllama_b200.so launched inside the timed region (graph nodes x replays).

The 15 GB of weights streamed per step are far larger than the 126 MB L2, so no explicit L2 flush is needed
between timed iterations (stated in ``config.l2``).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

BASELINE_PUBLISHED = None      # the reference publishes no performance numbers


def reference_arm(args) -> None:
    """`--impl reference`: the unmodified reference from baseline/_ref through its own API (baseline/reference_runner.py).
    On this image the import fails (no jax / flax; transformers 5.x ships no Flax generation mixin — DESIGN.md,
    "Reference arm"), and the line says so."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "baseline"))
    import reference_runner
    reference_runner.main(args)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index: int, period: float = 0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.error = None

    def run(self) -> None:
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
                getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
            }
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = get_reasons(h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
                time.sleep(self.period)
        except Exception as e:  # noqa: BLE001
            self.error = f"{type(e).__name__}: {e}"

    def stop(self) -> dict:
        self._stop_evt.set()
        self.join(timeout=2)
        out = {"sm_mhz": statistics.median(self.samples) if self.samples else None,
               "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}
        if self.error:
            out["error"] = self.error
        return out


class SyntheticTokenizer:
    """Stands in for a tokenizer file: only the ids the generation front-end needs."""

    def __init__(self, vocab: int, bos: int, eos: int):
        self.n_words, self.bos_id, self.eos_id, self.pad_id = vocab, bos, eos, -1

    def __len__(self) -> int:
        return self.n_words


def tp_check(args, cfg, eng, dev, rank: int, world: int) -> dict:
    """Correctness of the tensor-parallel engine, measured before the timed region and printed in the JSON line
    (the reference trusts its sharded model only after /root/reference/jax_test.py:427-522 compared logits and
    greedy generations over all GPUs).  The single-GPU engine built from the same seed is the same model
    (`B200Engine.random_init` slices full matrices), so on every rank we compare
      * last-token prefill logits, TP vs single GPU (max error relative to the largest logit);
      * 16 greedy tokens, TP vs single GPU, free-running and teacher-forced (each TP token must be the single-GPU
        arg-max or within bf16 noise of it — random-init logits are nearly flat, so exact ties do occur);
      * that all ranks produced identical tokens;
      * the fused sequence-parallel prefill against the all-reduce prefill.
    Models that do not fit twice in one GPU (70B) are checked on a 4-layer model of the same shapes."""
    import copy

    import torch
    import torch.distributed as dist

    from jax_llama_b200.models.engine import B200Engine

    ccfg, eng_tp, own = cfg, eng, False
    if cfg.num_hidden_layers * cfg.hidden_size * cfg.intermediate_size * 6 > 40e9:      # > ~40 GB of bf16 weights
        ccfg = copy.deepcopy(cfg)
        ccfg.num_hidden_layers = 4
        eng_tp = B200Engine.random_init(ccfg, dev, seed=0, tp=world, rank=rank, comm=eng.comm)
        own = True
    full = B200Engine.random_init(ccfg, dev, seed=0)
    P, n_new, B = 1024, 16, 1
    g = torch.Generator().manual_seed(4321)
    ids = torch.randint(0, ccfg.vocab_size, (B, P), generator=g).to(dev)
    want = full.forward(ids, last_only=True).logits[:, 0].float()
    got = eng_tp.forward(ids, last_only=True).logits[:, 0].float()
    scale = float(want.abs().max())
    err = float((got - want).abs().max()) / (scale + 1e-9)
    seq_tp = eng_tp.generate(ids, None, max_length=P + n_new, eos_token_id=-1)
    seq_1 = full.generate(ids, None, max_length=P + n_new, eos_token_id=-1)
    equal = bool(torch.equal(seq_tp, seq_1))
    # teacher-forced: the single-GPU engine scores the TP engine's tokens
    forced = full.forward(seq_tp[:, :-1], last_only=False).logits.float()
    # A token the TP engine picked may trail the single-GPU arg-max by what two bf16 engines with different reduction
    # orders differ by on these (nearly flat, random-init) logits: twice the measured logits error, at least 2 %.
    tol = max(2e-2, 2.0 * err)
    ok_forced = True
    for t in range(P, P + n_new):
        row = forced[:, t - 1]
        chosen = row.gather(-1, seq_tp[:, t:t + 1])[:, 0]
        ok_forced &= bool(((row.max(-1).values - chosen) <= tol * row.abs().max()).all())
    gathered = [torch.empty_like(seq_tp) for _ in range(world)]
    dist.all_gather(gathered, seq_tp)
    agree = all(bool(torch.equal(gathered[0], t)) for t in gathered)
    out = {"layers": ccfg.num_hidden_layers, "prompt_len": P, "new_tokens": n_new,
           "logits_max_rel_err": round(err, 5), "greedy_tokens_equal": equal,
           "greedy_tokens_teacher_forced_ok": ok_forced, "teacher_forced_tolerance_rel": round(tol, 4),
           "ranks_agree": agree}
    # fused sequence-parallel prefill vs the all-reduce prefill (same engine)
    if P % (128 * world) == 0 and hasattr(eng_tp, "prefill_paths"):
        outs = {}
        for path in eng_tp.prefill_paths():
            eng_tp.set_prefill_path(path)
            outs[path] = eng_tp.forward(ids, last_only=True).logits[:, 0].float()
        names = sorted(outs)
        if len(names) == 2:
            out["sp_vs_allreduce_prefill_max_rel_err"] = round(
                float((outs[names[0]] - outs[names[1]]).abs().max()) / (scale + 1e-9), 5)
        eng_tp.set_prefill_path(None)
    flags = torch.tensor([int(equal), int(ok_forced), int(agree)], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)                 # the line reports what held on EVERY rank
    out["greedy_tokens_equal"], out["greedy_tokens_teacher_forced_ok"], out["ranks_agree"] = (bool(int(x)) for x in flags)
    e = torch.tensor([err], device=dev)
    dist.all_reduce(e, op=dist.ReduceOp.MAX)
    out["logits_max_rel_err"] = round(float(e.item()), 5)
    del full
    if own:
        del eng_tp
    torch.cuda.empty_cache()
    return out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama3-8b")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--prompt-len", type=int, default=2048)
    ap.add_argument("--prefill-reps", type=int, default=3)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-baseline", action="store_true", help="skip the same-box PyTorch (cuBLAS + SDPA + NCCL) arm")
    ap.add_argument("--no-tp-check", action="store_true", help="skip the pre-run TP-vs-single-GPU correctness check")
    ap.add_argument("--sweep-batches", default="", help="comma-separated extra batch sizes measured with the same engine "
                    "after the headline run (device-timed decode + prefill); reported under \"batch_sweep\"")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
        return
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist

    from jax_llama_b200 import LLaMA, LLaMAForCausalLM, preset
    from jax_llama_b200.models.engine import B200Engine
    from jax_llama_b200.ops import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    cfg = preset(args.model)
    B, P = args.batch, args.prompt_len
    max_len = P + 1 + args.warmup + args.steps + 8
    cfg.max_sequence_length = max(cfg.max_sequence_length, (max_len + 1) // 2)
    kw = {}
    if world > 1:
        kw = dict(tp=world, rank=rank)
    eng = B200Engine.random_init(cfg, dev, seed=0, use_cuda_graph=not args.no_graph, **kw)
    check = None
    if world > 1 and not args.no_tp_check:
        check = tp_check(args, cfg, eng, dev, rank, world)
    model = LLaMAForCausalLM(cfg, _do_init=False)
    params = {"engine": eng}                              # handle understood by LLaMAForCausalLM._engine_for
    tok = SyntheticTokenizer(cfg.vocab_size, cfg.bos_token_id, -1)      # eos never fires: fixed-length runs
    llama = LLaMA(params, model, tok)

    g = torch.Generator().manual_seed(1234)
    prompt = torch.randint(0, cfg.vocab_size, (B, P), generator=g).pin_memory()
    mask = torch.ones(B, P, dtype=torch.int32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------------------------------------------------------- prefill (reported alongside)
    ids_dev = prompt.to(dev, non_blocking=True)
    st = eng._decode_state(B, max_len)
    prefill_ms = []
    for i in range(1 + args.prefill_reps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.prepare(st, ids_dev, None, 0)
        e1.record()
        torch.cuda.synchronize()
        if i > 0:
            prefill_ms.append(max_over_ranks(e0.elapsed_time(e1)))
    prefill_best = min(prefill_ms)
    # the other prefill paths of the same engine, timed for the record: GEMM + in-switch all-reduce kernel, and the
    # same with the all-reduce done by NCCL (the baseline arm; nothing in the engine calls NCCL by itself)
    PATH_NAMES = {"single": "single GPU", "sequence_parallel": "fused sequence-parallel (NVLS reduce-scatter / all-gather)",
                  "allreduce": "GEMM + in-switch all-reduce kernel (multimem.ld_reduce)",
                  "nccl": "GEMM + NCCL all-reduce (baseline arm)"}
    default_path = ("single" if world == 1 else
                    "sequence_parallel" if (eng.sequence_parallel_prefill and (B * P) % (128 * world) == 0) else "allreduce")
    prefill_alt = None
    if world > 1:
        prefill_alt = []
        for path in ("sequence_parallel", "allreduce", "nccl"):
            if path == default_path or (path == "sequence_parallel" and (B * P) % (128 * world)):
                continue
            eng.set_prefill_path("allreduce" if path == "nccl" else path)
            nv = eng.comm.nvls_state(B * P, cfg.hidden_size)
            nv.use_nccl_baseline = path == "nccl"
            alt = []
            for i in range(1 + args.prefill_reps):
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                eng.prepare(st, ids_dev, None, 0)
                e1.record()
                torch.cuda.synchronize()
                if i > 0:
                    alt.append(max_over_ranks(e0.elapsed_time(e1)))
            nv.use_nccl_baseline = False
            prefill_alt.append({"path": PATH_NAMES[path], "ms": round(min(alt), 3),
                                "tokens_per_sec": round(B * P / (min(alt) / 1e3), 1)})
        eng.set_prefill_path(None)
        eng.prepare(st, ids_dev, None, 0)

    # ---------------------------------------------------------------- device-timed decode steps
    stream = llama.stream(prompt, mask, max_gen_len=max_len - P, temperature=0.0, top_p=1.0)
    for _ in range(args.warmup):
        stream.graph.replay() if stream.graph is not None else eng._decode_step(
            stream.st, stream.cache, stream.sampling, stream.eos, stream.pad)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        if stream.graph is not None:
            stream.graph.replay()
        else:
            eng._decode_step(stream.st, stream.cache, stream.sampling, stream.eos, stream.pad)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_per_step = ms_total / args.steps
    launches_eager = _lib.launch_count() - launches0
    per_step_launches = eng.launches_per_step(stream.st) if hasattr(eng, "launches_per_step") else None
    gpu_launches = launches_eager if stream.graph is None else (per_step_launches or 0) * args.steps
    used_graph = stream.graph is not None

    # ---------------------------------------------------------------- end to end through the public API
    stream2 = llama.stream(prompt, mask, max_gen_len=max_len - P, temperature=0.0, top_p=1.0)
    feed = torch.empty(B, dtype=torch.int32).pin_memory()
    feed.copy_(stream2.last_tokens())
    for _ in range(args.warmup):
        feed.copy_(stream2.step(feed))
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = stream2.step(feed)          # H2D of this step's tokens (pinned) + D2H of the new tokens + sync
        feed.copy_(out)
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    h2d_bytes, d2h_bytes = stream2.h2d_bytes, stream2.d2h_bytes

    # ---------------------------------------------------------------- other batch sizes, same engine (optional)
    sweep = None
    if args.sweep_batches:
        sweep = []
        del stream, stream2, st
        for b in [int(v) for v in args.sweep_batches.split(",") if v]:
            eng._decode_states.clear()                     # drop the previous batch's KV cache and graphs
            torch.cuda.empty_cache()
            pr = torch.randint(0, cfg.vocab_size, (b, P), generator=g).pin_memory()
            mk = torch.ones(b, P, dtype=torch.int32).pin_memory()
            barrier()
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record()
            sb = llama.stream(pr, mk, max_gen_len=max_len - P, temperature=0.0, top_p=1.0)       # prefill + first token
            p1.record()
            torch.cuda.synchronize()
            pre_ms = max_over_ranks(p0.elapsed_time(p1))
            step = (sb.graph.replay if sb.graph is not None else
                    (lambda sb=sb: eng._decode_step(sb.st, sb.cache, sb.sampling, sb.eos, sb.pad)))
            for _ in range(args.warmup):
                step()
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            for _ in range(args.steps):
                step()
            s1.record()
            barrier()
            ms = max_over_ranks(s0.elapsed_time(s1)) / args.steps
            kv_bytes = 2 * len(eng.layers) * b * eng.nkv * eng.hd * 2 * (P + args.warmup + args.steps // 2)
            sweep.append({"global_batch": b, "ms_per_step": round(ms, 4), "tokens_per_sec": round(b / (ms / 1e3), 1),
                          "prefill_ms_incl_first_token_and_graph_capture": round(pre_ms, 2),
                          "kv_bytes_per_step_per_rank": kv_bytes,
                          "launches_per_step": eng.launches_per_step(sb.st)})
            del sb

    # ---------------------------------------------------------------- same-box library yardstick
    baseline = None
    if not args.no_baseline:
        try:
            from jax_llama_b200.models.baseline import TorchBaseline
            tb = TorchBaseline(eng, B, max_len)
            barrier()
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b0.record()
            tb.prefill(ids_dev)
            b1.record()
            torch.cuda.synchronize()
            base_prefill_ms = max_over_ranks(b0.elapsed_time(b1))
            for _ in range(args.warmup):
                tb.decode_step()
            barrier()
            b0.record()
            for _ in range(args.steps):
                tb.decode_step()
            b1.record()
            barrier()
            base_ms = max_over_ranks(b0.elapsed_time(b1)) / args.steps
            baseline = {"impl": "torch F.linear (cuBLAS) + SDPA + NCCL, CUDA-graphed, same weights",
                        "decode_tokens_per_sec": round(B / (base_ms / 1e3), 2), "ms_per_step": round(base_ms, 4),
                        "prefill_tokens_per_sec_first_call": round(B * P / (base_prefill_ms / 1e3), 1)}
            del tb
        except Exception as e:  # noqa: BLE001  (the yardstick must never take the product number down with it)
            baseline = {"error": f"{type(e).__name__}: {str(e)[:160]}"}

    if rank == 0:
        value = B * args.steps / (ms_total / 1e3)
        weight_bytes = sum(t.numel() * 2 for l in eng.layers for t in (l.wqkv, l.wo, l.w13, l.w2)) + eng.lm_head.numel() * 2
        peaks = {}
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        kv_main = 2 * len(eng.layers) * B * eng.nkv * eng.hd * 2 * (P + args.warmup + args.steps // 2)
        for sw in sweep or []:
            sw["hbm_roofline_incl_kv_frac"] = round(
                ((weight_bytes + sw["kv_bytes_per_step_per_rank"]) / (sw["ms_per_step"] / 1e3)) / (hbm * 1e9), 4)
        # prefill roofline: the slower of (FLOPs of this rank at the measured sustained cuBLAS bf16 rate) and (bytes
        # this rank must move over NVLink at the measured 770 GB/s per direction), assuming perfect overlap
        M_tok = B * P
        layer_w = sum(t.numel() for l in eng.layers[:1] for t in (l.wqkv, l.wo, l.w13, l.w2)) * len(eng.layers)
        flops = 2.0 * M_tok * layer_w + 2.0 * 2.0 * B * (P * (P + 1) / 2) * eng.nh * eng.hd * len(eng.layers)
        t_compute = flops / (peaks.get("bf16_tflops_sustained", 1400.0) * 1e12) * 1e3
        nvl_bytes = 0.0 if world == 1 else 4.0 * len(eng.layers) * (world - 1) / world * M_tok * cfg.hidden_size * 2
        t_nvlink = nvl_bytes / 770e9 * 1e3
        prefill_roofline = {"compute_ms_at_sustained_peak": round(t_compute, 3), "nvlink_ms_at_770GBs": round(t_nvlink, 3),
                            "frac_of_measured": round(max(t_compute, t_nvlink) / prefill_best, 4)}
        line = {
            "metric": "decode_tokens_per_sec", "value": round(value, 2), "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": BASELINE_PUBLISHED, "dtype": "bf16",
            "data": "synthetic token ids, random-init weights",
            "config": {"model": args.model, "global_batch": B, "seq_len": P, "parallelism": f"tp{world}",
                       "decode_context": f"{P}+{args.warmup}..{P + args.warmup + args.steps}",
                       "l2": "no flush needed: per-step weight stream (%.1f GB/rank) >> 126 MB L2" % (weight_bytes / 1e9),
                       "cuda_graph": used_graph},
            "p50_ms_per_token": round(ms_per_step, 4),
            "hbm_roofline_frac_of_measured": round((weight_bytes / (ms_per_step / 1e3)) / (hbm * 1e9), 4),
            # the same with the K/V cache the step must read (what bounds batched decode; negligible at batch 1 x 2k)
            "hbm_roofline_incl_kv_frac": round(((weight_bytes + kv_main) / (ms_per_step / 1e3)) / (hbm * 1e9), 4),
            "prefill": {"tokens_per_sec": round(B * P / (prefill_best / 1e3), 1), "ms": round(prefill_best, 3),
                        "prompt_len": P, "includes": "embedding..lm_head(last token) of the prompt",
                        "roofline": prefill_roofline,
                        "path": PATH_NAMES[default_path],
                        "other_path": prefill_alt},
            "e2e": {"value": round(B * args.steps / (e2e_ms / 1e3), 2), "unit": "tokens/s",
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "api": "LLaMA.stream(...).step(feed)"},
            "gpu_launches": int(gpu_launches), "clocks": clocks, "tp_check": check,
            "batch_sweep": sweep,
            "same_box_library_baseline": baseline,
            "speedup_vs_library_baseline": (round(baseline["ms_per_step"] / ms_per_step, 3)
                                            if baseline and "ms_per_step" in baseline else None),
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
