"""profiles/r2/*.json (one bench.py line each) → profiles/summary.json (headline fields only)."""
import glob
import json
import os

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
runs = {}
for f in sorted(glob.glob(os.path.join(root, "profiles", "r2", "bench_*.json"))):
    line = open(f).read().strip().splitlines()[-1]
    try:
        d = json.loads(line)
    except json.JSONDecodeError:
        continue
    cfg = d.get("config", {})
    runs[os.path.basename(f)[:-5]] = {
        "model": cfg.get("model"), "parallelism": cfg.get("parallelism"), "batch": cfg.get("global_batch"),
        "prompt_len": cfg.get("seq_len"), "decode_tok_s": d.get("value"), "ms_per_step": d.get("ms_per_step"),
        "hbm_roofline_frac_of_measured": d.get("hbm_roofline_frac_of_measured"),
        "prefill_ms": d.get("prefill", {}).get("ms"), "prefill_tok_s": d.get("prefill", {}).get("tokens_per_sec"),
        "prefill_path": d.get("prefill", {}).get("path"),
        "prefill_other_paths": d.get("prefill", {}).get("other_path"),
        "e2e_tok_s": d.get("e2e", {}).get("value"),
        "gpu_launches_per_step": (d.get("gpu_launches") or 0) / max(d.get("steps", 1), 1),
        "tp_check": d.get("tp_check"), "clocks": d.get("clocks"),
        "batch_sweep": [{k: s[k] for k in ("global_batch", "ms_per_step", "tokens_per_sec", "launches_per_step")}
                        for s in (d.get("batch_sweep") or [])] or None,
        "eager_arm_ms_per_step": (d.get("same_box_library_baseline") or {}).get("ms_per_step"),
    }
out = {"note": "Round-2 headline numbers (one JSON line per bench run is next to this file under profiles/r2/); "
               "device-timed, max over ranks; peaks: MEASURED_PEAKS.json", "runs": runs}
json.dump(out, open(os.path.join(root, "profiles", "summary.json"), "w"), indent=1)
print(len(runs), "runs")
