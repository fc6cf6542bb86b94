#!/bin/bash
# A/B sweep of one environment switch on the same box: bash scripts/sweep_env.sh VAR "v1 v2 ..." [bench args]
VAR="$1"; VALUES="$2"; shift 2
mkdir -p gpurun_out
for v in $VALUES; do
  tag="${v//\//_}"
  env "$VAR=$v" timeout 600 python bench.py --steps 64 --warmup 8 --no-baseline "$@" > "gpurun_out/sweep_${VAR}_${tag}.json" 2> "gpurun_out/sweep_${VAR}_${tag}.err"
  python - "$VAR" "$tag" <<'PY'
import json, sys
var, v = sys.argv[1:3]
try:
    d = json.loads(open(f"gpurun_out/sweep_{var}_{v}.json").read().strip().splitlines()[-1])
    print(f"{var}={v}: {d['ms_per_step']:.4f} ms/step  {d['value']:.1f} tok/s  prefill {d['prefill']['tokens_per_sec']:.0f}  e2e {d['e2e']['value']:.1f}")
except Exception as e:
    print(f"{var}={v}: failed ({e})")
PY
done
