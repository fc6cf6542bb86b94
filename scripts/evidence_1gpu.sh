#!/bin/bash
# One-GPU evidence run (through gpurun): full GPU test suite, the single-GPU benchmark configs, an L2-prefetch sweep of
# the decode kernel and ncu --set full captures of the hot kernels (each capture only after the same command exited 0
# without ncu).  Outputs under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for spec in "llama3-8b 1" "llama2-7b 1" "llama2-7b 8" "llama2-7b 32" "llama3-8b 8" "llama3-8b 32"; do
  set -- $spec
  timeout 600 python bench.py --model $1 --batch $2 --steps 64 --warmup 8 --prompt-len 2048 > gpurun_out/bench_$1_b$2.json 2> gpurun_out/bench_$1_b$2.err
  python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads(open(f"gpurun_out/bench_{sys.argv[1]}_b{sys.argv[2]}.json").read().strip().splitlines()[-1])
    print(sys.argv[1], "batch", sys.argv[2], d["ms_per_step"], "ms/step", d["value"], "tok/s", "roofline", d["hbm_roofline_frac_of_measured"],
          "prefill", d["prefill"]["tokens_per_sec"], "lib", (d.get("same_box_library_baseline") or {}).get("ms_per_step"))
except Exception as e:
    print(sys.argv[1], sys.argv[2], "failed", e)
PY
done
bash scripts/sweep_env.sh JLLAMA_MEGA_L2_AHEAD "0 8 16 32 48"
python scripts/profile_run.py --mode prefill --layers 2 --steps 1 > gpurun_out/plain_prefill.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemm_kernel|flash_prefill|rmsnorm|rope_kv|embed' -c 14 \
    -o gpurun_out/prof_prefill python scripts/profile_run.py --mode prefill --layers 2 --steps 1 > gpurun_out/ncu_prefill.log 2>&1
python scripts/profile_run.py --mode decode --layers 4 --steps 2 > gpurun_out/plain_decode.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'decode_mega|mega_select|sample_full|argmax|advance' -c 6 \
    -o gpurun_out/prof_decode python scripts/profile_run.py --mode decode --layers 4 --steps 2 > gpurun_out/ncu_decode.log 2>&1
python scripts/profile_run.py --mode decode --layers 2 --steps 1 --batch 16 --prompt 512 > gpurun_out/plain_decode16.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gemv|decode_attn|swiglu|sample_full' -c 12 \
    -o gpurun_out/prof_decode16 python scripts/profile_run.py --mode decode --layers 2 --steps 1 --batch 16 --prompt 512 > gpurun_out/ncu_decode16.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_prefill.log gpurun_out/ncu_decode.log gpurun_out/ncu_decode16.log
