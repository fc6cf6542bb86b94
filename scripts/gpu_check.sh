#!/bin/bash
# Runs the GPU test groups in separate processes (a faulting kernel poisons its CUDA context, not the rest).
# Usage (on the GPU box, from the repo root):  bash scripts/gpu_check.sh [quick]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
run() {  # name, pytest -k expression, file
  timeout 600 python -m pytest "$3" -m gpu -q --timeout 300 -k "$2" > "gpurun_out/test_$1.log" 2>&1
  echo "$1: exit $? :: $(tail -1 gpurun_out/test_$1.log)"
}
run basic "library or embed or rmsnorm or rope or swiglu" tests/test_gpu_kernels.py
run gemv "gemv_modes" tests/test_gpu_kernels.py
run gemvtma "gemv_tma" tests/test_gpu_kernels.py
run gemm "gemm" tests/test_gpu_kernels.py
run attn "decode_attention and not fused" tests/test_gpu_kernels.py
run attnfused "decode_attention_fused" tests/test_gpu_kernels.py
run flash "flash" tests/test_gpu_kernels.py
run sampling "argmax or sampling" tests/test_gpu_kernels.py
run engine "" tests/test_gpu_engine.py
for f in tests/test_gpu_*.py; do
  case "$f" in tests/test_gpu_kernels.py|tests/test_gpu_engine.py) ;; *)
    n=$(basename "$f" .py); run "$n" "" "$f";; esac
done
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke: exit $? :: $(tail -1 gpurun_out/smoke.log)"
