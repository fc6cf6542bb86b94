#!/bin/bash
# compute-sanitizer pass over the single-GPU kernel tests (run on the GPU box AFTER the same tests passed
# plainly).  memcheck by default; TOOL=racecheck|synccheck|initcheck selects another tool.
# Usage: bash scripts/sanitize.sh [pytest -k expression]
TOOL="${TOOL:-memcheck}"
EXPR="${1:-rmsnorm or rope or swiglu or embed or gemv_modes or decode_attention or argmax or sampling or flash or gemm}"
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool "$TOOL" --error-exitcode 9 --print-limit 20 \
  --log-file "gpurun_out/sanitizer_${TOOL}.log" \
  python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 1400 -k "$EXPR" \
  > "gpurun_out/sanitizer_${TOOL}_pytest.log" 2>&1
rc=$?
echo "sanitizer($TOOL): exit $rc :: $(tail -1 gpurun_out/sanitizer_${TOOL}_pytest.log)"
grep -c "========= Invalid\|========= Race\|========= Error" "gpurun_out/sanitizer_${TOOL}.log" 2>/dev/null
tail -3 "gpurun_out/sanitizer_${TOOL}.log"
exit $rc
