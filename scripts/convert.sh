#!/usr/bin/env bash
# Meta checkpoints → per-rank safetensors cache (counterpart of the reference's convert.sh, which calls a
# converter CLI the reference never shipped; here it exists: python -m jax_llama_b200.convert_weights).
#   SOURCE_FOLDER=/data/llama TARGET_FOLDER=/data/llama_b200 MODEL_SIZE=8B,70B TP=8 IS_LLAMA3=true scripts/convert.sh
set -euo pipefail
: "${SOURCE_FOLDER:?set SOURCE_FOLDER (directory holding tokenizer.model and one folder per model size)}"
: "${TARGET_FOLDER:?set TARGET_FOLDER}"
MODEL_SIZE="${MODEL_SIZE:-7B}"
TP="${TP:-1}"
IS_LLAMA3="${IS_LLAMA3:-false}"
cd "$(dirname "$0")/.."
mkdir -p "${TARGET_FOLDER}"
cp "${SOURCE_FOLDER}/tokenizer.model" "${TARGET_FOLDER}/"
[ -f "${SOURCE_FOLDER}/tokenizer_checklist.chk" ] && cp "${SOURCE_FOLDER}/tokenizer_checklist.chk" "${TARGET_FOLDER}/"
for size in ${MODEL_SIZE//,/ }; do
  python -m jax_llama_b200.convert_weights --ckpt_dir "${SOURCE_FOLDER}/${size}/" \
    --out_dir "${TARGET_FOLDER}/${size}-tp${TP}/" --tokenizer_path "${SOURCE_FOLDER}/tokenizer.model" \
    --is_llama3 "${IS_LLAMA3}" --tp "${TP}" --verbose
done
