#!/usr/bin/env bash
# Fetch Meta-format LLaMA checkpoints with the presigned URL Meta e-mails after the licence form
# (counterpart of the reference's download.sh).  Needs network access and wget; verifies md5 checklists.
#   PRESIGNED_URL='https://…/*?…' MODEL_SIZE=7B,13B TARGET_FOLDER=/data/llama scripts/download.sh
set -euo pipefail
: "${PRESIGNED_URL:?set PRESIGNED_URL to the presigned URL (with the * placeholder)}"
: "${TARGET_FOLDER:?set TARGET_FOLDER}"
MODEL_SIZE="${MODEL_SIZE:-7B}"
declare -A LAST_SHARD=( [7B]=0 [13B]=1 [30B]=3 [33B]=3 [65B]=7 [70B]=7 [8B]=0 )

fetch() { wget --continue "${PRESIGNED_URL/'*'/$1}" -O "${TARGET_FOLDER}/$1"; }

mkdir -p "${TARGET_FOLDER}"
fetch tokenizer.model
fetch tokenizer_checklist.chk
(cd "${TARGET_FOLDER}" && md5sum -c tokenizer_checklist.chk)
for size in ${MODEL_SIZE//,/ }; do
  last="${LAST_SHARD[$size]:?unknown model size $size}"
  mkdir -p "${TARGET_FOLDER}/${size}"
  for s in $(seq 0 "${last}"); do
    fetch "${size}/consolidated.$(printf '%02d' "$s").pth"
  done
  fetch "${size}/params.json"
  fetch "${size}/checklist.chk"
  (cd "${TARGET_FOLDER}/${size}" && md5sum -c checklist.chk)
done
