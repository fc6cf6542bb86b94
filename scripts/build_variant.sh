#!/bin/bash
# Builds build/ab/NAME.so = the current objects with ONE source recompiled with extra -D flags (kernel A/B runs:
# JLLAMA_LIB=build/ab/NAME.so).  Usage: scripts/build_variant.sh NAME csrc/file.cu -DFLAG [-DFLAG2 ...]
set -e
NAME="$1"; SRC="$2"; shift 2
cd "$(dirname "$0")/.."
mkdir -p build/ab build/obj_ab
python -m jax_llama_b200.ops.build > /dev/null
OBJ="build/obj_ab/${NAME}_$(basename "${SRC%.cu}").o"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --use_fast_math -Xcompiler -fPIC \
  -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr -Iinclude -Icsrc "$@" -c "$SRC" -o "$OBJ"
OTHERS=$(ls build/obj/*.o | grep -v "/$(basename "${SRC%.cu}").o")
nvcc -shared -o "build/ab/${NAME}.so" $OTHERS "$OBJ" -gencode arch=compute_100a,code=sm_100a
echo "build/ab/${NAME}.so"
