#!/bin/bash
# Builds hmm_layer_b200/libhmm_b200.so for sm_100a (B200).  nvcc cross-compiles without a GPU.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libhmm_b200.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
SRCS=("$HERE"/*.cu)
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC -Xcompiler -O3 -shared -o "$OUT" "${SRCS[@]}" ${HMM_NVCC_EXTRA:-}
echo "built $OUT"
