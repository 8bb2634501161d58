#!/usr/bin/env bash
# Builds libhectr_b200.so in-tree for sm_100a (the GPU box receives the built .so with the snapshot).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${OUT:-$HERE/../libhectr_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --shared \
  -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v -DHB_STREAM \
  "$HERE/hb_api.cu" "$HERE/hb_algo.cpp" "$HERE/hb_wire.cpp" "$HERE/hb_tables.cpp" "$HERE/hb_encode.c" -o "$OUT" "$@"
echo "built $OUT"
