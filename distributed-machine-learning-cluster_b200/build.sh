#!/bin/bash
# Builds libb200infer.so (sm_100a only) in-tree. nvcc cross-compiles without a GPU.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="$HERE/csrc"
OUT="$HERE/libb200infer.so"
OBJ="$HERE/build"
mkdir -p "$OBJ"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(${B2I_EXTRA_NVCC_FLAGS:-} -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off,-Wall,-Wno-unused-function)
pids=()
for f in conv_umma.cu stem.cu stem2.cu stem_alex.cu conv_fp32.cu misc_kernels.cu resize.cu ot_reader.cpp engine.cpp predict.cpp jpeg_decode.cpp capi.cpp; do
  o="$OBJ/${f%.*}.o"
  if [ ! -f "$o" ] || [ "$SRC/$f" -nt "$o" ] || [ -n "$(find "$SRC" "$HERE/../include" -name '*.h' -newer "$o" -o -name '*.cuh' -newer "$o")" ]; then
    "$NVCC" "${FLAGS[@]}" -c "$SRC/$f" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
"$NVCC" -shared -o "$OUT" "$OBJ"/*.o -lcudart_static -ldl -lpthread -lrt
echo "built $OUT"
