#!/usr/bin/env bash
# Compiles the reference's own decode-path kernels (append, estimate, sparse decode + merge) for sm_100a from
# /root/reference into oracle/_ref/libquest_ref.so.  Authoring container only (the GPU box uses the prebuilt file).
set -euo pipefail
cd "$(dirname "$0")"
REF=${QUEST_REFERENCE:-/root/reference}
OUT=../_ref
mkdir -p "$OUT"
if [ -f "$OUT/libquest_ref.so" ] && [ "$OUT/libquest_ref.so" -nt ref_harness.cu ]; then echo "up to date: $OUT/libquest_ref.so"; exit 0; fi
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -w \
  -I "$REF/kernels/include" -I "$REF/kernels/3rdparty/flashinfer/include" \
  -Xcompiler -fPIC -shared ref_harness.cu -o "$OUT/libquest_ref.so" -lcudart
echo "built $(cd "$OUT" && pwd)/libquest_ref.so"
