#!/usr/bin/env bash
# Builds libquest_b200.so for sm_100a in-tree (the .so travels to the GPU box with the snapshot).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="${QUEST_EXTRA_FLAGS:-} -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -Xcompiler -fPIC -Xptxas -v"
SRCS="capi.cu estimate.cu page.cu topk.cu decode.cu rope_norm.cu prefill.cu"
mkdir -p build
pids=()
for s in $SRCS; do
  $NVCC $FLAGS -c "$s" -o "build/${s%.cu}.o" 2> "build/${s%.cu}.ptxas.log" &
  pids+=($!)
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=1; done
if [ $rc -ne 0 ]; then cat build/*.ptxas.log >&2; exit 1; fi
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libquest_b200.so build/*.o -lcudart
echo "built $(cd .. && pwd)/libquest_b200.so"
