#!/usr/bin/env bash
# Builds libquest_b200.so for sm_100a in-tree (the .so travels to the GPU box with the snapshot).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="${QUEST_EXTRA_FLAGS:-} -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -Xcompiler -fPIC -Xptxas -v"
SRCS="capi.cu estimate.cu page.cu topk.cu decode.cu rope_norm.cu prefill.cu prefill_tma.cu peer.cu"
# QUEST_LIB_SUFFIX=_dbg (with QUEST_EXTRA_FLAGS) builds an instrumented copy next to the product library
SUF="${QUEST_LIB_SUFFIX:-}"
B="build$SUF"
mkdir -p $B
pids=()
for s in $SRCS; do
  $NVCC $FLAGS -c "$s" -o "$B/${s%.cu}.o" 2> "$B/${s%.cu}.ptxas.log" &
  pids+=($!)
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=1; done
if [ $rc -ne 0 ]; then cat $B/*.ptxas.log >&2; exit 1; fi
OBJS=""
for s in $SRCS; do OBJS="$OBJS $B/${s%.cu}.o"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libquest_b200$SUF.so $OBJS -lcudart
echo "built $(cd .. && pwd)/libquest_b200$SUF.so"
