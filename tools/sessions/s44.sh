#!/usr/bin/env bash
# bin-total top-k with 1024 threads per row for launches of few long rows: parity, then the shapes it touches
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_batch_step_gpu.py tests/test_seq_shard_gpu.py tests/test_full_size_gpu.py -m gpu -x -q -k "topk or layer_step or batch or shard or full" 2>&1 | tail -2
for shape in "--hq 32 --hkv 32 --ctx 131072 --budget 8192" "--hq 32 --hkv 8 --ctx 131072 --budget 4096" "--hq 32 --hkv 32 --ctx 65536 --budget 4096"; do
  timeout 300 python tools/kernel_breakdown.py $shape 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["shape"]["pages"], "layer", d["layer_step"]["us"], "fp8", d["layer_step_fp8"]["us"])'
done
timeout 300 python tools/rank_emulate.py --world 8 2>&1 | tail -1
timeout 600 python tools/seqshard_emulate.py --world 8 --rank 7 2>&1 | grep -v Warning | tail -6
for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("us/layer", round(d["us_per_layer"],2))'; done
