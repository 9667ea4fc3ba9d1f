#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for i in 1 2 3; do timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("us/layer", round(d["us_per_layer"],2), "ms/step", round(d["ms_per_step"],3), "e2e ms", round(d["e2e"]["ms_per_step"],3), "e2e GB/s", round(d["e2e"]["value"]))'; done
timeout 900 python -m pytest tests/test_batch_step_gpu.py tests/test_seq_shard_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/e2e_textgen.py --graph 2>&1 | tail -3
