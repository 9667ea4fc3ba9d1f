#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python tools/four_ops_profile.py > gpurun_out/r2s36_four_ops.txt 2>&1; echo rc=$?; head -3 gpurun_out/r2s36_four_ops.txt; sed -n 8,30p gpurun_out/r2s36_four_ops.txt
for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("us/layer", round(d["us_per_layer"],2), "e2e ms", round(d["e2e"]["ms_per_step"],3))'; done
QUEST_E2E_FOUR_OPS=1 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("four ops: e2e ms", round(d["e2e"]["ms_per_step"],3))'
timeout 600 python -m pytest tests/test_reference_python_on_shim.py tests/test_gpu_parity.py -m gpu -x -q -k "shim or operator or surface or golden" 2>&1 | tail -2
