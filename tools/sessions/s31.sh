#!/usr/bin/env bash
# groups-of-4 estimate: ring of 8 stages with PRMT masks against the ring of 4 (one CTA per SM either way)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_fp8_metadata.py tests/test_batch_step_gpu.py -m gpu -x -q -k "estimate or fp8 or batch" 2>&1 | tail -2
for v in _r4 ""; do
  echo "== lib$v"
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so timeout 300 python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("estimate", d["estimate"]["us"], "layer", d["layer_step"]["us"], "est fp8", d["estimate_fp8"]["us"])'
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so timeout 300 python tools/bench_configs.py --config 4 --batched 1 --steps 20 --warmup 5 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("configs[3]", d["us_per_layer"], d["frac_of_hbm_peak_per_gpu"])'
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so timeout 300 python tools/rank_emulate.py --world 8 2>&1 | tail -1
done
