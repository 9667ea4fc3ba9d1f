#!/usr/bin/env bash
set -u
python -m pytest tests/test_gpu_parity.py -q -x -k "layer_step" 2>&1 | tail -3
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2s3_bench.json 2> gpurun_out/r2s3_bench.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('gpurun_out/r2s3_bench.json'));print({k:d[k] for k in ('us_per_layer','frac_of_hbm_peak','eager_us_per_layer')}, d['e2e']['ms_per_step'])"
QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s3_timeline_fused.txt
