#!/usr/bin/env bash
set -u
python -m pytest tests -m gpu -x -q > gpurun_out/r2s5_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2s5_pytest.log
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2s5_bench.json 2> gpurun_out/r2s5_bench.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('gpurun_out/r2s5_bench.json'));print({k:d[k] for k in ('us_per_layer','frac_of_hbm_peak','eager_us_per_layer','kernels')}, d['e2e']['ms_per_step'])"
QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s5_timeline.txt
