#!/bin/bash
# s51: bound four-operator path: test, host profile, bench e2e (four ops and one call)
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "four_operators_bound or equals_four_ops or layer_step_matches" 2>&1 | tail -5
python -m pytest tests/test_reference_python_on_shim.py tests/test_fp8_metadata.py tests/test_batch_step_gpu.py -x -q -m gpu 2>&1 | tail -3
python tools/four_ops_profile.py 2>&1 | head -45 | cut -c1-160 > gpurun_out/r2s51_four_ops.txt; head -1 gpurun_out/r2s51_four_ops.txt
QUEST_E2E_FOUR_OPS=1 python bench.py --steps 20 --warmup 5 2>gpurun_out/r2s51_b4.err | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('four ops e2e', d['e2e'], d['ms_per_step'])"
python bench.py --steps 20 --warmup 5 2>gpurun_out/r2s51_b1.err | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('one call e2e', d['e2e'], d['ms_per_step'])"
