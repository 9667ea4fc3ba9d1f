#!/usr/bin/env bash
# s60: the GPU suite as the round's end runs it, on the final build (cluster bin-total top-k included), then the headline bench line
set -u
mkdir -p gpurun_out
( time timeout 400 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2s60_pytest.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2s60_pytest.log
( time timeout 150 python bench.py --no-cpu-baseline --no-extras ) > gpurun_out/r2s60_bench.json 2> gpurun_out/r2s60_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r2s60_bench.json
