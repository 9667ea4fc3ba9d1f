#!/usr/bin/env bash
# s61: randomised top-k differential runs over every dispatch (cluster / single CTA / register radix), then the full default bench line
set -u
mkdir -p gpurun_out
( time timeout 110 python tools/fuzz_parity.py --n 40 --topk 400 --seed 61 ) > gpurun_out/r2s61_fuzz.log 2>&1; echo "fuzz rc=$?"; tail -3 gpurun_out/r2s61_fuzz.log
( time timeout 270 python bench.py ) > gpurun_out/r2s61_bench.json 2> gpurun_out/r2s61_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r2s61_bench.json; tail -5 gpurun_out/r2s61_bench.err
