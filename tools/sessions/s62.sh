#!/usr/bin/env bash
# s62: compute-sanitizer over one small instance of every product kernel path (tools/sanitize_run.py): memcheck on everything,
# racecheck (shared-memory hazards) on the top-k and decode kernels
set -u
mkdir -p gpurun_out
SAN=$(command -v compute-sanitizer || echo /usr/local/cuda/bin/compute-sanitizer)
export PYTORCH_NO_CUDA_MEMORY_CACHING=1  # every tensor its own cudaMalloc: out-of-bounds accesses are seen per tensor
( time timeout 175 $SAN --tool memcheck --error-exitcode 9 --print-limit 30 --log-file gpurun_out/r2s62_memcheck.log python tools/sanitize_run.py --budget-s 125 ) > gpurun_out/r2s62_memcheck_run.log 2>&1; echo "memcheck rc=$?"
tail -3 gpurun_out/r2s62_memcheck_run.log; tail -4 gpurun_out/r2s62_memcheck.log
( time timeout 90 $SAN --tool racecheck --error-exitcode 9 --print-limit 30 --log-file gpurun_out/r2s62_racecheck.log python tools/sanitize_run.py --only topk,decode --budget-s 50 ) > gpurun_out/r2s62_racecheck_run.log 2>&1; echo "racecheck rc=$?"
tail -3 gpurun_out/r2s62_racecheck_run.log; tail -4 gpurun_out/r2s62_racecheck.log
