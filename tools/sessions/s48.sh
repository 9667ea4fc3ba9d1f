#!/usr/bin/env bash
# prefill (row a9): the case that failed on the round-1 driver box and its neighbours, 30 consecutive fresh processes; then the
# prefill timing tool, which asserts <= 2e-3 against the reference kernel at every size it times
set -u
mkdir -p gpurun_out
ok=0
for i in $(seq 1 30); do
  if timeout 300 python -m pytest tests/test_vs_ref_kernels.py -m gpu -x -q -k "prefill and (4096 or 8192 or 1000)" > gpurun_out/r2s48_run.log 2>&1; then ok=$((ok+1)); else echo "run $i FAILED"; tail -5 gpurun_out/r2s48_run.log; fi
done
echo "prefill vs reference kernel: $ok of 30 consecutive runs green" | tee gpurun_out/r2s48_prefill_30.txt
tail -1 gpurun_out/r2s48_run.log | tee -a gpurun_out/r2s48_prefill_30.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "prefill" 2>&1 | tail -1 | tee -a gpurun_out/r2s48_prefill_30.txt
timeout 900 python tools/prefill_bench.py 2>&1 | tail -8 | tee -a gpurun_out/r2s48_prefill_30.txt
