#!/usr/bin/env bash
# ncu --set full of the TMA + tcgen05 prefill kernel and of the fused RoPE + append producer (8K tokens)
set -u
mkdir -p gpurun_out
timeout 600 python tools/prefill_bench.py 8192 > gpurun_out/r2s49_plain.log 2>&1; echo "plain rc=$?"; tail -2 gpurun_out/r2s49_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'prefill_tma_kernel' -s 3 -c 1 -o gpurun_out/r2s49_prefill python tools/prefill_bench.py 8192 > gpurun_out/r2s49_ncu.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --set full --clock-control none -k regex:'rope_append_prefill_kernel' -s 2 -c 1 -o gpurun_out/r2s49_producer python tools/prefill_producer_bench.py --lens 8192 > gpurun_out/r2s49_ncu2.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out | grep r2s49
