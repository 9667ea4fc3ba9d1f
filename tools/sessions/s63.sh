#!/usr/bin/env bash
# s63: speculative K/V prefetch by the waiting decode CTAs (QUEST_DBG_SPEC_PREFETCH): parity with the switch on, A/B of the headline,
# DRAM bytes of the decode launch with and without it (does a demand load ride on a prefetch in flight, or fetch the line again?)
set -u
mkdir -p gpurun_out
( time timeout 120 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "speculative or handoff_stress" ) > gpurun_out/r2s63_parity.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s63_parity.log
us() { python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['us_per_layer'], d['kernels']['decode']['us'])"; }
for rep in 1 2; do
  for sp in 0 1; do
    echo -n "spec=$sp rep=$rep us_per_layer, decode alone: "
    BENCH_SPEC_PREFETCH=$sp timeout 100 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras 2>gpurun_out/r2s63_bench_${sp}_${rep}.err | tee gpurun_out/r2s63_bench_${sp}_${rep}.json | us
  done
done
for sp in 0 1; do
  BENCH_SPEC_PREFETCH=$sp timeout 120 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none -k regex:decode_kernel -s 40 -c 6 --csv --log-file gpurun_out/r2s63_ncu_spec$sp.csv python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-extras > gpurun_out/r2s63_ncu_spec$sp.log 2>&1
  echo "ncu spec=$sp rc=$?"; grep -E "dram__bytes_read|gpu__time" gpurun_out/r2s63_ncu_spec$sp.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | head -6
done
