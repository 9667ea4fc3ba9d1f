#!/usr/bin/env bash
# sweeps on the reference's benchmark axes, then ncu launch list + --set full capture of the layer-step kernels
set -u
mkdir -p gpurun_out
python tools/ref_sweeps.py > gpurun_out/r2s10_ref_sweeps.txt 2> gpurun_out/r2s10_ref_sweeps.err; echo "sweeps rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/r2s10_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'estimate_tile|topk_hist|topk_fast|decode_kernel|append' -c 600 --csv --log-file gpurun_out/r2s10_launches.csv $CMD > gpurun_out/r2s10_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'estimate_tile|topk_hist|decode_kernel' -s 96 -c 6 -o gpurun_out/r2s10_full $CMD > gpurun_out/r2s10_ncu2.log 2>&1
echo "full rc=$?"; ls -la gpurun_out/ | tail -8
