#!/bin/bash
# s52: ncu --set full of the GQA estimate (Hq 32 / Hkv 8, 8191 scored pages) -- where its 10 us go
python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 --layers 2 --steps 5 2>&1 | tail -2
ncu --set full --clock-control none --import-source on -k regex:estimate_tile -s 6 -c 2 -o gpurun_out/r2s52_gqa_est -f \
  python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 --layers 2 --steps 2 > gpurun_out/r2s52_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/r2s52*
