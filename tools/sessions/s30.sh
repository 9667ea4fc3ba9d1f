#!/usr/bin/env bash
# what one rank of the 8-GPU configs[3] run executes (1 kv head / 4 q heads x batch 8), per kernel; and the hand-over stress test
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "handoff_stress" 2>&1 | tail -2
timeout 300 python tools/rank_emulate.py --world 8 2>&1 | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'estimate_tile|topk|decode_kernel' -c 40 --csv --log-file gpurun_out/r2s30_rank8.csv python tools/rank_emulate.py --world 8 --steps 2 --warmup 1 > gpurun_out/r2s30_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/r2s30_rank8.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); gi = hdr.index('Grid Size')
agg = collections.defaultdict(list)
for r in rows[1:]:
    try: agg[(r[ki][:60], r[gi])].append(float(r[vi].replace(',', '')))
    except Exception: pass
for k, v in agg.items():
    v.sort(); print(k, len(v), 'median ns', v[len(v)//2])
PY
