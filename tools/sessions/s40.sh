#!/usr/bin/env bash
# where the sequence-sharded step's time goes on one rank of 8 (all ranks emulated in one process, exchanges stood in for)
set -u
mkdir -p gpurun_out
for r in 0 7; do timeout 600 python tools/seqshard_emulate.py --world 8 --rank $r 2>&1 | grep -v Warning | tail -7; done
timeout 600 python tools/seqshard_emulate.py --world 1 --rank 0 2>&1 | grep -v Warning | tail -7
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'estimate|topk|decode_kernel|merge' -c 400 --csv --log-file gpurun_out/r2s40_rank7.csv python tools/seqshard_emulate.py --world 8 --rank 7 --steps 2 > gpurun_out/r2s40_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows = [r for r in csv.reader(open('gpurun_out/r2s40_rank7.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); gi = hdr.index('Grid Size')
for r in rows[-12:]:
    print(r[ki][:60], r[gi], r[vi])
PY
