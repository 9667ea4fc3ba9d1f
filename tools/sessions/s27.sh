#!/usr/bin/env bash
# bin-total top-k for rows of up to 16384 pages (configs[3]'s row bound): parity of the batched step, then configs[3]
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_batch_step_gpu.py tests/test_gpu_parity.py -m gpu -x -q -k "batch or topk or layer_step" > gpurun_out/r2s27_pytest.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r2s27_pytest.log
timeout 600 python tools/bench_configs.py --config 4 --batched 1 --steps 20 --warmup 5 2>&1 | tail -3
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'estimate_tile|topk|decode_kernel' -c 40 --csv --log-file gpurun_out/r2s27_cfg3_launches.csv python tools/bench_configs.py --config 4 --batched 1 --steps 2 --warmup 1 > gpurun_out/r2s27_cfg3_ncu.log 2>&1; echo "cfg3 ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/r2s27_cfg3_launches.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); gi = hdr.index('Grid Size')
agg = collections.defaultdict(list)
for r in rows[1:]:
    try: agg[(r[ki][:60], r[gi])].append(float(r[vi].replace(',', '')))
    except Exception: pass
for k, v in agg.items():
    v.sort(); print(k, len(v), 'median ns', v[len(v)//2])
PY
