#!/usr/bin/env bash
# fp8 (e4m3) metadata variant: parity, then estimate / layer step timings against the fp16 pool; configs[3] kernel breakdown
set -u
mkdir -p gpurun_out
python -m pytest tests/test_fp8_metadata.py -x -q > gpurun_out/r2s22_fp8.log 2>&1; echo "fp8 rc=$?"; tail -4 gpurun_out/r2s22_fp8.log
python -m pytest tests/test_gpu_parity.py tests/test_prefill_fused_gpu.py tests/test_batch_step_gpu.py -m gpu -x -q -k "not prefill_vs and not multiwave" > gpurun_out/r2s22_pytest.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s22_pytest.log
python tools/kernel_breakdown.py --hq 32 --hkv 32 --ctx 32768 --budget 2048 2>&1 | tail -1 | tee gpurun_out/r2s22_breakdown.txt
python tools/kernel_breakdown.py --hq 32 --hkv 32 --ctx 131072 --budget 8192 2>&1 | tail -1 | tee -a gpurun_out/r2s22_breakdown.txt
python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1 | tee -a gpurun_out/r2s22_breakdown.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:quest -c 40 --csv --log-file gpurun_out/r2s22_cfg3_launches.csv python tools/bench_configs.py --config 4 --batched 1 --steps 2 --warmup 1 > gpurun_out/r2s22_cfg3_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/r2s22_cfg3_launches.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); gi = hdr.index('Grid Size')
agg = collections.defaultdict(list)
for r in rows[1:]:
    try: agg[(r[ki][:70], r[gi])].append(float(r[vi].replace(',', '')))
    except Exception: pass
for k, v in agg.items():
    v.sort(); print(k, len(v), 'median ns', v[len(v)//2])
PY
