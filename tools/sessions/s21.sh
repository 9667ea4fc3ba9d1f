#!/usr/bin/env bash
# which estimate CTAs finish late (per-blockIdx timeline), and a skewed page partition (more pages to the CTAs that start first)
set -u
mkdir -p gpurun_out
QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s21_timeline.txt
for sk in 0 8 16 24 32 48; do
  BENCH_EST_SKEW=$sk python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s21_bench_$sk.json 2> gpurun_out/r2s21_bench_$sk.err
  python - "$sk" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s21_bench_{v}.json"))
    k = d["kernels"]
    print(f"skew {v:3s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
EST_SKEW=24 QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s21_timeline_skew24.txt
# breakdown of the configs[3] batched step (ncu launch list: shares only)
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2s21_cfg3_launches.csv python tools/bench_configs.py --config 4 --batched 1 --steps 2 --warmup 1 > gpurun_out/r2s21_cfg3_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/r2s21_cfg3_launches.csv')) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); gi = hdr.index('Grid Size')
agg = collections.defaultdict(list)
for r in rows[1:]:
    try: agg[(r[ki][:60], r[gi])].append(float(r[vi].replace(',', '')))
    except Exception: pass
for k, v in agg.items():
    v.sort(); print(k, len(v), 'median ns', v[len(v)//2])
PY
