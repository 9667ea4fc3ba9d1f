#!/usr/bin/env bash
# GQA estimate through the shared-memory cp.async ring (parity, then timing against the register ring), and an interleaved
# repeat of the headline A/B (previous build / new / new without FHFMA)
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "estimate" > gpurun_out/r2s16_pytest.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s16_pytest.log
for v in _g1 ""; do
  echo "== lib$v"
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 32768 --budget 2048 2>&1 | tail -1
done
for rep in 1 2 3; do
for v in _base "" _nof; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s16_bench$v.json 2> gpurun_out/r2s16_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s16_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'new':8s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
