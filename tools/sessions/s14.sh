#!/usr/bin/env bash
# mixed-precision FMA (FHFMA) in estimate + decode, GQA groups of 4 scored from registers: parity then A/B against the previous build
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_vs_ref_kernels.py tests/test_full_size_gpu.py -m gpu -x -q -k "not prefill" > gpurun_out/r2s14_pytest.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s14_pytest.log
for v in _base "" _m8; do
  echo "== lib$v"
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 32 --ctx 32768 --budget 2048 2>&1 | tail -6
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -6
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 32 --ctx 131072 --budget 8192 2>&1 | tail -6
done
for v in _base "" _m8; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s14_bench$v.json 2> gpurun_out/r2s14_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s14_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'new':6s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
