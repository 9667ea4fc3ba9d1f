#!/usr/bin/env bash
# L2 prefetch up front (bulk rows in the estimate before the launch wait; per-line K/V rows in decode): parity, then A/B
# pf = both, pfe = estimate only, pfd = decode only, pf0 = neither (new source), base = round-2 HEAD build
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_full_size_gpu.py tests/test_batch_step_gpu.py -m gpu -x -q -k "not prefill" > gpurun_out/r2s18_pytest.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s18_pytest.log
for rep in 1 2; do
for v in _base _pf0 _pfe _pfd ""; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s18_bench$v.json 2> gpurun_out/r2s18_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s18_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'pf':8s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
for v in _base ""; do
  echo "== lib$v"
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python tools/kernel_breakdown.py --hq 32 --hkv 32 --ctx 131072 --budget 8192 2>&1 | tail -1
done
