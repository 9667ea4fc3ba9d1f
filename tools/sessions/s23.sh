#!/usr/bin/env bash
# page-id hand-over top-k -> decode (no grid dependency): parity, A/B, timeline
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_full_size_gpu.py tests/test_fp8_metadata.py tests/test_batch_step_gpu.py tests/test_seq_shard_gpu.py -m gpu -x -q -k "not prefill" > gpurun_out/r2s23_pytest.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s23_pytest.log
for rep in 1 2; do
for v in 1 ""; do
  BENCH_NO_HANDOFF=$v timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s23_bench_$v.json 2> gpurun_out/r2s23_bench_$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s23_bench_{v}.json"))
    k = d["kernels"]
    print(f"{'no handoff' if v else 'handoff   '} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f} | e2e {d['e2e']['ms_per_step']:.3f} ms")
except Exception as e:
    print(v, "failed", e)
PY
done
done
QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so timeout 300 python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s23_timeline.txt
timeout 300 python tools/kernel_breakdown.py --hq 32 --hkv 32 --ctx 131072 --budget 8192 2>&1 | tail -1
timeout 300 python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1
