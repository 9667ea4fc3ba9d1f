#!/usr/bin/env bash
# first ring of the estimate pinned before the launch wait (non-.nc loads), ordered q / id loads; fused RoPE + append + metadata
# producer and the prefill layer step: parity, headline A/B against the round-2 HEAD build, producer timing at 32K
set -u
mkdir -p gpurun_out
python -m pytest tests/test_prefill_fused_gpu.py -m gpu -x -q > gpurun_out/r2s19_fused.log 2>&1; echo "fused rc=$?"; tail -4 gpurun_out/r2s19_fused.log
python -m pytest tests/test_gpu_parity.py tests/test_full_size_gpu.py tests/test_batch_step_gpu.py tests/test_vs_ref_kernels.py -m gpu -x -q > gpurun_out/r2s19_pytest.log 2>&1; echo "parity rc=$?"; tail -3 gpurun_out/r2s19_pytest.log
for rep in 1 2; do
for v in _base ""; do
  QUEST_B200_LIB=quest_b200/libquest_b200$v.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2s19_bench$v.json 2> gpurun_out/r2s19_bench$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/r2s19_bench{v}.json"))
    k = d["kernels"]
    print(f"{v or 'new':8s} us/layer {d['us_per_layer']:.2f} frac {d['frac_of_hbm_peak']:.3f} | est {k['estimate']['us']:.2f} topk {k['topk']['us']:.2f} dec {k['decode']['us']:.2f}")
except Exception as e:
    print(v, "failed", e)
PY
done
done
python tools/prefill_producer_bench.py 2>&1 | tee gpurun_out/r2s19_producer.txt
