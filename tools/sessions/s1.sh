#!/usr/bin/env bash
# GPU session 1 (round 2): full GPU suite, prefill parity loop, fused vs unfused layer step
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2s1_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2s1_pytest.log
tail -5 gpurun_out/r2s1_pytest.log
for i in 1 2 3 4 5 6 7 8 9 10; do
  python -m pytest tests/test_vs_ref_kernels.py -q -k "prefill" -x > gpurun_out/r2s1_pf_loop_$i.log 2>&1; echo "prefill loop $i rc=$?"
done
python bench.py --steps 30 --warmup 5 > gpurun_out/r2s1_bench_fused.json 2> gpurun_out/r2s1_bench_fused.err; echo "bench fused rc=$?"
BENCH_UNFUSED=1 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2s1_bench_unfused.json 2> gpurun_out/r2s1_bench_unfused.err; echo "bench unfused rc=$?"
python -c "import json;[print(n, {k:d[k] for k in ('us_per_layer','frac_of_hbm_peak','kernels','eager_us_per_layer')}, d['e2e']['ms_per_step']) for n in ('fused','unfused') for d in [json.load(open(f'gpurun_out/r2s1_bench_{n}.json'))]]"
python __graft_entry__.py smoke 2>&1 | tail -2
