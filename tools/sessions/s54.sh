#!/usr/bin/env bash
# round-2 final validation after the GQA split layout, the bound four operators and the record updater: GPU suite, bench line, ncu launch list + --set full capture of the chain
set -u
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2s54_pytest.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2s54_pytest.log
( time timeout 900 python bench.py ) > gpurun_out/r2s54_bench.json 2> gpurun_out/r2s54_bench.err; echo "bench rc=$?"
( time timeout 600 python bench.py --impl reference ) > gpurun_out/r2s54_bench_ref.json 2> gpurun_out/r2s54_bench_ref.err; echo "ref rc=$?"; tail -c 600 gpurun_out/r2s54_bench_ref.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/r2s54_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'estimate_tile|topk_hist|topk_fast|decode_kernel|append' -c 600 --csv --log-file gpurun_out/r2s54_launches.csv $CMD > gpurun_out/r2s54_ncu1.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'estimate_tile|topk_hist|decode_kernel' -s 96 -c 6 -o gpurun_out/r2s54_full $CMD > gpurun_out/r2s54_ncu2.log 2>&1
echo "full rc=$?"; ls -la gpurun_out/ | grep r2s54
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'estimate_tile|topk|decode_kernel' -c 40 --csv --log-file gpurun_out/r2s54_cfg3_launches.csv python tools/bench_configs.py --config 4 --batched 1 --steps 2 --warmup 1 > gpurun_out/r2s54_cfg3_ncu.log 2>&1; echo "cfg3 ncu rc=$?"
QUEST_B200_LIB=quest_b200/libquest_b200_dbg.so timeout 300 python tools/step_timeline.py 2>&1 | grep -v Warning | tee gpurun_out/r2s54_timeline.txt
for v in 0 3; do QUEST_EST_PER_Q=$v python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('gqa128k layout $v', {k:d[k]['us'] for k in d if k!='shape'})"; done
