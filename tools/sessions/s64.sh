#!/usr/bin/env bash
# s64: ncu --set full of the bin-total top-k over a 4-CTA cluster (GQA 128K shape: 32 q heads, 8,191 scored pages, budget 4096) and of
# its single-CTA counterpart (QUEST_DBG 4=1), each after the same command has exited 0 without ncu
set -u
mkdir -p gpurun_out
CMD="python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 --layers 2 --steps 3"
for dbg in "" "4=1"; do
  tag=$([ -z "$dbg" ] && echo cluster || echo single)
  QUEST_DBG=$dbg $CMD 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$tag', {k:d[k]['us'] for k in d if k!='shape'})"
  QUEST_DBG=$dbg timeout 100 ncu --set full --clock-control none --import-source on -k regex:topk_hist -s 4 -c 2 -o gpurun_out/r2s64_topk_$tag -f $CMD > gpurun_out/r2s64_ncu_$tag.log 2>&1
  echo "ncu $tag rc=$?"
done
ls -la gpurun_out | grep r2s64
