#!/bin/bash
# s53: A/B of the GQA estimate layouts: group of 4 in registers (1 CTA/SM) vs two lane rows x two q rows (2 CTAs/SM)
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gqa_layouts" 2>&1 | tail -3
for v in 0 2 1 0 2; do
  echo "EST_PER_Q=$v"
  QUEST_EST_PER_Q=$v python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print({k:d[k]['us'] for k in d if k!='shape'})"
done
QUEST_EST_PER_Q=2 python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 32768 --budget 2048 2>&1 | tail -1 | cut -c1-400
QUEST_EST_PER_Q=0 python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 32768 --budget 2048 2>&1 | tail -1 | cut -c1-400
