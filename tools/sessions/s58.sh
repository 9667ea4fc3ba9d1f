#!/bin/bash
# s58: A/B cluster (4 CTAs) vs single-CTA hist top-k at 8193..16384-page rows (QUEST_DBG 4=1: no cluster variants)
for dbg in "" "4=1"; do
  echo "QUEST_DBG=$dbg"
  QUEST_DBG=$dbg python tools/rank_emulate.py 2>&1 | tail -1 | cut -c1-120
  QUEST_DBG=$dbg python tools/kernel_breakdown.py --ctx 262144 --budget 8192 --layers 2 --steps 20 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('mha256k', {k:d[k]['us'] for k in d if k!='shape'})"
  QUEST_DBG=$dbg python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096 --layers 2 --steps 20 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('gqa128k', {k:d[k]['us'] for k in d if k!='shape'})"
done
