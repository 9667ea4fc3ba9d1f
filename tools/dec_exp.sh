for dbg in 0 1 2 3; do for ppc in 4 8; do echo "== dbg=$dbg ppc=$ppc"; QUEST_DEC_DBG=$dbg QUEST_DEC_PPC=$ppc timeout 200 python tools/chain_probe.py 2>&1 | grep -E "^decode  "; done; done
