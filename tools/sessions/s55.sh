#!/bin/bash
# s55: probe -- the layer step as S parallel head-group chains (stand-in problems of 32/S heads)
for s in 1 2 4; do timeout 300 python tools/experiments/two_chain_probe.py --split $s 2>&1 | tail -1; done
