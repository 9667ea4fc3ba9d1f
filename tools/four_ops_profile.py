#!/usr/bin/env python
"""Host-side profile of the reference's four-operator decode sequence (append_kv, decode_estimate, decode_topk,
decode_sparse_attn per layer: quest/models/QuestAttention.py:100-150) through quest_b200.utils at the headline shape:
where an unchanged QuestAttention.py spends its host time on this library.  python tools/four_ops_profile.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["QUEST_E2E_FOUR_OPS"] = "1"
import bench  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    wl = bench.Workload(dev, 1234)
    for _ in range(5):
        wl.e2e_step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 20
    for _ in range(n):
        wl.e2e_step()
        torch.cuda.synchronize()
    print(f"four-operator path: {(time.perf_counter() - t0) / n * 1e3:.3f} ms per token (32 layers)")
    # host-bound or device-bound?  enqueue 20 tokens without waiting: the host's share is the time until the last call returned
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        wl.e2e_step()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"back to back: host enqueue {(t1 - t0) / n * 1e3:.3f} ms per token, device done after {(t2 - t0) / n * 1e3:.3f} ms per token")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(n):
        wl.e2e_step()
    torch.cuda.synchronize()
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats("cumulative").print_stats(28)
    st.sort_stats("tottime").print_stats(18)


if __name__ == "__main__":
    main()
