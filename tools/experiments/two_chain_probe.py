#!/usr/bin/env python
"""Probe: would a layer step split into S head groups, each its own estimate -> top-k -> decode chain on its own stream, fill the
HBM idle time of the single chain (~6 us of dependent latency per 18 us step)?  Stand-in: S independent problems
of 32/S heads each (own pools: 4 KB rows instead of 4 KB pieces of 8 KB rows), same total bytes as the headline, as parallel
branches of one CUDA graph -- joined after every layer (what an entry point with an internal fork/join would do) or free-running.
    python tools/experiments/two_chain_probe.py --split 2"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402
import quest_b200.utils as qu  # noqa: E402
from tools.bench_configs import fill, timed  # noqa: E402

P, D = 16, 128


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--split", type=int, default=2)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=32768)
    ap.add_argument("--budget", type=int, default=2048)
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib = K.load_library()
    S, L = a.split, a.layers
    H = a.heads // S
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    groups = []
    for g in range(S):
        ic = qu.InferenceController(L, H, D, P, a.budget // P, a.ctx, torch.float16, dev)
        fill(ic, L, a.ctx - 1, H, dev)
        ic.prepare_metadata(1)
        ic.begin_forward(1)
        n_pages = len(ic.kv_cache.indicies)
        n, k = n_pages - 1, ic.inference_page_budget - 1
        ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
        assert lib.quest_decode_plan(k, H, H, D, P, sms, ctypes.byref(ppc), ctypes.byref(nch)) == 0
        groups.append(dict(ic=ic, n_pages=n_pages, n=n, k=k, ppc=ppc.value,
                           q=torch.randn(L, 1, H, D, device=dev).half(), kn=torch.randn(L, 1, H, D, device=dev).half(),
                           vn=torch.randn(L, 1, H, D, device=dev).half(), o=torch.empty(L, 1, H, D, dtype=torch.float16, device=dev),
                           scores=torch.empty(L, H, n, dtype=torch.float16, device=dev),
                           sel=torch.zeros(L, H, k, dtype=torch.int32, device=dev),
                           ws=torch.zeros(lib.quest_decode_workspace_bytes(H, D, nch.value), dtype=torch.uint8, device=dev)))

    def step(g, l, stream):
        ic = g["ic"]
        rc = lib.quest_decode_layer_step(g["q"][l].data_ptr(), g["kn"][l].data_ptr(), g["vn"][l].data_ptr(), g["o"][l].data_ptr(),
                                         ic.kv_cache.buf_layer(l).data_ptr(), ic.kv_indices_with_last.data_ptr(), g["n_pages"],
                                         ic.kv_last_page_idx, ic.kv_cache.last_page_len, ic.metadata_cache.buf_layer(l).data_ptr(),
                                         ic.metadata_indices.data_ptr(), ic.metadata_indices.numel(), ic.metadata_cache.last_page_len,
                                         g["scores"][l].data_ptr(), g["sel"][l].data_ptr(), g["k"], H, H, D, P, g["ppc"],
                                         g["ws"].data_ptr(), g["ws"].numel(), stream.cuda_stream)
        assert rc == 0, rc

    side = [torch.cuda.Stream() for _ in range(S)]

    def run(join_every_layer):
        main_s = torch.cuda.current_stream()
        if join_every_layer:
            for l in range(L):
                for s_, g in zip(side, groups):
                    s_.wait_stream(main_s)
                    step(g, l, s_)
                for s_ in side:
                    main_s.wait_stream(s_)
        else:
            for s_, g in zip(side, groups):
                s_.wait_stream(main_s)
                for l in range(L):
                    step(g, l, s_)
            for s_ in side:
                main_s.wait_stream(s_)

    res = {"split": S, "heads_per_chain": H, "plan": {"k": groups[0]["k"], "ppc": groups[0]["ppc"]}}
    for name, j in (("joined_per_layer", True), ("free_running", False)):
        cap = torch.cuda.Stream()
        cap.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(cap):
            for _ in range(2):
                run(j)
        torch.cuda.current_stream().wait_stream(cap)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            run(j)
        res[name + "_us_per_layer"] = round(timed(gr.replay, a.steps, 5, False) / L * 1e6, 2)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
