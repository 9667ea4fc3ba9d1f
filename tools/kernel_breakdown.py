#!/usr/bin/env python
"""Per-kernel device time of one layer step at an arbitrary shape (bench.py does this for the headline shape only):
estimate, top-k, decode alone and the chained layer step, each as a CUDA graph over `--layers` rotated layer pools.
    python tools/kernel_breakdown.py --hq 32 --hkv 8 --ctx 131072 --budget 4096"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402
import quest_b200.utils as qu  # noqa: E402
from tools.bench_configs import fill, timed, try_graph  # noqa: E402

P, D = 16, 128


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=32768)
    ap.add_argument("--budget", type=int, default=2048, help="token budget")
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--steps", type=int, default=50)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib = K.load_library()
    if os.environ.get("QUEST_NO_HANDOFF"):  # A/B: the layer step's decode launch waits for the top-k launch
        lib.quest_debug_set_option(K.DBG_NO_HANDOFF, 1)
    if os.environ.get("QUEST_EST_PER_Q"):  # A/B of the GQA estimate layouts (include/quest_b200.h, QUEST_DBG_EST_PER_Q)
        lib.quest_debug_set_option(K.DBG_EST_PER_Q, int(os.environ["QUEST_EST_PER_Q"]))
    from tools.bench_configs import apply_debug_env
    apply_debug_env()
    Hq, Hkv, L = a.hq, a.hkv, a.layers
    ic = qu.InferenceController(L, Hq, D, P, a.budget // P, a.ctx, torch.float16, dev, num_kv_heads=Hkv)
    fill(ic, L, a.ctx - 1, Hkv, dev)
    ic.prepare_metadata(1)
    ic.begin_forward(1)
    n_pages = len(ic.kv_cache.indicies)
    n, k = n_pages - 1, ic.inference_page_budget - 1
    q = torch.randn(L, 1, Hq, D, device=dev).half()
    kn = torch.randn(L, 1, Hkv, D, device=dev).half()
    vn = torch.randn(L, 1, Hkv, D, device=dev).half()
    o = torch.empty(L, 1, Hq, D, dtype=torch.float16, device=dev)
    scores = torch.empty(L, Hq, n, dtype=torch.float16, device=dev)
    sel = torch.zeros(L, Hq, k, dtype=torch.int32, device=dev)
    ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    assert lib.quest_decode_plan(k, Hq, Hkv, D, P, sms, ctypes.byref(ppc), ctypes.byref(nch)) == 0
    ws = torch.zeros(lib.quest_decode_workspace_bytes(Hq, D, nch.value), dtype=torch.uint8, device=dev)
    kvi, mi = ic.kv_indices_with_last, ic.metadata_indices
    last_idx, last_len, mlast = ic.kv_last_page_idx, ic.kv_cache.last_page_len, ic.metadata_cache.last_page_len

    def st():
        return torch.cuda.current_stream().cuda_stream

    def f_est():
        for l in range(L):
            assert lib.quest_estimate(q[l].data_ptr(), scores[l].data_ptr(), n, ic.metadata_cache.buf_layer(l).data_ptr(),
                                      mi.data_ptr(), Hq, Hkv, D, P, 3, st()) == 0

    def f_topk():
        for l in range(L):
            assert lib.quest_topk(scores[l].data_ptr(), kvi.data_ptr(), 0, None, sel[l].data_ptr(), Hq, n, k, st()) == 0

    def f_dec():
        for l in range(L):
            assert lib.quest_decode_run(q[l].data_ptr(), o[l].data_ptr(), None, ic.kv_cache.buf_layer(l).data_ptr(),
                                        sel[l].data_ptr(), k, k, last_idx, last_len, Hq, Hkv, D, P, ppc.value,
                                        ws.data_ptr(), ws.numel(), st()) == 0

    def f_layer():
        for l in range(L):
            assert lib.quest_decode_layer_step(q[l].data_ptr(), kn[l].data_ptr(), vn[l].data_ptr(), o[l].data_ptr(),
                                               ic.kv_cache.buf_layer(l).data_ptr(), kvi.data_ptr(), n_pages, last_idx, last_len,
                                               ic.metadata_cache.buf_layer(l).data_ptr(), mi.data_ptr(), mi.numel(), mlast,
                                               scores[l].data_ptr(), sel[l].data_ptr(), k, Hq, Hkv, D, P, ppc.value,
                                               ws.data_ptr(), ws.numel(), st()) == 0

    # flag-gated fp8 metadata (section 8 row f-4): e4m3 copy of every layer's metadata pool, estimate / layer step over it
    meta8 = [torch.zeros(ic.metadata_cache.buf_layer(l).shape, dtype=torch.uint8, device=dev) for l in range(L)]
    for l in range(L):
        assert lib.quest_metadata_to_fp8(ic.metadata_cache.buf_layer(l).data_ptr(), mi.data_ptr(), 0, n_pages, meta8[l].data_ptr(),
                                         Hkv, D, P, st()) == 0

    def f_est8():
        for l in range(L):
            assert lib.quest_estimate_fp8(q[l].data_ptr(), scores[l].data_ptr(), n, meta8[l].data_ptr(), mi.data_ptr(), Hq, Hkv, D, P,
                                          1, st()) == 0

    def f_layer8():
        for l in range(L):
            assert lib.quest_decode_layer_step_fp8(q[l].data_ptr(), kn[l].data_ptr(), vn[l].data_ptr(), o[l].data_ptr(),
                                                   ic.kv_cache.buf_layer(l).data_ptr(), kvi.data_ptr(), n_pages, last_idx, last_len,
                                                   ic.metadata_cache.buf_layer(l).data_ptr(), meta8[l].data_ptr(), mi.data_ptr(),
                                                   mi.numel(), mlast, scores[l].data_ptr(), sel[l].data_ptr(), k, Hq, Hkv, D, P,
                                                   ppc.value, ws.data_ptr(), ws.numel(), st()) == 0

    by = {"estimate": n * Hkv * D * 4 + Hq * n * 2, "topk": Hq * n * 2 + Hq * k * 4, "decode": Hq * (k + 1) * P * 2 * D * 2}
    by["layer_step"] = sum(by.values())
    by["estimate_fp8"] = n * Hkv * D * 2 + Hq * n * 2
    by["layer_step_fp8"] = by["layer_step"] - n * Hkv * D * 2
    res = {"shape": {"Hq": Hq, "Hkv": Hkv, "ctx": a.ctx, "pages": n_pages, "k": k, "ppc": ppc.value, "chunks": nch.value}}
    for name, fn in (("estimate", f_est), ("topk", f_topk), ("decode", f_dec), ("layer_step", f_layer), ("estimate_fp8", f_est8),
                     ("layer_step_fp8", f_layer8)):
        fn()
        torch.cuda.synchronize()
        g = try_graph(fn)
        us = timed(g, a.steps, 5, False) / L * 1e6
        res[name] = {"us": round(us, 2), "GBps": round(by[name] / us / 1e3, 1), "bytes": by[name]}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
