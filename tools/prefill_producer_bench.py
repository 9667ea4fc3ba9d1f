#!/usr/bin/env python
"""Prompt-side producer at the headline shape (32 heads x 128, page 16): the reference's sequence apply_rope_in_place ->
append_kv_cache_prefill (two launches, k read twice and written once more) against quest_rope_append_prefill (one launch),
and the whole prompt layer (producer + causal prefill attention) as three operator calls against quest_prefill_layer_step.
Device times with CUDA events; pools are rewritten every iteration (inputs 0.8 GB at 32K, larger than L2)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200._kernels as K  # noqa: E402
import quest_b200.utils as qu  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lens", default="2048,8192,32768")
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=32)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    D, P = 128, 16
    for n in [int(x) for x in a.lens.split(",")]:
        ic = qu.InferenceController(1, a.hq, D, P, 2048 // P, n + 64, torch.float16, dev, num_kv_heads=a.hkv)
        ic.prepare_metadata(n), ic.begin_forward(n)
        q = torch.randn(n, a.hq, D, device=dev).half()
        k = torch.randn(n, a.hkv, D, device=dev).half()
        v = torch.randn(n, a.hkv, D, device=dev).half()
        kv, meta = ic.kv_cache.buf_layer(0), ic.metadata_cache.buf_layer(0)
        kvi, mi, last = ic.kv_indices_with_last, ic.metadata_indices, ic.kv_cache.last_page_len

        def two_calls():
            qu.apply_rope_in_place(q, k, 0)
            qu.append_kv(k, v, ic, 0)

        def fused():
            K.rope_append_kv_cache_prefill(q, k, v, kv, kvi, last, meta, mi, 1.0, 1e4)

        def three_ops():
            two_calls()
            return qu.prefill_forward(q, ic, 0)

        def layer():
            return qu.prefill_layer_step(q, k, v, ic, 0)

        iters = 20 if n <= 8192 else 4
        # interleaved rounds, medians: the 32K attention kernel runs for milliseconds at full tensor load and the clocks
        # settle over the first rounds, so back-to-back blocks of one variant would favour whichever ran first
        r2, r1, r3, rl = [], [], [], []
        for _ in range(5):
            r2.append(timed(two_calls, iters)), r1.append(timed(fused, iters))
            r3.append(timed(three_ops, iters)), rl.append(timed(layer, iters))
        t2, t1, t3, tl = (sorted(x)[len(x) // 2] for x in (r2, r1, r3, rl))
        moved2 = (q.numel() * 2 + k.numel() * 2) * 2 + (k.numel() + v.numel()) * 2 * 2  # rope rw of q,k + append r/w of k,v
        moved1 = q.numel() * 2 * 2 + (k.numel() + v.numel()) * 2 * 2                      # q rw + k,v read, pool write
        print(f"n={n:6d} Hq={a.hq} Hkv={a.hkv}: rope+append {t2:9.1f} us ({moved2 / t2 / 1e3:6.0f} GB/s) | fused producer {t1:9.1f} us "
              f"({moved1 / t1 / 1e3:6.0f} GB/s) {t2 / t1:4.2f}x | layer: three ops {t3 / 1e3:8.3f} ms, one call {tl / 1e3:8.3f} ms "
              f"({t3 / tl:4.3f}x)")
        ic.end_forward()
        del ic, q, k, v
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
