#!/usr/bin/env python
"""Randomised differential runs of the GPU parity tests' own bodies (tests/test_gpu_parity.py, tests/test_prefill_fused_gpu.py,
tests/test_fp8_metadata.py) over shapes the parametrised lists do not name: the layer step (with and without the page-id
hand-over, with and without bin totals) against the four operator calls, the fused prefill producer against the three-call
sequence, the fp8 layer step against its oracle.  --topk N adds N random rows of 2 ... 16384 scores through the two top-k test bodies (quest_topk on self-counted bin
totals, the layer step's selection from published totals) under every dispatch setting (cluster / single CTA / register radix).
python tools/fuzz_parity.py [--n 150] [--topk 0] [--seed 0]"""
import argparse
import os
import random
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=150)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--topk", type=int, default=0)
    a = ap.parse_args()
    import tests.test_fp8_metadata as t8
    import tests.test_gpu_parity as tp
    import tests.test_prefill_fused_gpu as tf
    rnd = random.Random(a.seed)
    fails = 0
    for i in range(a.n):
        D = rnd.choice([64, 128, 128, 128])
        Hkv = rnd.choice([1, 2, 4, 8, 32])
        G = rnd.choice([1, 1, 2, 4, 8])
        Hq = min(Hkv * G, 64)
        while Hq % Hkv:
            Hq -= 1
        P = rnd.choice([8, 16, 16, 16, 32])
        kv_len = rnd.randint(3 * P + 2, 9000)
        n_pages = (kv_len + P - 1) // P
        budget = rnd.randint(2, min(n_pages, 300))
        kind = rnd.choice(["layer", "layer", "producer", "fp8"])
        try:
            if kind == "layer":
                tp.test_decode_layer_step_equals_four_operator_calls(Hq, Hkv, kv_len, budget, P, D, rnd.choice([True, "nohandoff", False]))
            elif kind == "producer":
                rounds = [rnd.randint(2, 700) for _ in range(rnd.randint(1, 3))]
                tf.test_fused_producer_bitexact_vs_three_calls(rounds, Hq, Hkv, D, P if P >= 8 else 16, rnd.choice([False, True]))
            else:
                kv8 = max(kv_len, 40 * 16)  # that test fixes page size 16 and head_dim 128
                if D != 128 or budget > kv8 // 16 - 2:
                    continue
                t8.test_fp8_layer_step(kv8, budget, Hq, Hkv)
        except BaseException as e:  # pytest.skip raises too: report and go on
            if type(e).__name__ == "Skipped":
                continue
            fails += 1
            print(f"FAIL #{i} {kind} Hq={Hq} Hkv={Hkv} D={D} P={P} kv_len={kv_len} budget={budget}: {type(e).__name__}: {str(e)[:200]}")
            traceback.print_exc(limit=3)
    import quest_b200._kernels as K
    for i in range(a.topk):
        n = rnd.choice([rnd.randint(2, 4096), rnd.randint(4097, 8192), rnd.randint(8193, 16384), rnd.choice([4097, 8191, 8192, 8193, 16383, 16384])])
        k = rnd.choice([1, n, rnd.randint(1, n), rnd.randint(1, min(n, 1024))])
        dist = rnd.choice(["normal", "normal", "narrow", "ties", "const"])
        try:
            for no_hist, no_cluster in ((0, 0), (0, 1), (1, 0)):
                with K.debug_option(K.DBG_TOPK_NO_HIST, no_hist), K.debug_option(K.DBG_TOPK_NO_CLUSTER, no_cluster):
                    tp._topk_exact(n, k, dist)
            for no_cluster in (0, 1):
                with K.debug_option(K.DBG_TOPK_NO_CLUSTER, no_cluster):
                    tp._topk_from_bin_totals_exact(n, k, dist)
        except BaseException as e:
            if type(e).__name__ == "Skipped":
                continue
            fails += 1
            print(f"FAIL topk #{i} n={n} k={k} dist={dist}: {type(e).__name__}: {str(e)[:200]}")
            traceback.print_exc(limit=3)
    print(f"fuzz: {a.n} + {a.topk} top-k cases, {fails} failures")
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
