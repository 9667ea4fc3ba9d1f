#!/usr/bin/env python
"""Randomised differential runs of the GPU parity tests' own bodies (tests/test_gpu_parity.py, tests/test_prefill_fused_gpu.py,
tests/test_fp8_metadata.py) over shapes the parametrised lists do not name: the layer step (with and without the page-id
hand-over, with and without bin totals) against the four operator calls, the fused prefill producer against the three-call
sequence, the fp8 layer step against its oracle.  python tools/fuzz_parity.py [--n 150] [--seed 0]"""
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
    print(f"fuzz: {a.n} cases, {fails} failures")
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
