#!/usr/bin/env python
"""One small instance of every product kernel path, through the parity tests' own bodies, for a run under compute-sanitizer:

    compute-sanitizer --tool memcheck --error-exitcode 9 --log-file memcheck.log python tools/sanitize_run.py
    compute-sanitizer --tool racecheck ... python tools/sanitize_run.py --only topk,decode,layer
    compute-sanitizer --tool synccheck ... python tools/sanitize_run.py --only topk

Shapes are the smallest ones of the test lists that still reach each kernel variant (tile / rows estimate, GQA layouts, the three
top-k dispatches incl. the cluster kernels, hand-over and waiting layer step, split-KV decode, batched step, fp8 metadata, fused
producer, prefill kernels, sequence-sharded step, static-scratch decode).  The bodies assert parity as in the suite, so a run also
says whether the sanitizer's serialised execution changed a result.  Prints one line per case and a summary; exit code 1 on any
failed case (the sanitizer's own findings come through its exit code / log).

Status: compute-sanitizer is closed on the GPU pool this repository was developed on (the wrapper there refuses to start), so no
sanitizer run is on record under profiles/; without it the script is a 70-case smoke run of every kernel path (12 s on a B200 box, 0 failures on the round's final build)."""
import argparse
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def cases():
    import quest_b200._kernels as K
    import tests.test_batch_step_gpu as tb
    import tests.test_fp8_metadata as t8
    import tests.test_gpu_parity as tp
    import tests.test_prefill_fused_gpu as tf
    import tests.test_seq_shard_gpu as ts
    import tests.test_zz_no_alloc as tz

    def topk_own(n, k, dist, no_hist, no_cluster):
        with K.debug_option(K.DBG_TOPK_NO_HIST, no_hist), K.debug_option(K.DBG_TOPK_NO_CLUSTER, no_cluster):
            tp._topk_exact(n, k, dist)

    def topk_pub(n, k, dist, no_cluster):
        with K.debug_option(K.DBG_TOPK_NO_CLUSTER, no_cluster):
            tp._topk_from_bin_totals_exact(n, k, dist)

    c = []
    add = lambda group, f, *a: c.append((group, f, a))
    add("append", tp.test_append_multi_round_mixed, 16, 32, 128)
    add("append", tp.test_append_multi_round_mixed, 7, 4, 64)
    add("append", tp.test_append_prefill_then_decode_bitexact, 409, 8, 128, 7)
    add("estimate", tp.test_estimate_bitexact_reference_sweep, 577)
    add("estimate", tp.test_estimate_bitexact_reference_sweep, 2047)
    add("estimate", tp.test_estimate_bitexact_shapes_and_gqa, 32, 8, 128, 16, 1000, K.EST_META_ORDERED)
    add("estimate", tp.test_estimate_bitexact_shapes_and_gqa, 6, 3, 64, 1, 40, 0)
    add("estimate", tp.test_estimate_bitexact_shapes_and_gqa, 8, 2, 128, 7, 150, 0)
    for per_q in (0, 1, 2, 3):
        add("estimate", tp.test_estimate_bitexact_gqa_layouts, *((16, 8, 128, 16, 1300) if per_q < 2 else (32, 8, 128, 16, 5000)), per_q)
    add("estimate", tp.test_estimate_unordered_random_metadata, 8, 2, 64, 3, 77)
    for n, k in ((311, 69), (2047, 127), (4095, 300), (4097, 333), (8193, 511), (16384, 1023), (16385, 64)):
        for dist in ("normal", "ties") if n == 8193 else ("normal",):
            add("topk", topk_own, n, k, dist, 0, 0)
            if 4096 < n <= 16384:
                add("topk", topk_own, n, k, dist, 0, 1)
            if n <= 16384:
                add("topk", topk_own, n, k, dist, 1, 0)
    for n, k in ((2047, 127), (4097, 333), (8193, 511), (12289, 12288), (16384, 1023)):
        for dist in ("normal", "narrow") if n == 8193 else ("normal",):
            add("topk", topk_pub, n, k, dist, 0)
            if n > 4096:
                add("topk", topk_pub, n, k, dist, 1)
    add("decode", tp.test_sparse_decode_random_page_subsets, 577, 64, 8, 8, 64, 16)
    add("decode", tp.test_sparse_decode_random_page_subsets, 1541, 128, 32, 8, 128, 16)
    add("decode", tp.test_sparse_decode_random_page_subsets, 33, 1, 32, 32, 128, 3)
    add("decode", tz.test_static_scratch_decode_equals_workspace_decode, 8, 8, 64, 8, 300, 40, 8)
    add("decode", ts.test_decode_run_negative_ids_and_no_last_page)
    add("decode", tp.test_merge_states)
    add("layer", tp.test_layer_step_matches_oracle, 1011, 31)
    for hist in (True, "nohandoff", False):
        add("layer", tp.test_decode_layer_step_equals_four_operator_calls, 32, 32, 2500, 32, 16, 128, hist)
    add("layer", tp.test_decode_layer_step_equals_four_operator_calls, 32, 8, 3000, 24, 16, 128, True)
    add("layer", tp.test_decode_layer_step_equals_four_operator_calls, 8, 8, 700, 9, 8, 64, True)
    add("layer", tp.test_decode_layer_step_equals_four_operator_calls, 6, 3, 1000, 12, 16, 128, True)
    add("layer", tp.test_layer_step_equals_separate_calls, 16, 9000, 100, 16, True)
    add("layer", tb.test_batch_step_equals_per_sequence_path, (3000, 2047), 32, 8, 128, 16, 40)
    add("layer", t8.test_fp8_pool_and_estimate, 300, 16, 8, 2, 64, 1.0)
    add("layer", t8.test_fp8_layer_step, 2500, 32, 32, 32)
    add("shard", ts.test_sharded_step_equals_single_gpu, 1000, 8, 8, 64, 8, 12, 3)
    add("shard", ts.test_sharded_step_equals_single_gpu, 700, 32, 32, 128, 16, 30, 4)
    add("rope", tp.test_rope, 37, 100, 8, 2, 64)
    add("rope", tp.test_rms_norm)
    add("producer", tf.test_fused_producer_bitexact_vs_three_calls, [5, 300], 8, 2, 128, 16, False)
    add("producer", tf.test_fused_producer_bitexact_vs_three_calls, [130, 2, 40], 4, 4, 64, 16, True)
    for impl in ("tma", "umma", "simt"):
        add("prefill", tp._prefill_case, 130, 333, 4, 4, 128, 16, True, impl)
    add("prefill", tp._prefill_case, 257, 321, 6, 2, 64, 32, True, "")
    return c


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="", help="comma-separated groups: append,estimate,topk,decode,layer,shard,rope,producer,prefill")
    ap.add_argument("--budget-s", type=float, default=1e9, help="stop starting new cases after this many seconds")
    a = ap.parse_args()
    only = set(x for x in a.only.split(",") if x)
    t0 = time.time()
    ran = fails = skipped = 0
    for group, f, args in cases():
        if only and group not in only:
            continue
        if time.time() - t0 > a.budget_s:
            skipped += 1
            continue
        t1 = time.time()
        try:
            f(*args)
            res = "ok"
        except BaseException as e:  # pytest.skip raises too
            if type(e).__name__ == "Skipped":
                res = "skipped by the test body"
            else:
                fails += 1
                res = f"FAIL {type(e).__name__}: {str(e)[:200]}"
                traceback.print_exc(limit=3)
        ran += 1
        print(f"[{group}] {f.__name__}{args}: {res} ({time.time() - t1:.1f} s)", flush=True)
    print(f"sanitize_run: {ran} cases, {fails} failures, {skipped} not started (time budget), {time.time() - t0:.0f} s")
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
