#!/usr/bin/env python
"""Multi-GPU workloads of BASELINE.json beyond the headline one (bench.py measures configs[1]):

  --config 4   Llama-3.1-8B GQA shape (32 q / 8 kv heads, D 128), 128K context, token budget 4096, batch 8,
               KV-HEAD sharded over the ranks: rank r holds kv heads [r*8/R, (r+1)*8/R) of every sequence and their
               q heads; no collective on the path.  Strong scaling: total work fixed.
  --config 5   Yarn-Llama2-7B-128K shape (32 heads MHA, D 128), token budget 8192, SEQUENCE sharded: every rank holds
               8192 pages (128K tokens) of one sequence, so R = 8 is the 1M-token context of configs[4]; smaller R run
               the same per-GPU share of a shorter context (weak scaling).  Two all-gathers per layer step
               (quest_b200.parallel.SeqShardedDecode): --exchange nccl | peer (quest_peer_allgather over symmetric memory).
  --batched G  (config 4) G groups of sequences, each one descriptor-driven batched launch chain (BatchDecodeStep).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_configs.py --config 5
Prints one JSON line per run on rank 0.  Times are CUDA-event device times, max over ranks; a step walks `--layers`
layer pools so that consecutive launches never find their inputs in L2."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quest_b200.utils as qu  # noqa: E402
from quest_b200 import parallel as par  # noqa: E402

P, D = 16, 128


def peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6650.0


def apply_debug_env():
    """QUEST_DBG="4=1,7=2": library A/B switches (include/quest_b200.h QUEST_DBG_*) for the tools; the library itself never reads
    the environment."""
    import quest_b200._kernels as K
    for kv in filter(None, os.environ.get("QUEST_DBG", "").split(",")):
        o, v = kv.split("=")
        K.load_library().quest_debug_set_option(int(o), int(v))


def timed(fn, steps, warmup, dist_on):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) / 1e3 / steps
    if dist_on:
        t = torch.tensor([sec], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
    return sec


def try_graph(fn):
    try:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                fn()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return g.replay
    except Exception as e:  # NCCL capture can be refused by the runtime: report eager only
        sys.stderr.write(f"graph capture failed: {e}\n")
        return None


def fill(ic, layers, tokens, Hkv, dev, chunk=32768):
    """Prompt of `tokens` N(0,1) rows per layer through append_kv, in chunks (bounded scratch)."""
    done = 0
    while done < tokens:
        n = min(chunk, tokens - done)
        ic.prepare_metadata(n)
        ic.begin_forward(n)
        for l in range(layers):
            k = torch.randn(n, Hkv, D, dtype=torch.float16, device=dev)
            v = torch.randn(n, Hkv, D, dtype=torch.float16, device=dev)
            qu.append_kv(k, v, ic, l)
        ic.end_forward()
        done += n


def config4(args, rank, world, dev):
    Hq_all, Hkv_all, ctx, budget_tok, batch, L = 32, 8, 131072, 4096, 8, args.layers
    hs, he = par.head_shard(Hkv_all, world, rank)
    Hkv, Hq = he - hs, (he - hs) * (Hq_all // Hkv_all)
    budget = budget_tok // P
    ics = []
    for b in range(batch):
        torch.manual_seed(1234 + 4 + 100 * b + rank)
        ic = qu.InferenceController(L, Hq, D, P, budget, ctx, torch.float16, dev, num_kv_heads=Hkv)
        fill(ic, L, ctx - 1, Hkv, dev)
        ic.prepare_metadata(1)
        ic.begin_forward(1)
        assert ic.need_estimate()
        ics.append(ic)
    q = torch.randn(batch, L, 1, Hq, D, device=dev).half()
    kn = torch.randn(batch, L, 1, Hkv, D, device=dev).half()
    vn = torch.randn(batch, L, 1, Hkv, D, device=dev).half()
    out = torch.empty(batch, L, 1, Hq, D, dtype=torch.float16, device=dev)

    side = [torch.cuda.Stream() for _ in range(min(args.streams, batch))]
    bsteps = []
    if args.batched:
        # descriptor-driven batched launches (quest_decode_layer_step_batch): `--batched` groups of sequences, each group
        # three launches per layer on its own stream
        from quest_b200.utils.batch import BatchDecodeStep
        ng = max(1, min(args.batched, batch))
        groups = [list(range(g, batch, ng)) for g in range(ng)]
        for grp in groups:
            bs = BatchDecodeStep([ics[b] for b in grp])
            bs.update_host()
            bs.upload()
            idx = torch.tensor(grp, device=dev)
            bsteps.append((bs, q[idx].transpose(0, 1).contiguous(), kn[idx].transpose(0, 1).contiguous(),
                           vn[idx].transpose(0, 1).contiguous(), torch.empty(L, len(grp), Hq, D, dtype=torch.float16, device=dev)))
        side = [torch.cuda.Stream() for _ in range(ng)] if ng > 1 else []

    def step():
        # The sequences of a batch are independent: per layer they are forked over `--streams` streams (inside the
        # captured graph these become parallel branches) and joined before the next layer, as a model would (the next
        # layer's q/k/v of all sequences come out of batched GEMMs).  One sequence's top-k latency then hides under
        # the others' metadata / KV streams.
        cur = torch.cuda.current_stream()
        for l in range(L):
            if bsteps:
                for st in side:
                    st.wait_stream(cur)
                for g, (bs, bq, bk, bv, bo) in enumerate(bsteps):
                    with torch.cuda.stream(side[g] if side else cur):
                        bs.layer_step(bq[l, :, 0], bk[l, :, 0], bv[l, :, 0], l, out=bo[l])
                for st in side:
                    cur.wait_stream(st)
                continue
            if not side:
                for b in range(batch):
                    qu.decode_layer_step(q[b, l], kn[b, l], vn[b, l], ics[b], l, out=out[b, l])
                continue
            for st in side:
                st.wait_stream(cur)
            for b in range(batch):
                with torch.cuda.stream(side[b % len(side)]):
                    qu.decode_layer_step(q[b, l], kn[b, l], vn[b, l], ics[b], l, out=out[b, l])
            for st in side:
                cur.wait_stream(st)

    n_pages = ctx // P
    k = budget - 1
    by_meta = (n_pages - 1) * Hkv * D * 2 * 2
    by_kv = Hq * budget * P * 2 * D * 2
    by_small = Hq * (n_pages - 1) * 2 * 2 + Hq * k * 4 * 2 + 2 * Hq * D * 2
    per_layer_rank = batch * (by_meta + by_kv + by_small)
    return step, L, per_layer_rank, {
        "workload": "Llama-3.1-8B GQA shape (32q/8kv x 128, fp16), 128K ctx, page 16, budget 4096, batch 8",
        "parallelism": f"kv-head sharded x{world} ({Hkv} kv / {Hq} q heads per rank), no collective", "scaling": "strong",
        "streams": len(side), "batched_groups": len(bsteps),
        "selection": "per q head (reference semantics, evaluation/quest_attention.py:139-140)"}


def config5(args, rank, world, dev):
    H, pages_per_rank, budget_tok, L = 32, 8192, 8192, args.layers
    budget = budget_tok // P
    tokens = pages_per_rank * P - (1 if rank == world - 1 else 0)  # the decode token completes the last rank's last page
    torch.manual_seed(1234 + 5 + rank)
    ic = qu.InferenceController(L, H, D, P, budget, pages_per_rank * P, torch.float16, dev)
    fill(ic, L, tokens, H, dev)
    if rank == world - 1:
        ic.prepare_metadata(1)
    shard = par.SeqShardedDecode(ic, H, world, rank, budget, exchange=args.exchange)
    n_scored = [pages_per_rank] * (world - 1) + [pages_per_rank - 1]
    shard.begin_step(n_scored)
    torch.manual_seed(77)  # the same q on every rank
    q = torch.randn(L, 1, H, D, device=dev).half()
    kn = torch.randn(L, 1, H, D, device=dev).half()
    vn = torch.randn(L, 1, H, D, device=dev).half()

    def step():
        for l in range(L):
            shard.layer_step(q[l], kn[l], vn[l], l)

    def compute_only():  # the same kernels with the two exchanges replaced by local copies: the collectives' share
        for l in range(L):
            msg = shard.candidates(q[l], kn[l], vn[l], l)
            shard.gath[rank].copy_(msg)
            shard.select(shard.gath)
            rec = shard.attend(q[l], l)
            shard.rec_all[rank].copy_(rec)
            shard.merge(shard.rec_all)

    k = budget - 1
    by_meta = n_scored[rank] * H * D * 2 * 2
    by_kv_total = H * budget * P * 2 * D * 2            # the selected pages, wherever they live
    per_layer_rank = by_meta + by_kv_total // world     # uniform split of the selection (data dependent in reality)
    extra = {"compute_only": compute_only, "shard": shard,
             "exchange_bytes_per_rank": {"candidates": H * min(k, pages_per_rank) * 2, "partials": (H * D + H) * 4}}
    return step, L, per_layer_rank, {
        "workload": f"Yarn-Llama2-7B-128K shape (32 heads x 128, fp16), {world * pages_per_rank * P} ctx "
                    f"({pages_per_rank} pages per GPU), page 16, budget 8192, batch 1, one sequence over all ranks",
        "parallelism": f"sequence (page) sharded x{world}, 2 all-gathers per layer step ("
                       + ("NCCL" if args.exchange == "nccl" else "quest_peer_allgather over symmetric memory") + ")",
        "exchange": args.exchange, "scaling": "weak"}, extra


def measure(config, rank, world, dev, layers=2, streams=8, batched=0, exchange="nccl", steps=30, warmup=5,
            graph_collectives=False):
    """One workload on the ranks of an initialised process group (world > 1; config 5 needs one for world 1 as well).
    Returns the result dict (complete on rank 0)."""
    args = argparse.Namespace(config=config, layers=layers, streams=streams, batched=batched, exchange=exchange, steps=steps,
                              warmup=warmup, graph_collectives=graph_collectives)
    dist_on = dist.is_initialized()
    extra = {}
    if config == 4:
        step, L, by_rank, cfg = config4(args, rank, world, dev)
    else:
        step, L, by_rank, cfg, extra = config5(args, rank, world, dev)
    step()
    torch.cuda.synchronize()

    def agreed_graph(fn, allowed=True):
        """Capture fn on every rank or on none: a rank whose capture failed must not fall out of step with the others."""
        g = try_graph(fn) if allowed else None
        ok = torch.tensor([1 if g is not None else 0], device=dev)
        if dist_on:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        return g if int(ok.item()) else None

    def note(k, v):
        res[k] = v
        if rank == 0:
            sys.stderr.write(f"[bench_configs] {k} = {v:.2f}\n")
            sys.stderr.flush()

    res = {}
    note("eager_us_per_layer", timed(step, args.steps, args.warmup, dist_on) / L * 1e6)
    # collectives inside a captured graph only on request (NCCL capture is all-or-nothing across the ranks)
    g = agreed_graph(step, allowed=(args.config == 4 or world == 1 or args.graph_collectives or args.exchange == "peer"))
    if g is not None:
        note("graph_us_per_layer", timed(g, args.steps, args.warmup, dist_on) / L * 1e6)
    if "compute_only" in extra:
        co = extra.pop("compute_only")
        note("compute_only_eager_us_per_layer", timed(co, args.steps, args.warmup, dist_on) / L * 1e6)
        gc = agreed_graph(co)
        if gc is not None:
            note("compute_only_graph_us_per_layer", timed(gc, args.steps, args.warmup, dist_on) / L * 1e6)
    best = min(v for k, v in res.items() if k in ("eager_us_per_layer", "graph_us_per_layer"))
    by = torch.tensor([float(by_rank)], device=dev)
    if dist_on:
        dist.all_reduce(by)
    total_bytes = float(by.item())
    if args.config == 5:
        shard_obj = extra.pop("shard")
        shard_obj.check_peer()
        co_best = min(v for k, v in res.items() if k.startswith("compute_only"))
        res["exchange_us_per_layer"] = max(best - co_best, 0.0)
    gbps = total_bytes / (best * 1e-6) / 1e9
    return {"config": args.config, "n_gpus": world, "us_per_layer": best, "algorithmic_bytes_per_layer_all_ranks": total_bytes,
            "GBps_all_ranks": gbps, "frac_of_hbm_peak_per_gpu": gbps / world / peak(), "layers_rotated": L, **res, **cfg, **extra}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True, choices=[4, 5])
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--streams", type=int, default=8, help="config 4: concurrent sequences per layer (0 = one stream)")
    ap.add_argument("--batched", type=int, default=0, help="config 4: N groups of sequences, one batched launch chain each")
    ap.add_argument("--exchange", default="nccl", choices=["nccl", "peer"], help="config 5: how the two all-gathers are done")
    ap.add_argument("--graph-collectives", action="store_true", help="config 5, N > 1: also capture the NCCL step in a CUDA graph")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    elif args.config == 5:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    out = measure(args.config, rank, world, dev, args.layers, args.streams, args.batched, args.exchange, args.steps, args.warmup,
                  args.graph_collectives)
    if rank == 0:
        print(json.dumps(out))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
