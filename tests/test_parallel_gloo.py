"""world_size-2 tests of the sharding glue (quest_b200/parallel.py) on CPU with the gloo backend: the oracle plays the
part of the per-rank kernels, the collectives and the selection / merge logic are the code under test.
  * kv-head sharding: the ranks' outputs concatenated equal the single-process step bit for bit (no collective on
    the data path; the gather here is only the test's comparison);
  * sequence sharding: candidate all-gather -> the same global top-k as the single-process selection (including ties),
    partial states all-gather + LSE merge -> the single-process output."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import quest_oracle as orc
from quest_b200 import parallel as par
from tests.helpers import OracleCache, golden_inputs

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _case(seed, kv_len, H, D, P, tie_scores=False):
    q, k, v = golden_inputs(seed, kv_len, H, D)
    if tie_scores:  # many equal keys -> many equal scores: the tie rule decides
        half = kv_len // 2
        k[:half] = np.tile(k[:P], (half // P + 1, 1, 1))[:half]
    oc = OracleCache((kv_len + P - 1) // P + 2, P, H, D, seed=seed)
    oc.append(k[:-1], v[:-1])
    oc.append(k[-1:], v[-1:])
    return q[0], oc


def _partial(q, oc, phys_pages_per_head, with_last):
    """(o fp32 [H, D], lse base-2 [H]) over each head's own page list (+ the last page on the rank that holds it)."""
    H, D = q.shape
    o = np.zeros((H, D), np.float32)
    lse = np.full(H, -np.inf, np.float32)
    for h in range(H):
        pages = np.asarray(phys_pages_per_head[h], np.int64)
        if len(pages) == 0 and not with_last:
            continue
        pool_h = oc.kv_pool[:, :, :, h:h + 1]
        if with_last:
            oh, lh = orc.sparse_decode_attention(q[h:h + 1], pool_h, pages.reshape(1, -1), oc.kv_indices[-1], oc.last_page_len)
        else:  # the oracle always adds a "last page": hand it the rank's final page as such
            oh, lh = orc.sparse_decode_attention(q[h:h + 1], pool_h, pages[:-1].reshape(1, -1), int(pages[-1]), oc.P)
        o[h], lse[h] = oh[0], lh[0]
    return o, lse


def _worker(rank, port, seed, kv_len, budget, tie_scores, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        H, D, P = 4, 128, 16
        q, oc = _case(seed, kv_len, H, D, P, tie_scores)
        want_o, want_s, want_sel = orc.quest_decode_step(q, oc.kv_pool, oc.kv_indices, oc.last_page_len, oc.meta_pool,
                                                         oc.meta_indices, budget)
        n, k = len(oc.kv_indices) - 1, budget - 1
        table = np.asarray(oc.kv_indices, np.int64)

        # ---- kv-head sharding: the single-GPU path on a slice of the heads
        hs, he = par.head_shard(H, WORLD, rank)
        o_h, s_h, sel_h = orc.quest_decode_step(q[hs:he], oc.kv_pool[:, :, :, hs:he], oc.kv_indices, oc.last_page_len,
                                                oc.meta_pool[:, :, :, hs:he], oc.meta_indices, budget)
        parts = [torch.empty(he - hs, D) for _ in range(WORLD)]
        dist.all_gather(parts, torch.from_numpy(np.ascontiguousarray(o_h)))
        np.testing.assert_array_equal(torch.cat(parts).numpy(), want_o)
        np.testing.assert_array_equal(np.sort(sel_h, axis=1), np.sort(want_sel[hs:he], axis=1))

        # ---- sequence sharding
        a, b = par.page_shard(n, WORLD, rank)
        local = torch.from_numpy(np.ascontiguousarray(want_s[:, a:b]))           # this rank's fp16 scores
        sel = par.select_global_topk(local, a, k, pad_to=k).numpy()              # [H, k] ascending positions
        np.testing.assert_array_equal(sel, np.sort(want_sel, axis=1))
        mine = [table[row[(row >= a) & (row < b)]] for row in sel]                # my share, physical page ids
        o_p, lse_p = _partial(q, oc, mine, with_last=(rank == WORLD - 1))
        o = par.combine_partials(torch.from_numpy(o_p), torch.from_numpy(lse_p)).numpy()
        assert np.abs(o - want_o).max() <= 1e-5
        out[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kv_len,budget,tie_scores", [(400, 8, False), (1030, 20, False), (600, 12, True)])
def test_two_rank_sharding_matches_single_process(kv_len, budget, tie_scores):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, port, 77 + kv_len, kv_len, budget, tie_scores, out)) for r in range(WORLD)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(180)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        assert dict(out) == {0: "ok", 1: "ok"}


def test_shard_arithmetic():
    assert [par.head_shard(8, 4, r) for r in range(4)] == [(0, 2), (2, 4), (4, 6), (6, 8)]
    with pytest.raises(ValueError):
        par.head_shard(6, 4, 0)
    cuts = [par.page_shard(2047, 8, r) for r in range(8)]
    assert cuts[0][0] == 0 and cuts[-1][1] == 2047 and all(cuts[i][1] == cuts[i + 1][0] for i in range(7))
    assert max(b - a for a, b in cuts) - min(b - a for a, b in cuts) <= 1


def test_order_key_matches_selection_rule():
    s = torch.tensor([[1.0, -2.0, 1.0, 0.0, -0.0, 3.5]], dtype=torch.float16)
    pos = torch.arange(6).unsqueeze(0)
    order = torch.argsort(par._order_key(s, pos), dim=1, descending=True)[0].tolist()
    assert order[:3] == [5, 0, 2]          # 3.5, then the two 1.0 by position
    assert order[-1] == 1                  # -2.0 last
