"""Sequence (page) sharded decode step on the GPU (quest_b200.parallel.SeqShardedDecode) against the
single-GPU layer step on the same seeded inputs:
  * the global page selection is the single-GPU one exactly (every shard's share united), ties included;
  * the merged output is within the fp16 attention tolerance of the unsharded output (2e-3 max abs);
  * the shards' KV / metadata pools hold exactly the rows the single-GPU pools hold.
One process drives all shards on cuda:0 and stands in for the two all-gathers (torch.stack); the same phases under
NCCL with one process per GPU are covered by test_two_rank_nccl (needs 2 GPUs; `gpurun --gpus 2`)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import quest_b200.utils as qu
from quest_b200 import parallel as par
from tests import seqshard_common as sc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATTN_ATOL = 2e-3  # north_star: fp16 attention output max abs error


def _run_case(seed, kv_len, Hq, Hkv, D, P, budget, world, steps=3, ties=False, exchange="emulated", fused=True):
    dev = torch.device(DEV)
    k, v, dec = sc.seeded_case(seed, kv_len, Hq, Hkv, D, steps)
    if ties:  # repeat a block of keys: equal metadata -> equal scores across shards, the tie rule decides
        blk = k[:4 * P].clone()
        for s in range(0, k.shape[0] - 4 * P, 4 * P):
            k[s:s + 4 * P] = blk
    full = sc.filled_controller(k, v, Hq, Hkv, D, P, budget, kv_len + steps + P, dev)
    ranges = sc.shard_token_ranges(kv_len, P, world)
    ics = [sc.shard_controller(k, v, ranges[r], r, world, Hq, Hkv, D, P, budget, steps, dev) for r in range(world)]
    if exchange == "peer":
        # all ranks in this process: their exchange areas are plain device buffers, every rank's kernels run on its own
        # stream so that the ranks' waits can be met by one another (the kernels are single-CTA and co-resident)
        probe = par.SeqShardedDecode(ics[0], Hq, 1, 0, budget)
        nbytes = par.SeqShardedDecode._peer_layout(probe.lib, Hq, D, budget - 1, world)[2]
        areas = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(world)]
        shards = [par.SeqShardedDecode(ics[r], Hq, world, r, budget, exchange="peer", peer_buffers=areas, fused_candidates=fused) for r in range(world)]
        streams = [torch.cuda.Stream() for _ in range(world)]
    else:
        shards = [par.SeqShardedDecode(ics[r], Hq, world, r, budget, fused_candidates=fused) for r in range(world)]
    for q, k1, v1 in dec:
        q, k1, v1 = q.to(dev), k1.to(dev), v1.to(dev)
        # single-GPU truth
        full.prepare_metadata(1)
        full.begin_forward(1)
        assert full.need_estimate()
        want = qu.decode_layer_step(q, k1, v1, full, 0).clone()
        full.end_forward()
        table = full.kv_cache.indicies
        lut = {int(p): i for i, p in enumerate(table)}
        want_sel = [sorted(lut[int(x)] for x in row) for row in full.topk_dindices_buffer.cpu().numpy()]
        # sharded
        ics[-1].prepare_metadata(1)
        n_scored = [len(ic.kv_cache.indicies) - (1 if r == world - 1 else 0) for r, ic in enumerate(ics)]
        for s in shards:
            s.begin_step(n_scored)
        if exchange == "peer":
            torch.cuda.synchronize()
            outs = [None] * world
            for r, s in enumerate(shards):  # enqueue every rank's whole step before waiting for any of them
                with torch.cuda.stream(streams[r]):
                    outs[r] = s.layer_step(q, k1, v1, 0)
            torch.cuda.synchronize()
            for s in shards:
                s.check_peer()
            outs = [o.clone() for o in outs]
            sels = [s.sel[:, :s.k].clone() for s in shards]
        else:
            gath = torch.stack([s.candidates(q, k1, v1, 0) for s in shards])
            sels = [s.select(gath).clone() for s in shards]
            recs = torch.stack([s.attend(q, 0) for s in shards])
            outs = [s.merge(recs).clone() for s in shards]
            torch.cuda.synchronize()
        first = np.concatenate([[0], np.cumsum(n_scored)[:-1]])
        got_sel = [[] for _ in range(Hq)]
        for r in range(world):
            tab = ics[r].kv_cache.indicies[:n_scored[r]]
            for h, row in enumerate(sc.selected_positions(sels[r], tab, int(first[r]))):
                got_sel[h] += row
        assert [sorted(x) for x in got_sel] == want_sel
        for o in outs:
            assert torch.equal(o, outs[0])  # every rank ends with the same bits
        err = (outs[0].float() - want.float()).abs().max().item()
        assert err <= ATTN_ATOL, err
    # the shards' pools hold the single-GPU rows: compare page by page through the tables
    fkv, fmeta = full.kv_cache.buf_layer(0), full.metadata_cache.buf_layer(0)
    page = 0
    for r in range(world):
        kv = ics[r].kv_cache.buf_layer(0)
        meta = ics[r].metadata_cache.buf_layer(0)
        ids = ics[r].kv_cache.indicies
        n_tok_last = ics[r].kv_cache.last_page_len
        for j, pid in enumerate(ids):
            fp = full.kv_cache.indicies[page]
            ntok = n_tok_last if j == len(ids) - 1 else P
            assert torch.equal(kv[pid, :, :ntok].view(torch.int16), fkv[fp, :, :ntok].view(torch.int16))
            m_l = meta[ics[r].metadata_cache.indicies[j // P], :, j % P]
            m_f = fmeta[full.metadata_cache.indicies[page // P], :, page % P]
            assert torch.equal(m_l.view(torch.int16), m_f.view(torch.int16))
            page += 1
    assert page == len(full.kv_cache.indicies)


@pytest.mark.parametrize("kv_len,Hq,Hkv,D,P,budget,world", [
    (3000, 32, 32, 128, 16, 24, 2),
    (2049, 32, 32, 128, 16, 16, 2),     # the decode token opens a new page on the last rank
    (5000, 32, 8, 128, 16, 40, 4),      # GQA
    (1000, 8, 8, 64, 8, 12, 3),
    (700, 32, 32, 128, 16, 30, 4),      # k = 29 > pages per rank (10-11): every local page is a candidate
    (32768, 32, 32, 128, 16, 128, 8),   # BASELINE configs[1] shape cut in eight
])
def test_sharded_step_equals_single_gpu(kv_len, Hq, Hkv, D, P, budget, world):
    _run_case(kv_len + world, kv_len, Hq, Hkv, D, P, budget, world)


@pytest.mark.parametrize("kv_len,Hq,Hkv,D,P,budget,world", [
    (3000, 32, 32, 128, 16, 24, 2),
    (5000, 32, 8, 128, 16, 40, 4),
    (700, 32, 32, 128, 16, 30, 4),      # ragged candidate counts: -inf padding travels
    (32768, 32, 32, 128, 16, 128, 8),
    (2100, 32, 32, 128, 16, 24, 1),     # world 1: the exchange degenerates to a local copy
])
def test_sharded_step_peer_memory_exchange(kv_len, Hq, Hkv, D, P, budget, world):
    """The same step with both exchanges done by quest_peer_allgather (push into every rank's staging area, flags,
    wait, unpack) instead of being stood in for: all ranks' kernels run concurrently on one GPU."""
    _run_case(kv_len + world, kv_len, Hq, Hkv, D, P, budget, world, steps=4, exchange="peer")


def test_peer_allgather_kernel_layout_and_epochs():
    """quest_peer_allgather alone: dst = [rows][world][row_bytes], repeated calls (epoch parity), a missing rank (timeout instead of a hang), argument checks."""
    import ctypes
    import quest_b200._kernels as K
    lib = K.load_library()
    R, rows, row_bytes = 3, 5, 48
    nst = lib.quest_peer_staging_bytes(rows, row_bytes, R)
    stg = [torch.zeros(nst, dtype=torch.uint8, device=DEV) for _ in range(R)]
    flg = [torch.zeros(R, dtype=torch.int32, device=DEV) for _ in range(R)]
    state = [torch.zeros(1 + R, dtype=torch.int32, device=DEV) for _ in range(R)]
    src = [torch.zeros(rows, row_bytes, dtype=torch.uint8, device=DEV) for _ in range(R)]
    dst = [torch.zeros(rows, R, row_bytes, dtype=torch.uint8, device=DEV) for _ in range(R)]
    arr = ctypes.c_void_p * R
    ps, pf = arr(*[x.data_ptr() for x in stg]), arr(*[x.data_ptr() for x in flg])
    streams = [torch.cuda.Stream() for _ in range(R)]

    def round_():
        for r in range(R):
            with torch.cuda.stream(streams[r]):
                assert lib.quest_peer_allgather(src[r].data_ptr(), rows, row_bytes, ps, pf, state[r].data_ptr(), dst[r].data_ptr(),
                                                R, r, 500, torch.cuda.current_stream().cuda_stream) == 0

    g = torch.Generator().manual_seed(0)
    for it in range(5):
        msgs = [torch.randint(0, 256, (rows, row_bytes), generator=g, dtype=torch.uint8) for _ in range(R)]
        for r in range(R):
            src[r].copy_(msgs[r].to(DEV))
        torch.cuda.synchronize()
        round_()
        torch.cuda.synchronize()
        want = torch.stack(msgs, dim=1)  # [rows, R, row_bytes]
        for r in range(R):
            assert torch.equal(dst[r].cpu(), want), (it, r)
            assert state[r].tolist() == [0] + [it + 1] * R
    # a rank missing from the exchange: the others time out and say so instead of hanging
    with torch.cuda.stream(streams[0]):
        assert lib.quest_peer_allgather(src[0].data_ptr(), rows, row_bytes, ps, pf, state[0].data_ptr(), dst[0].data_ptr(),
                                        R, 0, 50, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert state[0].tolist()[0] == 2
    # ... loudly: the lost ranks' slots are poisoned (all-ones = NaN as fp16 / fp32), the sticky device error is raised and the
    # next exchange is refused until it is cleared
    d0 = dst[0].cpu()
    assert bool((d0[:, 1:] == 255).all()) and torch.equal(d0[:, 0], src[0].cpu())
    assert lib.quest_device_error() & K.DEVERR_PEER
    assert lib.quest_peer_allgather(src[0].data_ptr(), rows, row_bytes, ps, pf, state[0].data_ptr(), dst[0].data_ptr(), R, 0, 50, 0) == -4
    lib.quest_device_error_clear()
    assert lib.quest_device_error() == 0
    # argument checks
    assert lib.quest_peer_allgather(src[0].data_ptr(), rows, 40, ps, pf, state[0].data_ptr(), dst[0].data_ptr(), R, 0, 50, 0) == -1
    assert lib.quest_peer_allgather(src[0].data_ptr(), rows, row_bytes, ps, pf, state[0].data_ptr(), dst[0].data_ptr(), 9, 0, 50, 0) == -2


def test_sharded_step_ties_across_shards():
    _run_case(5, 4000, 32, 32, 128, 16, 20, 4, ties=True)


def test_decode_run_negative_ids_and_no_last_page():
    """quest_decode_run: ids < 0 contribute nothing, last_page_len = 0 drops the last page, a head with no rows at all
    yields o = 0 and lse = -inf."""
    import ctypes
    import quest_b200._kernels as K
    lib = K.load_library()
    H, D, P, npg, ksel = 32, 128, 16, 40, 12
    g = torch.Generator().manual_seed(3)
    kv = torch.randn(npg, 2, P, H, D, generator=g).half().to(DEV)
    q = torch.randn(1, H, D, generator=g).half().to(DEV)
    ids = torch.stack([torch.randperm(npg - 1, generator=g)[:ksel] for _ in range(H)]).int()
    masked = ids.clone()
    masked[:, ::3] = -1
    masked[5] = -1  # head 5 owns nothing
    ppc, nch = ctypes.c_uint32(), ctypes.c_uint32()
    assert lib.quest_decode_plan(ksel, H, H, D, P, 148, ctypes.byref(ppc), ctypes.byref(nch)) == 0
    ws = torch.zeros(lib.quest_decode_workspace_bytes(H, D, nch.value), dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream

    def run(idx, n, last_len):
        o = torch.empty(H, D, dtype=torch.float16, device=DEV)
        lse = torch.empty(H, dtype=torch.float32, device=DEV)
        di = idx.to(DEV).contiguous()
        assert lib.quest_decode_run(q.data_ptr(), o.data_ptr(), lse.data_ptr(), kv.data_ptr(), di.data_ptr(), idx.shape[1], n,
                                    npg - 1, last_len, H, H, D, P, ppc.value, ws.data_ptr(), ws.numel(), st) == 0
        torch.cuda.synchronize()
        return o, lse

    o_m, l_m = run(masked, ksel, 0)
    assert torch.all(o_m[5] == 0) and torch.isinf(l_m[5]) and l_m[5] < 0
    # reference: per head, dense softmax over the kept pages' rows
    qf = q[0].float() * (1.0 / np.sqrt(D))
    for h in (0, 7, 31):
        keep = [int(x) for x in masked[h] if x >= 0]
        kk = kv[keep, 0, :, h].float().reshape(-1, D)
        vv = kv[keep, 1, :, h].float().reshape(-1, D)
        s = kk @ qf[h]
        want = torch.softmax(s, 0) @ vv
        assert (o_m[h].float() - want).abs().max().item() <= ATTN_ATOL
        want_lse = torch.logsumexp(s, 0).item() / np.log(2.0)
        assert abs(l_m[h].item() - want_lse) <= 1e-3


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_nccl():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(root, "tests", "seqshard_nccl_worker.py")]
    r = subprocess.run(cmd, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "seqshard nccl ok" in r.stdout


@pytest.mark.parametrize("exchange", ["emulated", "peer"])
def test_separate_operator_candidates_path(exchange):
    """fused_candidates=False: phase 1 as append + estimate + top-k + copies (the path shapes outside the bin-total top-k take);
    the default (one quest_estimate_select call) is what every other test here runs."""
    _run_case(5, 3000, 32, 32, 128, 16, 24, 3, exchange=exchange, fused=False)
    import quest_b200._kernels as K
    with K.debug_option(K.DBG_NO_ROW_COMPACT, 1):  # ... and the attend phase walking the sparse list chunk by chunk
        _run_case(6, 2049, 32, 8, 128, 16, 16, 2, ties=True, exchange=exchange, fused=False)
