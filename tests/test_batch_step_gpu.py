"""quest_decode_layer_step_batch / BatchDecodeStep (device-resident per-sequence records) against the
per-sequence operator path (quest_b200.utils.decode_layer_step, itself pinned to the oracle and the reference kernels in
test_gpu_parity.py / test_vs_ref_kernels.py) on twin caches with the same seeded contents:
  * KV and metadata pools bit-equal after every token, scores bit-equal, selected page sets equal;
  * attention output within 2e-3 (the split-KV chunking differs with the batch, so the fp32 merge order does);
  * ONE captured CUDA graph (record upload + RoPE + the layer steps) replayed token after token, across page and
    metadata-page boundaries, gives the same results as eager per-token calls."""
import numpy as np
import pytest
import torch

import quest_b200.utils as qu
from quest_b200.utils.batch import BatchDecodeStep

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
ATTN_ATOL = 2e-3


def _twin_controllers(lens, L, Hq, Hkv, D, P, budget, grow, seed):
    """Two sets of controllers holding the same random prompts (lens[i] tokens each)."""
    g = torch.Generator().manual_seed(seed)
    cap = max(lens) + grow + P
    a, b = [], []
    for n in lens:
        ia = qu.InferenceController(L, Hq, D, P, budget, cap, torch.float16, DEV, num_kv_heads=Hkv)
        ib = qu.InferenceController(L, Hq, D, P, budget, cap, torch.float16, DEV, num_kv_heads=Hkv)
        for ic in (ia, ib):
            ic.prepare_metadata(n)
            ic.begin_forward(n)
        for l in range(L):
            k = torch.randn(n, Hkv, D, generator=g).half().to(DEV)
            v = torch.randn(n, Hkv, D, generator=g).half().to(DEV)
            qu.append_kv(k, v, ia, l)
            qu.append_kv(k, v, ib, l)
        ia.end_forward(), ib.end_forward()
        a.append(ia), b.append(ib)
    return a, b


def _reference_token(ics, q, k, v, L, rope_scale):
    """Per-sequence operator path for one token; returns o [B, L, Hq, D], scores and selections per (b, l)."""
    B = len(ics)
    outs, sels, scs = [], [], []
    for b, ic in enumerate(ics):
        ic.prepare_metadata(1)
        ic.begin_forward(1)
        row, srow, crow = [], [], []
        for l in range(L):
            qq, kk = q[b, l:l + 1].clone(), k[b, l:l + 1].clone()
            qu.apply_rope_in_place(qq, kk, ic.kv_cache.seqlen - 1, rope_scale=rope_scale)
            o = qu.decode_layer_step(qq, kk, v[b, l:l + 1].contiguous(), ic, l)
            row.append(o[0].clone())
            if ic.need_estimate():
                tab = {int(p): i for i, p in enumerate(ic.kv_cache.indicies)}
                srow.append([sorted(tab[int(x)] for x in r) for r in ic.topk_dindices_buffer.cpu().numpy()])
                n = len(ic.kv_cache.indicies) - 1
                crow.append(ic._scores_buffer(n).clone())
            else:
                n = len(ic.kv_cache.indicies) - 1
                srow.append([list(range(n))] * ic.num_heads)
                crow.append(None)
        ic.end_forward()
        outs.append(torch.stack(row)), sels.append(srow), scs.append(crow)
    return torch.stack(outs), sels, scs


def _check_token(step, ics_a, ics_b, got_o, L, want_o, want_sel, want_sc, topk_per_layer, scores_per_layer):
    torch.cuda.synchronize()
    B = len(ics_a)
    err = (got_o.float() - want_o.float()).abs().max().item()
    assert err <= ATTN_ATOL, err
    for b in range(B):
        tab = {int(p): i for i, p in enumerate(ics_b[b].kv_cache.indicies)}
        n = len(ics_b[b].kv_cache.indicies) - 1
        for l in range(L):
            got = [sorted(tab[int(x)] for x in r if x >= 0) for r in topk_per_layer[l][b].cpu().numpy()]
            assert got == want_sel[b][l], (b, l)
            if want_sc[b][l] is not None:
                sc = scores_per_layer[l][b, :step.Hq * n].view(step.Hq, n)
                assert torch.equal(sc.view(torch.int16), want_sc[b][l].view(torch.int16)), (b, l)
        # pools: the rows / entries in use (everything else is uninitialised memory in both)
        ids_a, ids_b = ics_a[b].kv_cache.indicies, ics_b[b].kv_cache.indicies
        assert ids_a == ids_b
        for j, pid in enumerate(ids_a):
            ntok = ics_a[b].kv_cache.last_page_len if j == len(ids_a) - 1 else step.P
            ka = ics_a[b].kv_cache.pool.buf[:, pid, :, :ntok]
            kb = ics_b[b].kv_cache.pool.buf[:, pid, :, :ntok]
            assert torch.equal(ka.view(torch.int16), kb.view(torch.int16)), (b, j)
        for j in range(len(ids_a)):
            ma = ics_a[b].metadata_cache.pool.buf[:, ics_a[b].metadata_cache.indicies[j // step.P], :, j % step.P]
            mb = ics_b[b].metadata_cache.pool.buf[:, ics_b[b].metadata_cache.indicies[j // step.P], :, j % step.P]
            assert torch.equal(ma.view(torch.int16), mb.view(torch.int16)), (b, j)


@pytest.mark.parametrize("lens,Hq,Hkv,D,P,budget", [
    ((2500, 1029, 4000), 32, 32, 128, 16, 24),
    ((3000, 2047), 32, 8, 128, 16, 40),          # GQA
    ((700, 300, 90, 1023), 8, 8, 64, 8, 30),     # 90 tokens = 12 pages < budget: attends to everything
    ((2040,), 32, 32, 128, 16, 16),              # batch of one
    ((3000, 2047, 5000), 4, 1, 128, 16, 40),     # one kv head per rank (configs[3] sharded over 8 GPUs)
    ((1500, 1501), 16, 4, 128, 16, 20),
])
def test_batch_step_equals_per_sequence_path(lens, Hq, Hkv, D, P, budget):
    L, T, rope_scale = 2, 4, 2.0
    ics_a, ics_b = _twin_controllers(lens, L, Hq, Hkv, D, P, budget, T, seed=len(lens) * 7 + Hkv)
    B = len(lens)
    step = BatchDecodeStep(ics_b)
    g = torch.Generator().manual_seed(5)
    for t in range(T):
        q = torch.randn(B, L, Hq, D, generator=g).half().to(DEV)
        k = torch.randn(B, L, Hkv, D, generator=g).half().to(DEV)
        v = torch.randn(B, L, Hkv, D, generator=g).half().to(DEV)
        want_o, want_sel, want_sc = _reference_token(ics_a, q, k, v, L, rope_scale)
        for ic in ics_b:
            ic.prepare_metadata(1)
        step.update_host()
        step.upload()
        outs, topks, scs = [], [], []
        for l in range(L):
            ql, kl = q[:, l].contiguous(), k[:, l].contiguous()
            step.rope(ql, kl, rope_scale=rope_scale)
            outs.append(step.layer_step(ql, kl, v[:, l].contiguous(), l).clone())
            topks.append(step.topk.clone()), scs.append(step.scores.clone())
        _check_token(step, ics_a, ics_b, torch.stack(outs, dim=1), L, want_o, want_sel, want_sc, topks, scs)


def test_one_graph_replayed_across_tokens_and_page_boundaries():
    """The whole step of a 2-layer, 3-sequence batch captured once; 40 tokens replayed (every sequence opens new pages, one
    crosses a metadata-page boundary at 4096 tokens) against eager per-sequence calls."""
    lens, Hq, Hkv, D, P, budget, L, T, rope_scale = (4090, 1000, 2500), 32, 32, 128, 16, 32, 2, 40, 8.0
    ics_a, ics_b = _twin_controllers(lens, L, Hq, Hkv, D, P, budget, T, seed=3)
    B = len(lens)
    step = BatchDecodeStep(ics_b)
    q = torch.zeros(B, L, Hq, D, dtype=torch.float16, device=DEV)
    k = torch.zeros(B, L, Hkv, D, dtype=torch.float16, device=DEV)
    v = torch.zeros(B, L, Hkv, D, dtype=torch.float16, device=DEV)
    qr, kr = torch.empty_like(q), torch.empty_like(k)
    out = torch.empty(B, L, Hq, D, dtype=torch.float16, device=DEV)
    topks = [torch.empty_like(step.topk) for _ in range(L)]
    scs = [torch.empty_like(step.scores) for _ in range(L)]

    def body():
        step.upload()
        qr.copy_(q), kr.copy_(k)
        for l in range(L):
            ql, kl = qr[:, l].contiguous(), kr[:, l].contiguous()
            step.rope(ql, kl, rope_scale=rope_scale)
            out[:, l].copy_(step.layer_step(ql, kl, v[:, l].contiguous(), l))
            topks[l].copy_(step.topk), scs[l].copy_(step.scores)

    graph = None
    g = torch.Generator().manual_seed(11)
    for t in range(T):
        q.copy_(torch.randn(B, L, Hq, D, generator=g).half())
        k.copy_(torch.randn(B, L, Hkv, D, generator=g).half())
        v.copy_(torch.randn(B, L, Hkv, D, generator=g).half())
        want_o, want_sel, want_sc = _reference_token(ics_a, q, k, v, L, rope_scale)
        for ic in ics_b:
            ic.prepare_metadata(1)
        step.update_host()
        torch.cuda.synchronize()
        if graph is None:   # token 0 runs eagerly inside the capture's warm-up; the graph serves every later token
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                body()
            torch.cuda.current_stream().wait_stream(s)
            _check_token(step, ics_a, ics_b, out, L, want_o, want_sel, want_sc, topks, scs)
            # capture on scratch copies of nothing: the append of token 0 is idempotent, so capturing re-runs it harmlessly
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                body()
            continue
        assert step.fits()
        graph.replay()
        _check_token(step, ics_a, ics_b, out, L, want_o, want_sel, want_sc, topks, scs)
    assert len(ics_b[0].metadata_cache.indicies) == 17  # sequence 0 crossed 4096 tokens = 256 pages = 16 metadata pages


def test_run_ahead_graph_with_device_side_record_updater():
    """40 tokens reserved in advance (run_ahead), then ONE captured graph -- advance() (quest_seq_desc_advance), RoPE, the layer
    steps -- replayed 40 times with no host work on the records: outputs and selected pages of every token match the
    eager per-sequence path, and once the run is over the pools do too."""
    lens, Hq, Hkv, D, P, budget, L, T, rope_scale = (4090, 1000, 2500), 32, 32, 128, 16, 32, 2, 40, 8.0
    ics_a, ics_b = _twin_controllers(lens, L, Hq, Hkv, D, P, budget, T, seed=3)
    B = len(lens)
    step = BatchDecodeStep(ics_b, max_kv_pages=max(lens) // P + T // P + 4)
    q = torch.zeros(B, L, Hq, D, dtype=torch.float16, device=DEV)
    k = torch.zeros(B, L, Hkv, D, dtype=torch.float16, device=DEV)
    v = torch.zeros(B, L, Hkv, D, dtype=torch.float16, device=DEV)
    qr, kr = torch.empty_like(q), torch.empty_like(k)
    out = torch.empty(B, L, Hq, D, dtype=torch.float16, device=DEV)
    topks = [torch.empty_like(step.topk) for _ in range(L)]

    def body():
        step.advance()
        qr.copy_(q), kr.copy_(k)
        for l in range(L):
            ql, kl = qr[:, l].contiguous(), kr[:, l].contiguous()
            step.rope(ql, kl, rope_scale=rope_scale)
            out[:, l].copy_(step.layer_step(ql, kl, v[:, l].contiguous(), l))
            topks[l].copy_(step.topk)

    step.run_ahead(T)
    torch.cuda.synchronize()
    graph = None
    g = torch.Generator().manual_seed(11)
    for t in range(T):
        q.copy_(torch.randn(B, L, Hq, D, generator=g).half())
        k.copy_(torch.randn(B, L, Hkv, D, generator=g).half())
        v.copy_(torch.randn(B, L, Hkv, D, generator=g).half())
        want_o, want_sel, want_sc = _reference_token(ics_a, q, k, v, L, rope_scale)
        if graph is None:  # capture only (a captured launch does not run); every token, the first included, is a replay
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=s):
                body()
            torch.cuda.current_stream().wait_stream(s)
        graph.replay()
        torch.cuda.synchronize()
        err = (out.float() - want_o.view(B, L, Hq, D).float()).abs().max().item()
        assert err <= ATTN_ATOL, (t, err)   # the batched plan splits the pages differently from the per-sequence one
        for b in range(B):   # selected pages, in list positions: the device tables hold the whole run, so map through them
            tab = {int(p): i for i, p in enumerate(ics_b[b].kv_cache.indicies)}
            for l in range(L):
                got = [sorted(tab[int(x)] for x in r if x >= 0) for r in topks[l][b].cpu().numpy()]
                assert got == want_sel[b][l], (t, b, l)
    # the run is over: host and device agree again; the pools hold what the eager path wrote
    for b in range(B):
        ia, ib = ics_a[b], ics_b[b]
        assert ia.kv_cache.seqlen == ib.kv_cache.seqlen and len(ia.kv_cache.indicies) == len(ib.kv_cache.indicies)
        for j, (pa, pb) in enumerate(zip(ia.kv_cache.indicies, ib.kv_cache.indicies)):
            n = P if j < len(ia.kv_cache.indicies) - 1 else ia.kv_cache.last_page_len
            assert torch.equal(ia.kv_cache.pool.buf[:, pa, :, :n].view(torch.int16), ib.kv_cache.pool.buf[:, pb, :, :n].view(torch.int16)), (b, j)
