"""BASELINE.json's long-context shapes at FULL size (configs[3]: GQA 32q/8kv, 128K tokens, budget 4096; configs[4] per-GPU
share: MHA 32 heads, 128K tokens, budget 8192), checked through properties that do not need the oracle to walk the whole
cache:
  * metadata entries == per-page amin / amax of the pool's own K rows (exact);
  * scores bit-equal to the oracle's estimate (reference lane / butterfly order) on the whole row;
  * the selected set satisfies the top-k contract on those scores (everything above the k-th value, ties lowest position
    first) -- rows of 8191 pages take the cluster kernel;
  * the output equals the oracle's attention over exactly the selected pages + the last page (per head, on a compacted
    pool), within 2e-3."""
import numpy as np
import pytest
import torch

import quest_b200.utils as qu
from oracle import quest_oracle as orc

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
ATTN_ATOL = 2e-3


@pytest.mark.parametrize("Hq,Hkv,ctx,budget_tokens", [(32, 8, 131072, 4096), (32, 32, 131072, 8192)])
def test_long_context_layer_step_properties(Hq, Hkv, ctx, budget_tokens):
    D, P = 128, 16
    budget = budget_tokens // P
    g = torch.Generator(device=DEV).manual_seed(ctx + Hkv)
    ic = qu.InferenceController(1, Hq, D, P, budget, ctx, torch.float16, DEV, num_kv_heads=Hkv)
    done = 0
    while done < ctx - 1:  # prompt through append_kv in chunks
        n = min(32768, ctx - 1 - done)
        ic.prepare_metadata(n), ic.begin_forward(n)
        k = torch.randn(n, Hkv, D, generator=g, device=DEV, dtype=torch.float32).half()
        v = torch.randn(n, Hkv, D, generator=g, device=DEV, dtype=torch.float32).half()
        qu.append_kv(k, v, ic, 0)
        ic.end_forward()
        done += n
    q = torch.randn(1, Hq, D, generator=g, device=DEV, dtype=torch.float32).half()
    k1 = torch.randn(1, Hkv, D, generator=g, device=DEV, dtype=torch.float32).half()
    v1 = torch.randn(1, Hkv, D, generator=g, device=DEV, dtype=torch.float32).half()
    ic.prepare_metadata(1), ic.begin_forward(1)
    assert ic.need_estimate()
    o = qu.decode_layer_step(q, k1, v1, ic, 0)
    ic.end_forward()
    torch.cuda.synchronize()
    ids = ic.kv_cache.indicies
    n_pages, n = len(ids), len(ids) - 1
    assert n_pages == ctx // P
    table = torch.tensor(ids, device=DEV)
    kv, meta = ic.kv_cache.buf_layer(0), ic.metadata_cache.buf_layer(0)

    # metadata == amax / amin of the pages' own K rows (all pages are full: the decode token completed the last one)
    kpages = kv[table, 0]                                   # [n_pages, P, Hkv, D] in sequence order
    mtab = torch.tensor(ic.metadata_cache.indicies, device=DEV)
    m = meta[mtab]                                          # [n_meta, 2, P, Hkv, D]
    mx = m[:, 0].reshape(-1, Hkv, D)[:n_pages]
    mn = m[:, 1].reshape(-1, Hkv, D)[:n_pages]
    assert torch.equal(mx, kpages.amax(dim=1)) and torch.equal(mn, kpages.amin(dim=1))

    # scores: bit-equal to the oracle on the whole row
    scores = ic._scores_buffer(n)
    want_s = orc.estimate_scores(q[0].cpu().numpy(), meta.cpu().numpy(), ic.metadata_cache.indicies, n, num_qo_heads=Hq)
    assert np.array_equal(scores.cpu().numpy().view(np.int16), want_s.view(np.int16))

    # selection: the top-k contract on those scores
    k_sel = budget - 1
    lut = {int(p): i for i, p in enumerate(ids)}
    sel = ic.topk_dindices_buffer.cpu().numpy()
    pos = np.array([[lut[int(x)] for x in row] for row in sel], dtype=np.int64)
    assert (np.diff(pos, axis=1) > 0).all()                 # ascending positions, no duplicates
    assert orc.topk_contract_ok(want_s, pos, k_sel)
    np.testing.assert_array_equal(pos, orc.topk_select(want_s, k_sel))

    # output: the oracle over exactly the selected pages + the last page, head by head on a compacted pool
    G = Hq // Hkv
    got = o[0].float().cpu().numpy()
    qn = q[0].cpu().numpy()
    for h in (0, 5, Hq - 1):
        pages = torch.tensor(np.concatenate([sel[h], [ids[-1]]]), device=DEV)
        small = kv[pages][:, :, :, h // G:h // G + 1].cpu().numpy()   # [k+1, 2, P, 1, D]
        oh, _ = orc.sparse_decode_attention(qn[h:h + 1], small, np.arange(k_sel).reshape(1, -1), k_sel, ic.kv_cache.last_page_len)
        assert np.abs(got[h] - oh[0]).max() <= ATTN_ATOL
