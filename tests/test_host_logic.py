"""CPU: host-side page bookkeeping mirrors reference quest/utils/kv_cache.py and controller.py."""
import pytest
import torch

from quest_b200.utils import InferenceController, KvCache


def test_kvcache_append_matches_reference_token_loop():
    c = KvCache(num_layers=1, num_heads=2, head_dim=64, max_seq_len=200, page_size=16, dtype=torch.float16,
                device=torch.device("cpu"))
    seqlen, pages = 0, 0
    for n in (1, 15, 1, 16, 33, 0, 7):
        new = c.append_seq(n)
        # reference kv_cache.py:115-126: a page is allocated whenever the last page is full before a token
        exp_new = 0
        for _ in range(n):
            if (seqlen - 1) % 16 + 1 == 16:
                exp_new += 1
            seqlen += 1
        if seqlen and pages == 0 and exp_new == 0:
            exp_new = 0
        pages += new
        assert c.seqlen == seqlen and len(c.indicies) == (seqlen + 15) // 16
        assert c.last_page_len == (seqlen - 1) % 16 + 1
    assert c.indicies == list(range(len(c.indicies)))  # same allocation order as the reference pool
    c.release()
    assert c.seqlen == 0 and c.pool.num_free_blocks == c.pool.capacity
    with pytest.raises(ValueError):
        KvCache(1, 1, 64, 0, 16, torch.float16, torch.device("cpu"))


def test_controller_tables_and_budget():
    dev = torch.device("cpu")
    ic = InferenceController(num_layers=1, num_heads=4, head_dim=64, page_size=16, page_budget=3, max_seq_len=1024,
                             dtype=torch.float16, device=dev)
    ic.prepare_metadata(100)
    ic.begin_forward(100)
    assert ic.kv_indices_with_last.tolist() == list(range(7)) and ic.metadata_indices.tolist() == [0]
    assert ic.kv_indptr_for_append.tolist() == [0, 7] and ic.metadata_indptr_for_append.tolist() == [0, 1]
    assert ic.kv_last_page_idx == 6 and ic.metadata_cache.last_page_len == 7 and ic.kv_cache.last_page_len == 4
    assert not ic.need_estimate()
    ic.end_forward()
    ic.prepare_metadata(1)
    try:
        ic.begin_forward(1)  # builds the decode plan through the C-ABI host entry point
    except RuntimeError:
        pytest.skip("library not built")
    assert ic.inference_page_budget == 3 and ic.need_estimate()
    assert ic.kv_indices_without_last.shape == (4, 6) and ic.kv_indices_without_last.stride(0) == 0
    assert ic.kv_indices_without_last[3].tolist() == list(range(6))
    assert ic.kv_indptr_for_approx_decode.tolist() == [0, 2]
    assert ic.topk_dindices_buffer.shape == (4, 2) and ic.topk_dout_buffer.dtype == torch.float16
    ic.set_page_budget(1 << 20)  # the "skip first two layers" trick, reference llama.py:428-430
    ic.begin_forward(1, updateTensor=False)
    assert ic.inference_page_budget == 7 and not ic.need_estimate()
    ic.clean_states()
    assert ic.kv_cache.seqlen == 0
