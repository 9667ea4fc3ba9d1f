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


def test_argument_checks_of_the_batched_and_peer_entry_points_need_no_gpu():
    """quest_decode_layer_step_batch / quest_peer_allgather / quest_merge_states reject bad arguments on the host,
    before any CUDA call (include/quest_b200.h: negative return codes)."""
    import ctypes

    import quest_b200._kernels as K
    from quest_b200.parallel import SeqShardedDecode
    lib = K.load_library()
    one = ctypes.c_void_p(16)  # any non-null, 16-byte aligned address: never dereferenced on these paths
    # batched layer step: null records, k = 0, workspace too small
    args = [one, one, one, one, None, 1, 0, 0, 2048, one, 32 * 2047, one, 127, 32, 32, 128, 16, 4, one, 1 << 20, None]
    assert lib.quest_decode_layer_step_batch(*args) == -1
    args[4] = one
    args[12] = 0
    assert lib.quest_decode_layer_step_batch(*args) == -1
    args[12] = 127
    args[19] = 16
    assert lib.quest_decode_layer_step_batch(*args) == -3
    # peer all-gather: row size not a multiple of 16 bytes, world > 8, rank >= world
    arr = (ctypes.c_void_p * 8)(*[16] * 8)
    assert lib.quest_peer_allgather(one, 4, 40, arr, arr, one, one, 2, 0, 0, None) == -1
    assert lib.quest_peer_allgather(one, 4, 48, arr, arr, one, one, 9, 0, 0, None) == -2
    assert lib.quest_peer_allgather(one, 4, 48, arr, arr, one, one, 2, 2, 0, None) == -2
    assert lib.quest_peer_staging_bytes(4, 48, 3) == 2 * 3 * 4 * 48
    # merge of packed records: a stride smaller than one record's o block
    assert lib.quest_merge_states(one, one, 2, 100, one, None, 32, 128, None) == -1
    # exchange-area layout of the sequence-sharded step: staging for candidates and records, then the flag words
    off_rec, off_flags, total = SeqShardedDecode._peer_layout(lib, 32, 128, 511, 8)
    assert off_rec == 2 * 8 * 32 * 512 * 2 and off_flags >= off_rec + 2 * 8 * (32 * 128 + 32) * 4
    assert off_flags % 16 == 0 and total >= off_flags + 2 * 8 * 4


def test_argument_checks_of_the_round2_entry_points_need_no_gpu():
    """quest_rope_append_prefill / quest_prefill_layer_step (row f-3) and the fp8 metadata entry points (row f-4) reject bad
    arguments on the host, before any CUDA call."""
    import ctypes

    import quest_b200._kernels as K
    lib = K.load_library()
    one = ctypes.c_void_p(16)
    prod = [one, one, one, 2, one, one, 1, 2, one, one, 1]  # q k v append_len kv idx pages last_len meta midx meta_pages
    assert lib.quest_rope_append_prefill(*prod, 1, 1, 96, 16, 1.0, 1e4, 0, None) == -2   # head_dim not built
    assert lib.quest_rope_append_prefill(*prod, 1, 1, 128, 16, 0.0, 1e4, 0, None) == -1  # rope_scale must be positive
    assert lib.quest_rope_append_prefill(*prod, 1, 1, 128, 16, 1.0, -1.0, 0, None) == -1
    assert lib.quest_rope_append_prefill(None, *prod[1:], 1, 1, 128, 16, 1.0, 1e4, 0, None) == -1
    bad = list(prod)
    bad[3] = 40  # more new tokens than the cache holds
    assert lib.quest_rope_append_prefill(*bad, 1, 1, 128, 16, 1.0, 1e4, 0, None) == -1
    bad = list(prod)
    bad[7] = 17  # last page longer than a page
    assert lib.quest_rope_append_prefill(*bad, 1, 1, 128, 16, 1.0, 1e4, 0, None) == -1
    none_new = list(prod)
    none_new[3] = 0  # nothing to append: a no-op
    assert lib.quest_rope_append_prefill(*none_new, 1, 1, 128, 16, 1.0, 1e4, 0, None) == 0
    layer = [one, one, one, None, 2, one, one, 1, 2, one, one, 1]  # o missing
    assert lib.quest_prefill_layer_step(*layer, 1, 1, 128, 16, 1.0, 1e4, 0, None) == -1
    # fp8 metadata
    assert lib.quest_metadata_to_fp8(one, one, 3, 2, one, 1, 128, 16, None) == -1
    assert lib.quest_metadata_to_fp8(one, one, 0, 1, one, 1, 96, 16, None) == -2
    assert lib.quest_metadata_to_fp8(one, one, 5, 5, one, 1, 128, 16, None) == 0  # empty range
    assert lib.quest_estimate_fp8(one, one, 4, None, one, 1, 1, 128, 16, 0, None) == -1
    assert lib.quest_estimate_fp8(one, one, 4, one, one, 3, 2, 128, 16, 0, None) == -1
    assert lib.quest_estimate_fp8(one, one, 0, one, one, 1, 1, 128, 16, 0, None) == 0
    step = [one, one, one, one, one, one, 4, 0, 1, one, None, one, 1, 4, one, one, 2, 1, 1, 128, 16, 4, one, 1 << 20, None]
    assert lib.quest_decode_layer_step_fp8(*step) == -1  # no e4m3 pool
    # device-side record updater
    assert lib.quest_seq_desc_advance(None, 1, 16, None) == -1
    assert lib.quest_seq_desc_advance(one, 0, 16, None) == -1
    assert lib.quest_seq_desc_advance(one, 1, 0, None) == -1
