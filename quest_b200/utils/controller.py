"""InferenceController: mirror of reference quest/utils/controller.py:7-146.

Same constructor, methods and attribute names (QuestAttention.forward and the reference tests read them
directly), but the per-step index tensors are not rebuilt from Python lists (controller.py:84-120):
the page tables live in two device tensors sized for max_seq_len that grow by one small H2D copy when a
page is allocated, ``kv_indices_without_last`` is a stride-0 broadcast view instead of a [H, N-1] repeat,
the [0, n] indptr tensors are cached per n, and the top-k output buffers are cached per budget.
"""
import torch

from quest_b200.utils.decode_wrapper import BatchDecodeWithPagedKVCacheWrapper
from quest_b200.utils.kv_cache import KvCache
from quest_b200.utils.utils import TensorLayout


class InferenceController:
    def __init__(self, num_layers, num_heads, head_dim, page_size,
                 page_budget,  # Real page budget including the last page
                 max_seq_len,  # Real max for allocating kv / metadata
                 dtype, device, num_kv_heads=None, fp8_metadata=False):
        # num_kv_heads (extension): GQA pools hold num_kv_heads heads while estimate/top-k/decode run per q head.
        # fp8_metadata (extension, off by default): keep an e4m3 copy of the page metadata, rounded
        # outward, and score from it -- the estimate streams half the bytes; scores are upper bounds of the reference's, not
        # bit-identical to them.  The reference has fp16 metadata only.
        max_kv_pages_num = (max_seq_len + page_size - 1) // page_size
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.kv_cache = KvCache(num_layers=num_layers, num_heads=self.num_kv_heads, head_dim=head_dim,
                                max_seq_len=max_seq_len, page_size=page_size, dtype=dtype, device=device)
        self.metadata_cache = KvCache(num_layers=num_layers, num_heads=self.num_kv_heads, head_dim=head_dim,
                                      max_seq_len=max_kv_pages_num, page_size=page_size, dtype=dtype, device=device)
        self.fp8_metadata = bool(fp8_metadata)
        self.meta8 = torch.zeros((num_layers,) + tuple(self.metadata_cache.buf_layer(0).shape), dtype=torch.uint8,
                                 device=device) if self.fp8_metadata else None  # [layers, meta pages, 2, page, kv heads, dim]
        self.layout = TensorLayout.NHD
        self.device = device
        self.dtype = dtype
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.page_size = page_size

        self._page_budget = page_budget
        self._decode_handler = BatchDecodeWithPagedKVCacheWrapper(kv_layout="NHD")

        # device-resident page tables (capacity entries) + how many entries are valid on the device
        self._kv_table = torch.zeros(self.kv_cache.pool.capacity, dtype=torch.int32, device=device)
        self._meta_table = torch.zeros(self.metadata_cache.pool.capacity, dtype=torch.int32, device=device)
        self._kv_synced = (0, 0)    # (valid entries on the device, generation of the host list they mirror)
        self._meta_synced = (0, 0)
        self._indptr_cache = {}
        self._topk_cache = {}
        self._step = None  # bound arguments of decode_layer_step, valid until the next begin_forward
        self._ops = None   # bound arguments of the four decode operators (quest_b200.utils._BoundOps), same lifetime
        self._pool_ptrs = None
        self._step_meta = None

        self.kv_indices_with_last = None
        self.kv_indices_without_last = None
        self.metadata_indices = None
        self.kv_last_page_idx = None  # For decoding self-attention
        self.metadata_last_page_idx = None

        self.kv_indptr_for_append = None
        self.metadata_indptr_for_append = None
        self.kv_indptr_for_approx_decode = None

        self.inference_page_budget = None

        self.topk_dout_buffer = None
        self.topk_dindices_buffer = None
        self.topk_buf = None

    # Used for controlling the number of pages; layers 0,1 are skipped by manipulating this.
    def set_page_budget(self, page_budget: int):
        self._page_budget = page_budget

    # Called once per forwarding in all layers: reserve KV pages and metadata entries.
    def prepare_metadata(self, seq_len: int):
        self._ops = self._step = None  # lengths change: whatever was bound for the last token is void
        appended_new_pages = self.kv_cache.append_seq(seq_len)
        _ = self.metadata_cache.append_seq(appended_new_pages)

    def _sync_table(self, table, cache, synced):
        """Copies the entries the device does not have yet.  The list only grows between two release() calls; a
        release (directly on the cache or through clean_states) bumps the cache's generation, and a list that is
        shorter than what was copied can only mean the same, so both resynchronise from entry 0 (the reference
        rebuilds the tensors on every begin_forward, controller.py:84-95)."""
        ids = cache.indicies
        n, (have, gen) = len(ids), synced
        if gen != cache.generation or n < have:
            have = 0
        if n > have:
            table[have:n] = torch.tensor(ids[have:n], dtype=torch.int32)
        return (n, cache.generation)

    def _indptr(self, n):
        t = self._indptr_cache.get(n)
        if t is None:
            if len(self._indptr_cache) > 64:
                self._indptr_cache.clear()
            t = torch.tensor([0, n], dtype=torch.int32, device=self.device)
            self._indptr_cache[n] = t
        return t

    def sync_tables(self):
        """Bring the device page tables up to date with the host lists (one small H2D copy per newly allocated page)
        and refresh the views / host ids derived from them."""
        kv_ids, meta_ids = self.kv_cache.indicies, self.metadata_cache.indicies
        self._kv_synced = self._sync_table(self._kv_table, self.kv_cache, self._kv_synced)
        self._meta_synced = self._sync_table(self._meta_table, self.metadata_cache, self._meta_synced)
        self.kv_indptr_for_append = self._indptr(len(kv_ids))
        self.metadata_indptr_for_append = self._indptr(len(meta_ids))
        self.kv_last_page_idx = kv_ids[-1]
        self.metadata_last_page_idx = meta_ids[-1]
        self.kv_indices_with_last = self._kv_table[:len(kv_ids)]
        self.metadata_indices = self._meta_table[:len(meta_ids)]

    # Prepare metadata used for inference under certain PAGE_BUDGET (called per budget change).
    def begin_forward(self, seq_len: int, updateTensor: bool = True):
        self._step = None
        self._ops = None
        kv_ids = self.kv_cache.indicies
        if updateTensor:
            self.sync_tables()

        if seq_len > 1:
            return  # prefill: append_kv_cache_prefill and prefill_with_paged_kv_cache need nothing else

        cur_page_nums = len(kv_ids)
        assert cur_page_nums > 1  # at least two pages for excluding last page
        if updateTensor:
            # top-k payload ids / full-attention page list: every page but the last, same row for all heads
            self.kv_indices_without_last = self._kv_table[:cur_page_nums - 1].unsqueeze(0).expand(self.num_heads, -1)

        # used as page_budget for topk and approx kernel
        self.inference_page_budget = min(self._page_budget, cur_page_nums)
        k = self.inference_page_budget - 1  # exclude the last page
        self.kv_indptr_for_approx_decode = self._indptr(k)
        bufs = self._topk_cache.get(k)
        if bufs is None:
            if len(self._topk_cache) > 8:
                self._topk_cache.clear()
            bufs = (torch.zeros((self.num_heads, k), dtype=self.dtype, device=self.device),
                    torch.zeros((self.num_heads, k), dtype=torch.int32, device=self.device))
            self._topk_cache[k] = bufs
        self.topk_dout_buffer, self.topk_dindices_buffer = bufs
        if self.topk_buf is None:
            # raft scratch in the reference (controller.py:120); unused here, kept so callers can pass it
            self.topk_buf = torch.zeros((self.num_heads, 1), dtype=self.dtype, device=self.device)

        self._decode_handler.begin_forward(self.kv_indptr_for_approx_decode, self.num_heads, self.num_kv_heads,
                                           self.head_dim, self.page_size, self.dtype, num_pages=k)

    def _scores_buffer(self, n):
        """[num_heads, n] fp16 scratch for decode_layer_step (decode_estimate allocates a fresh tensor per call, as the
        reference does): a view of one flat buffer that grows geometrically."""
        flat = getattr(self, "_scores_flat", None)
        need = self.num_heads * n
        if flat is None or flat.numel() < need:
            self._scores_flat = flat = torch.empty(max(need, 2 * (flat.numel() if flat is not None else 0)),
                                                   dtype=self.dtype, device=self.device)
        return flat[:need].view(self.num_heads, n)

    # Release the split-KV plan (reference frees handler buffers here).
    def end_forward(self):
        self._ops = None
        self._decode_handler.end_forward()

    def layer_ptrs(self):
        """(kv pool pointers, metadata pool pointers) per layer; the pools never move."""
        if self._pool_ptrs is None:
            layers = range(self.kv_cache.pool.num_layers)
            self._pool_ptrs = ([self.kv_cache.buf_layer(i).data_ptr() for i in layers],
                               [self.metadata_cache.buf_layer(i).data_ptr() for i in layers])
        return self._pool_ptrs

    def need_estimate(self) -> bool:
        if self.inference_page_budget is None:
            return False
        cur_page_nums = len(self.kv_cache.indicies)
        return cur_page_nums > self.inference_page_budget

    def clean_states(self):
        self._ops = self._step = None
        self.kv_cache.release()
        self.metadata_cache.release()
