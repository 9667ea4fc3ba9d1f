"""Host-side paged pool bookkeeping: mirror of reference quest/utils/kv_cache.py:7-133.

Same classes, properties and allocation order (so a page table built here equals the reference's), but
``append_seq`` is O(new pages) instead of a Python loop over every token (kv_cache.py:120-125).
"""
import torch

from quest_b200.utils.utils import TensorLayout


class KvPool:
    """[num_layers, capacity, 2, block_len, num_heads, head_dim] pool + free list (kv_cache.py:7-67)."""

    def __init__(self, num_layers: int, num_heads: int, head_dim: int, capacity: int, block_len: int,
                 dtype: torch.dtype, device: torch.device):
        self._layout = TensorLayout.NHD
        self._buf = torch.empty((num_layers, capacity, 2, block_len, num_heads, head_dim), dtype=dtype, device=device)
        # The reference pops from set(range(capacity)) which yields 0,1,2,... for small ints; a sorted
        # stack keeps that order without depending on set internals.
        self._free = list(range(capacity - 1, -1, -1))
        self._in_use = set()

    @property
    def layout(self):
        return self._layout

    @property
    def buf(self):
        return self._buf

    @property
    def num_layers(self):
        return self._buf.shape[0]

    @property
    def block_len(self):
        return self._buf.shape[3]

    @property
    def num_free_blocks(self):
        return len(self._free)

    @property
    def capacity(self):
        return self._buf.shape[1]

    def alloc_block(self) -> int:
        if not self._free:
            raise KeyError("pop from an empty pool")  # the reference's set.pop() raises KeyError
        idx = self._free.pop()
        self._in_use.add(idx)
        return idx

    def free_block(self, idx: int):
        assert 0 <= idx < self.capacity
        assert idx in self._in_use
        self._in_use.remove(idx)
        self._free.append(idx)


class KvCache:
    """Key-value cache for one sequence (kv_cache.py:70-133)."""

    def __init__(self, num_layers, num_heads, head_dim, max_seq_len: int, page_size, dtype: torch.dtype,
                 device: torch.device):
        if max_seq_len <= 0:
            raise ValueError("init_len must be non-negative")
        self._pool = KvPool(num_layers=num_layers, num_heads=num_heads, head_dim=head_dim,
                            capacity=(max_seq_len + page_size - 1) // page_size, block_len=page_size, dtype=dtype,
                            device=device)
        self._indicies = []
        self._seqlen = 0
        self.generation = 0  # bumped by release(): device copies of the page table made before it are stale

    @property
    def pool(self) -> KvPool:
        return self._pool

    @property
    def seqlen(self) -> int:
        return self._seqlen

    @property
    def last_page_len(self) -> int:
        return (self.seqlen - 1) % self._pool.block_len + 1

    @property
    def indicies(self):
        return self._indicies

    def buf_layer(self, layer_idx: int):
        # one view per layer, made once: indexing the pool tensor builds a new view object on every call, four times per
        # layer and token on the op-by-op path (reference kv_cache.py:112-114 indexes every time)
        views = self.__dict__.get("_layer_views")
        if views is None:
            views = self._layer_views = [self._pool.buf[i] for i in range(self._pool.num_layers)]
        return views[layer_idx]

    def append_seq(self, seq_len: int) -> int:
        """Reserve space for tokens and return number of new pages"""
        if seq_len <= 0:
            return 0
        P = self._pool.block_len
        have = len(self._indicies)
        need = (self._seqlen + seq_len + P - 1) // P
        for _ in range(need - have):
            self._indicies.append(self._pool.alloc_block())
        self._seqlen += seq_len
        return need - have

    def release(self):
        """Release all blocks"""
        self._seqlen = 0
        for idx in self._indicies:
            self._pool.free_block(idx)
        self._indicies.clear()
        self.generation += 1
