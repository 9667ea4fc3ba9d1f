"""Shared builders for tests: seeded inputs and oracle-side paged caches."""
import numpy as np

from oracle import quest_oracle as orc


def golden_inputs(seed, kv_len, H, D, Hkv=None):
    """Must match tests/golden/make_golden.py:golden_inputs bit for bit."""
    rng = np.random.default_rng(seed)
    Hkv = H if Hkv is None else Hkv
    q = rng.standard_normal((1, H, D), dtype=np.float32).astype(np.float16)
    k = rng.standard_normal((kv_len, Hkv, D), dtype=np.float32).astype(np.float16)
    v = rng.standard_normal((kv_len, Hkv, D), dtype=np.float32).astype(np.float16)
    return q, k, v


class OracleCache:
    """One sequence in oracle pools with an arbitrary (shuffled) page table, as reference test_page.cu:37-40."""

    def __init__(self, max_pages, P, H, D, seed=0, shuffle=True):
        self.P, self.H, self.D = P, H, D
        max_meta = (max_pages + P - 1) // P
        rng = np.random.default_rng(seed)
        self.kv_free = list(rng.permutation(max_pages)) if shuffle else list(range(max_pages))
        self.meta_free = list(rng.permutation(max_meta)) if shuffle else list(range(max_meta))
        # NaN-poisoned pools: any read of an unwritten slot shows up
        self.kv_pool = orc.new_pool(max_pages, P, H, D, fill=np.nan)
        self.meta_pool = orc.new_pool(max_meta, P, H, D, fill=np.nan)
        self.kv_indices, self.meta_indices = [], []
        self.seqlen = 0

    @property
    def last_page_len(self):
        return (self.seqlen - 1) % self.P + 1

    @property
    def meta_last_page_len(self):
        return (len(self.kv_indices) - 1) % self.P + 1

    def reserve(self, n):
        need = (self.seqlen + n + self.P - 1) // self.P
        while len(self.kv_indices) < need:
            self.kv_indices.append(int(self.kv_free.pop(0)))
        need_m = (len(self.kv_indices) + self.P - 1) // self.P
        while len(self.meta_indices) < need_m:
            self.meta_indices.append(int(self.meta_free.pop(0)))
        self.seqlen += n

    def append(self, k, v):
        self.reserve(k.shape[0])
        orc.append_paged_kv(self.kv_pool, self.kv_indices, self.last_page_len, self.meta_pool, self.meta_indices, k, v)


PREFILL_IMPL = {"": 0, "simt": 1, "umma": 2, "tma": 3}


class prefill_impl:
    """``with prefill_impl("tma_smem"): ...`` -- selects a prefill kernel through the library's explicit debug switches
    (quest_debug_set_option; the library does not read the environment).  "": the library's own choice by shape."""

    def __init__(self, impl):
        self.impl = impl

    def __enter__(self):
        import quest_b200._kernels as K
        self.a = K.debug_option(K.DBG_PREFILL_IMPL, PREFILL_IMPL[self.impl.split("_")[0]])
        self.b = K.debug_option(K.DBG_PREFILL_P_SMEM, 1 if self.impl == "tma_smem" else 0)
        self.a.__enter__(), self.b.__enter__()
        return self

    def __exit__(self, *exc):
        self.b.__exit__(*exc), self.a.__exit__(*exc)
        return False
