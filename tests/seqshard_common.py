"""Builders shared by the sequence-sharding GPU tests (single-process emulation and the 2-rank NCCL worker)."""
import numpy as np
import torch

import quest_b200.utils as qu
from quest_b200 import parallel as par


def seeded_case(seed, kv_len, Hq, Hkv, D, steps):
    """fp16 K/V of the kv_len-1 prompt tokens plus `steps` decode tokens (q, k, v each)."""
    g = torch.Generator().manual_seed(seed)
    k = torch.randn(kv_len - 1, Hkv, D, generator=g).half()
    v = torch.randn(kv_len - 1, Hkv, D, generator=g).half()
    dec = [(torch.randn(1, Hq, D, generator=g).half(), torch.randn(1, Hkv, D, generator=g).half(),
            torch.randn(1, Hkv, D, generator=g).half()) for _ in range(steps)]
    return k, v, dec


def filled_controller(k, v, Hq, Hkv, D, P, budget, max_len, dev):
    """One-layer InferenceController holding the given token rows (through append_kv, as a prefill would)."""
    ic = qu.InferenceController(1, Hq, D, P, budget, max_len, torch.float16, dev, num_kv_heads=Hkv)
    n = k.shape[0]
    if n:
        ic.prepare_metadata(n)
        ic.begin_forward(n)
        qu.append_kv(k.to(dev), v.to(dev), ic, 0)
        ic.end_forward()
    return ic


def shard_token_ranges(kv_len, P, world):
    """Page ranges [a, b) of the scored pages per rank when the prompt holds kv_len-1 tokens and the first decode
    token brings the sequence to kv_len; the last rank also takes everything after its last scored page."""
    n_pages = (kv_len + P - 1) // P
    return [par.page_shard(n_pages - 1, world, r) for r in range(world)]


def shard_controller(k, v, rng, rank, world, Hq, Hkv, D, P, budget, extra_tokens, dev):
    a, b = rng
    t0 = a * P
    t1 = k.shape[0] if rank == world - 1 else b * P
    return filled_controller(k[t0:t1], v[t0:t1], Hq, Hkv, D, P, budget, (t1 - t0) + extra_tokens + P, dev)


def selected_positions(sel, table, first_page):
    """Global page positions of a shard's selected physical ids (sel: int32 [Hq, k] with -1 for other shards)."""
    lut = {int(p): first_page + i for i, p in enumerate(table)}
    return [sorted(lut[int(x)] for x in row if x >= 0) for row in sel.cpu().numpy()]
