"""Port of the reference's PyTorch "simulated Quest" decode step (evaluation/quest_attention.py:23-69, 72-240).
TEST / BASELINE INFRASTRUCTURE ONLY (see oracle/quest_oracle.py header): used as the timed CPU baseline
(bench.py cpu_baseline and --impl reference) and pinned against the real reference function by
tests/golden/make_golden.py -> tests/golden/torch_path_*.npz.

The reference file cannot travel to the GPU box, so this restates its dense algorithm op for op (same
tensor passes, so the timing is representative): full QK^T, sign trick, per-page amax of K*sign, second
matmul for the page bound, top-k over pages with k = min(max(3, budget // page), n_pages) where the last
page competes like any other (quest_attention.py:55-57), scatter to a token mask, masked fp32 softmax, @V.
Inputs are post-RoPE (the projections and RoPE are outside the attention path being measured).
"""
import math

import torch


def heavy_hitter_page_mask(page_bound, token_budget, chunk_size):
    """quest_attention.py:23-69 local_heavy_hitter_mask.  page_bound: [B, H, 1, T] per-token copy of the page
    upper bound.  Returns bool mask [B, H, 1, T]."""
    T = page_bound.shape[-1]
    pad = chunk_size - ((T - 1) % chunk_size + 1)
    filler = torch.full(page_bound.shape[:-1] + (pad,), torch.finfo(page_bound.dtype).min, dtype=page_bound.dtype)
    padded = torch.cat([page_bound, filler], dim=-1)
    per_page = padded.reshape(*padded.shape[:-1], padded.shape[-1] // chunk_size, chunk_size).amax(dim=-1)
    k = min(max(3, token_budget // chunk_size), per_page.size(-1))
    _, top = per_page.topk(k=k, dim=-1)
    tok = (top.unsqueeze(-1) * chunk_size + torch.arange(chunk_size)).reshape(*top.shape[:-1], -1)
    mask = torch.zeros_like(padded, dtype=torch.bool)
    mask.scatter_(-1, tok, True)
    return mask[..., :T]


def quest_torch_decode(q, k, v, token_budget, chunk_size, num_kv_groups=1):
    """quest_attention.py:139-224 for q_len == 1.  q: [B, Hq, 1, D]; k, v: [B, Hkv, T, D] (post-RoPE, whole
    cache incl. the new token).  Returns attention output [B, Hq, 1, D] (before o_proj)."""
    B, Hq, _, D = q.shape
    if num_kv_groups > 1:  # repeat_kv, :139-140
        k = k[:, :, None].expand(B, k.shape[1], num_kv_groups, k.shape[2], D).reshape(B, Hq, k.shape[2], D)
        v = v[:, :, None].expand(B, v.shape[1], num_kv_groups, v.shape[2], D).reshape(B, Hq, v.shape[2], D)
    T = k.shape[2]
    attn = torch.matmul(q, k.transpose(2, 3)) / math.sqrt(D)  # :142
    sign = (q > 0).to(q.dtype) * 2 - 1  # :146
    max_key = k * sign  # :147
    pos_q = q * sign  # :148
    pad = chunk_size - ((T - 1) % chunk_size + 1)
    filler = torch.full((B, Hq, pad, D), torch.finfo(max_key.dtype).min, dtype=max_key.dtype)
    mk = torch.cat([max_key, filler], dim=-2)  # :151-163
    page_max = mk.reshape(B, Hq, mk.shape[2] // chunk_size, chunk_size, D).amax(dim=-2)  # :166-172
    per_tok = page_max.unsqueeze(-2).repeat(1, 1, 1, chunk_size, 1).reshape(B, Hq, -1, D)[:, :, :T, :]  # :175-179
    bound = torch.matmul(pos_q.float(), per_tok.float().transpose(2, 3))  # :181-184
    budget = min(T, token_budget)  # :207
    mask = heavy_hitter_page_mask(bound, budget, chunk_size) if budget > 0 else torch.zeros_like(bound, dtype=torch.bool)
    attn = attn.masked_fill(~mask, torch.finfo(attn.dtype).min)  # :217-218 (tril is a no-op for the last position)
    p = torch.softmax(attn, dim=-1, dtype=torch.float32).to(q.dtype)  # :221-223
    return torch.matmul(p, v)  # :224
