"""CPU restatement of the reference's sparse decode-attention path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; the product (quest_b200/) never does and has no CPU path.

Parity pinning: the reference ships no golden vectors, so this oracle is pinned
against outputs of the reference's own Python code run in the authoring container
(tests/golden/make_golden.py -> tests/golden/*.npz; checked by tests/test_oracle_golden.py) and, on the
GPU box, against the reference's CUDA kernels compiled for sm_100a (oracle/_ref, tests/test_vs_ref_kernels.py).

Every function cites the reference file:line it restates.  Arrays are numpy; fp16 data stays np.float16,
arithmetic follows the reference's types (fp32 unless stated).
"""
import math

import numpy as np

MAX_HALF = np.float16(65504.0)  # CUDART_MAX_NORMAL_FP16, decode_page.cuh:430-431


# ----------------------------------------------------------------------------------------------------
# paged pool helpers (kernels/include/decode/decode_page.cuh:196-239; quest/utils/kv_cache.py:20-23)
# pool: np.float16 [pages, 2, P, H, D]   (NHD)
# ----------------------------------------------------------------------------------------------------
def new_pool(num_pages, page_size, num_heads, head_dim, fill=0.0):
    return np.full((num_pages, 2, page_size, num_heads, head_dim), fill, dtype=np.float16)


def append_paged_kv(kv_pool, kv_indices, kv_last_page_len, meta_pool, meta_indices, k, v):
    """Append the LAST len(k) tokens of the sequence and update per-page min/max metadata.

    K/V scatter: kernels/src/include/cpu_reference.h:294-327 (append_paged_kv_cache).
    Metadata: decode_page.cuh:398-449 (decode) and :471-562 (prefill): entry (meta_indices[j // P], j % P)
    of KV page j holds max in slab 0 and min in slab 1; initialised to -/+65504 when the page had no tokens
    before this call, otherwise resumed from the stored values (:424-432, :509-523).
    k, v: np.float16 [n, H, D].
    """
    P = kv_pool.shape[2]
    n = k.shape[0]
    seq_len = (len(kv_indices) - 1) * P + kv_last_page_len
    start = seq_len - n
    assert 0 <= start
    first_page = start // P
    for j in range(first_page, len(kv_indices)):
        page = kv_indices[j]
        e0 = max(0, start - j * P)
        e1 = min(P, seq_len - j * P)
        rows = slice(j * P + e0 - start, j * P + e1 - start)
        kv_pool[page, 0, e0:e1] = k[rows]
        kv_pool[page, 1, e0:e1] = v[rows]
        mp, me = meta_indices[j // P], j % P
        if e0 > 0:
            mx, mn = meta_pool[mp, 0, me], meta_pool[mp, 1, me]
        else:
            mx = np.full(k.shape[1:], -MAX_HALF, dtype=np.float16)
            mn = np.full(k.shape[1:], MAX_HALF, dtype=np.float16)
        meta_pool[mp, 0, me] = np.maximum(mx, k[rows].max(axis=0))
        meta_pool[mp, 1, me] = np.minimum(mn, k[rows].min(axis=0))


# ----------------------------------------------------------------------------------------------------
# estimate (decode_attn.cuh:137-168, 245-401): bit-exact emulation of the kernel's fp32 reduction order
# ----------------------------------------------------------------------------------------------------
def estimate_scores(q, meta_pool, meta_indices, num_scored_pages, num_qo_heads=None):
    """score[h, p] = fp16( sum_d max(q*max, q*min) ), p in [0, num_scored_pages).

    Order (decode_attn.cuh:152-160): each of D/8 lanes adds its 8 dims sequentially starting from 0.f,
    then an xor-butterfly over the lanes (offsets D/16 ... 1); the caller stores lane 0.
    q: np.float16 [Hq, D].  Returns np.float16 [Hq, num_scored_pages].
    """
    P, Hkv, D = meta_pool.shape[2], meta_pool.shape[3], meta_pool.shape[4]
    Hq = q.shape[0] if num_qo_heads is None else num_qo_heads
    G = Hq // Hkv
    n = num_scored_pages
    lpr = D // 8
    pages = np.arange(n)
    mp = np.asarray(meta_indices, dtype=np.int64)[pages // P]
    mx = meta_pool[mp, 0, pages % P].astype(np.float32)  # [n, Hkv, D]
    mn = meta_pool[mp, 1, pages % P].astype(np.float32)
    mx = np.repeat(mx, G, axis=1).reshape(n, Hq, lpr, 8)
    mn = np.repeat(mn, G, axis=1).reshape(n, Hq, lpr, 8)
    qf = q.astype(np.float32).reshape(1, Hq, lpr, 8)
    s = np.zeros((n, Hq, lpr), dtype=np.float32)
    for i in range(8):
        s = s + np.maximum(qf[..., i] * mx[..., i], qf[..., i] * mn[..., i])
    lanes = np.arange(lpr)
    off = lpr // 2
    while off > 0:
        s = s + s[..., lanes ^ off]
        off //= 2
    return np.ascontiguousarray(s[..., 0].astype(np.float16).T)


# ----------------------------------------------------------------------------------------------------
# fp8 (e4m3) metadata, SURVEY.md section 8 row f-4 (a variant of this framework; the reference's metadata is fp16 only,
# decode_page.cuh:79-352).  OCP FP8 E4M3: 1 sign, 4 exponent (bias 7), 3 mantissa bits, subnormals at exponent 0, largest
# finite 448 (0x7e), 0x7f / 0xff = NaN, no infinities.  The pool stores every max rounded UP and every min rounded DOWN.
# ----------------------------------------------------------------------------------------------------
def _e4m3_table():
    vals = np.zeros(256, dtype=np.float64)
    for b in range(256):
        sgn = -1.0 if b & 0x80 else 1.0
        e, m = (b >> 3) & 0xF, b & 7
        if e == 15 and m == 7:
            vals[b] = np.nan
        elif e == 0:
            vals[b] = sgn * (m / 8.0) * 2.0 ** -6
        else:
            vals[b] = sgn * (1.0 + m / 8.0) * 2.0 ** (e - 7)
    return vals


E4M3_VALUES = _e4m3_table()
_E4M3_FINITE = np.array([b for b in range(256) if not np.isnan(E4M3_VALUES[b]) and b != 0x80], dtype=np.uint8)  # -0 left out
_E4M3_ORDER = _E4M3_FINITE[np.argsort(E4M3_VALUES[_E4M3_FINITE], kind="stable")]
_E4M3_SORTED = E4M3_VALUES[_E4M3_ORDER]


def e4m3_round_up(x):
    """Byte of the smallest e4m3 value >= x (x fp16 array); values above 448 saturate to 448."""
    xf = np.asarray(x, dtype=np.float64)
    i = np.searchsorted(_E4M3_SORTED, xf, side="left")
    return _E4M3_ORDER[np.minimum(i, len(_E4M3_SORTED) - 1)]


def e4m3_round_down(x):
    """Byte of the largest e4m3 value <= x; values below -448 saturate to -448."""
    xf = np.asarray(x, dtype=np.float64)
    i = np.searchsorted(_E4M3_SORTED, xf, side="right") - 1
    return _E4M3_ORDER[np.maximum(i, 0)]


def metadata_to_fp8(meta_pool):
    """fp16 metadata pool [pages, 2, P, H, D] -> uint8 pool of the same shape: slab 0 (max) rounded up, slab 1 (min) down."""
    out = np.zeros(meta_pool.shape, dtype=np.uint8)
    out[:, 0] = e4m3_round_up(meta_pool[:, 0])
    out[:, 1] = e4m3_round_down(meta_pool[:, 1])
    return out


def fp8_pool_to_f16(meta8_pool):
    """The fp16 values an e4m3 pool stands for (every e4m3 value is exactly representable in fp16)."""
    return E4M3_VALUES[meta8_pool].astype(np.float16)


def estimate_scores_torch_ref(q, k, page_size):
    """The reference's own torch restatement (quest/tests/test_estimate.py:17-75), in numpy fp32:
    score[h, p] = sum_d |q_d| * max_{t in page p} (k_td * sign q_d), last page dropped.  Used to pin
    estimate_scores within the reference's tolerance (5e-3)."""
    qf = q.astype(np.float32)  # [H, D]
    kf = k.astype(np.float32)  # [T, H, D]
    T = kf.shape[0]
    sign = np.where(qf > 0, 1.0, -1.0).astype(np.float32)
    mk = kf * sign[None]
    n_pages = (T + page_size - 1) // page_size
    pad = n_pages * page_size - T
    if pad:
        mk = np.concatenate([mk, np.full((pad,) + mk.shape[1:], np.finfo(np.float32).min, np.float32)])
    pm = mk.reshape(n_pages, page_size, *mk.shape[1:]).max(axis=1)  # [n_pages, H, D]
    s = np.einsum("hd,phd->hp", qf * sign, pm)
    return s[:, : n_pages - 1].astype(np.float16)


# ----------------------------------------------------------------------------------------------------
# top-k (decode_select_k.cuh:25-62; raft select_radix.cuh:82-104 twiddle, :380-440 bucket/filter)
# ----------------------------------------------------------------------------------------------------
def radix_key(scores_f16):
    """Order-preserving 16-bit key of an fp16 bit pattern (raft twiddle_in, select_radix.cuh:82-90)."""
    bits = np.ascontiguousarray(scores_f16, dtype=np.float16).view(np.uint16).astype(np.uint32)
    return np.where(bits & 0x8000, (~bits) & 0xFFFF, bits | 0x8000).astype(np.uint32)


def topk_select(scores_f16, k):
    """Deterministic member of the reference's admissible top-k sets: all keys > tau, then ties at tau by
    lowest position; output positions ascending.  scores: [H, n] fp16 -> int32 [H, k]."""
    keys = radix_key(scores_f16)
    H, n = keys.shape
    out = np.empty((H, k), dtype=np.int32)
    for h in range(H):
        order = np.argsort(-keys[h].astype(np.int64), kind="stable")  # descending key, ties by position
        out[h] = np.sort(order[:k])
    return out


def topk_contract_ok(scores_f16, selected_pos, k):
    """The reference's contract: |S| = k distinct, {s > tau} subset S subset {s >= tau}."""
    keys = radix_key(scores_f16)
    for h in range(keys.shape[0]):
        sel = np.asarray(selected_pos[h], dtype=np.int64)
        if len(sel) != k or len(np.unique(sel)) != k or sel.min() < 0 or sel.max() >= keys.shape[1]:
            return False
        tau = np.sort(keys[h])[::-1][k - 1]
        chosen = np.zeros(keys.shape[1], dtype=bool)
        chosen[sel] = True
        if not np.all(chosen[keys[h] > tau]) or np.any(keys[h][chosen] < tau):
            return False
    return True


# ----------------------------------------------------------------------------------------------------
# sparse decode attention (cpu_reference.h:161-292 selected_single_mha; decode_attn.cuh:440-646)
# ----------------------------------------------------------------------------------------------------
def sparse_decode_attention(q, kv_pool, sel_pages, last_page_idx, last_page_len):
    """softmax(q.K/sqrt(D)) V over the tokens of sel_pages[h] (all full) + the last page (last_page_len
    tokens).  q: fp16 [Hq, D]; sel_pages: int [Hq, n_sel] physical page ids.  float64 accumulate.
    Returns (o fp32 [Hq, D], lse_base2 fp32 [Hq])."""
    P, Hkv, D = kv_pool.shape[2], kv_pool.shape[3], kv_pool.shape[4]
    Hq = q.shape[0]
    G = Hq // Hkv
    o = np.zeros((Hq, D), dtype=np.float64)
    lse = np.zeros(Hq, dtype=np.float64)
    sm = 1.0 / math.sqrt(D)
    sel_pages = np.asarray(sel_pages).reshape(Hq, -1)
    for h in range(Hq):
        kvh = h // G
        pages = sel_pages[h].astype(np.int64)
        K = kv_pool[pages, 0, :, kvh].reshape(-1, D)
        V = kv_pool[pages, 1, :, kvh].reshape(-1, D)
        K = np.concatenate([K, kv_pool[last_page_idx, 0, :last_page_len, kvh]]).astype(np.float64)
        V = np.concatenate([V, kv_pool[last_page_idx, 1, :last_page_len, kvh]]).astype(np.float64)
        s = K @ q[h].astype(np.float64) * sm
        mval = s.max()
        p = np.exp(s - mval)
        o[h] = (p / p.sum()) @ V
        lse[h] = (mval + math.log(p.sum())) * math.log2(math.e)
    return o.astype(np.float32), lse.astype(np.float32)


def dense_attention(q, k, v, causal=True):
    """cpu_reference.h:55-159 single_mha (NHD, kNone rotary).  q: [qo, Hq, D], k/v: [kv, Hkv, D] -> fp32 [qo, Hq, D]."""
    qo, Hq, D = q.shape
    kv_len, Hkv = k.shape[0], k.shape[1]
    G = Hq // Hkv
    out = np.zeros((qo, Hq, D), dtype=np.float32)
    sm = 1.0 / math.sqrt(D)
    for h in range(Hq):
        K = k[:, h // G].astype(np.float64)
        V = v[:, h // G].astype(np.float64)
        s = q[:, h].astype(np.float64) @ K.T * sm  # [qo, kv]
        if causal:
            pos = kv_len - qo + np.arange(qo)[:, None]
            s = np.where(np.arange(kv_len)[None, :] > pos, -np.inf, s)
        s = s - s.max(axis=1, keepdims=True)
        p = np.exp(s)
        out[:, h] = (p / p.sum(axis=1, keepdims=True)) @ V
    return out


def merge_states(o_parts, lse_parts):
    """flashinfer state.cuh:53-62 / cascade.cuh:279-340 in base 2: o_parts [n, H, D] normalised, lse [n, H]."""
    L = lse_parts.max(axis=0)
    w = np.exp2(lse_parts.astype(np.float64) - L)
    o = (o_parts.astype(np.float64) * w[..., None]).sum(axis=0) / w.sum(axis=0)[..., None]
    return o.astype(np.float32), (L + np.log2(w.sum(axis=0))).astype(np.float32)


# ----------------------------------------------------------------------------------------------------
# RoPE (cpu_reference.h:30-52 apply_llama_rope; decode_page.cuh:644-692) and RMSNorm (rms_norm.cu:82-158)
# ----------------------------------------------------------------------------------------------------
def rope(x, past_len, rope_scale=1.0, rope_theta=1e4):
    """x: fp16 [N, H, D]; position of row i = past_len + i.  Returns fp32."""
    N, H, D = x.shape
    xf = x.astype(np.float64)
    half = D // 2
    perm = np.concatenate([-xf[..., half:], xf[..., :half]], axis=-1)
    idx = np.arange(D) % half
    inv = 1.0 / (rope_theta ** (2.0 * idx / D))
    ang = ((past_len + np.arange(N)) / rope_scale)[:, None, None] * inv[None, None, :]
    return (np.cos(ang) * xf + np.sin(ang) * perm).astype(np.float32)


def rms_norm(x, w, eps):
    xf = x.astype(np.float64)
    inv = 1.0 / np.sqrt((xf * xf).mean(axis=-1, keepdims=True) + eps)
    return (xf * inv * w.astype(np.float64)).astype(np.float32)


# ----------------------------------------------------------------------------------------------------
# whole layer step with kernel-path semantics (quest/tests/test_approx_attention.py:17-110; controller.py:106-119)
# ----------------------------------------------------------------------------------------------------
def quest_decode_step(q, kv_pool, kv_indices, kv_last_page_len, meta_pool, meta_indices, page_budget):
    """estimate -> top-(B-1) among all pages but the last -> attention over them + the last page.
    Returns (o fp32, scores fp16, selected logical positions int32 [Hq, B-1])."""
    n_pages = len(kv_indices)
    budget = min(page_budget, n_pages)
    scores = estimate_scores(q, meta_pool, meta_indices, n_pages - 1)
    if n_pages > budget:
        sel = topk_select(scores, budget - 1)
    else:
        sel = np.tile(np.arange(n_pages - 1, dtype=np.int32), (q.shape[0], 1))
    phys = np.asarray(kv_indices, dtype=np.int32)[sel]
    o, _ = sparse_decode_attention(q, kv_pool, phys, kv_indices[-1], kv_last_page_len)
    return o, scores, sel
