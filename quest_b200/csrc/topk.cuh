// Cooperative top-k selection over 16-bit order-preserving keys held in shared memory, for a group of W warps.
// Used by the stand-alone top-k kernel (topk.cu) and by the fused layer-step kernel (fused.cu).
//
// Replaces raft radix_topk_one_block_kernel<half,int,8,512> behind decode_select_k
// (reference kernels/include/topk/decode_select_k.cuh:25-62; raft select_radix.cuh:82-104 key twiddle,
// :380-399 choose_bucket, :404-440 last_filter).  MSB radix select, 4 passes of 4 bits.  No shared-memory
// atomics (2 cycles per lane on this part): every thread counts its own slice into 16 packed 16-bit
// counters held in registers, warps reduce with REDUX, and one barrier per pass publishes the 16 bin totals.
// Selection rule (deterministic, unlike raft's atomicAdd race among ties): every key > tau is taken, keys
// == tau are taken lowest position first, output is in ascending input position.
#pragma once
#include "common.cuh"

namespace quest {

#ifdef QUEST_TOPK_CLOCKS
__device__ long long g_topk_clk[64];
#define TOPK_STAMP(i) do { if (t == 0 && blockIdx.x == 0) g_topk_clk[i] = clock64(); } while (0)
#else
#define TOPK_STAMP(i) do { } while (0)
#endif

__device__ __forceinline__ uint16_t key_of(uint16_t h) { return (h & 0x8000u) ? (uint16_t)~h : (uint16_t)(h | 0x8000u); }
__device__ __forceinline__ uint16_t half_of(uint16_t k) { return (k & 0x8000u) ? (uint16_t)(k & 0x7fffu) : (uint16_t)~k; }

__device__ __forceinline__ void group_bar(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <int W>
struct TopkScratch {
  uint32_t hist[2][W][8];  // per warp: 16 bins as packed 16-bit pairs
  uint32_t wgt[W], weq[W];
};

__device__ __forceinline__ uint64_t shl64_clamped(uint32_t s) {  // 1 << s, 0 for s >= 64 (PTX shl clamps)
  uint64_t r;
  asm("shl.b64 %0, %1, %2;" : "=l"(r) : "l"(1ull), "r"(s));
  return r;
}

#ifndef TOPK_FAST_STAMP
#define TOPK_FAST_STAMP(i) do { } while (0)  // fused.cu's timeline build stamps the passes
#endif

template <int W, int G = 1>
struct TopkFastScratch {
  uint8_t histT[2][16][W * G];  // [pass parity][bin][warp * G + key group]: bin totals of 128 keys (<= 128 each)
  uint32_t wgt[W], weq[W];
};

// Fast path for n <= 4 * G * 32 * W: every thread owns 4*G consecutive keys (2*G packed words) held in registers
// for the whole selection.  The counting loop is bound by the integer pipe (half rate on this part), so it is
// written for instruction count: per key a 128-bit one-hot (8-bit field per bin) is built with two clamped
// 64-bit shifts and added into 4 words per group of 4 keys (fields cannot carry: a thread adds at most 4 per
// group, a warp at most 128, so REDUX runs on the 8-bit fields directly); bin totals are published transposed so
// that lane b sums its bin over all warps and groups with 128-bit loads and dp4a.
// G > 1 (8 or 16 keys per thread: rows of up to 8192 / 16384 pages on 32 warps) needs the key array zero-padded to
// the full 4*G*32*W entries.
// Hook: called by every thread of the group once the selection is decided and before the ids are written, with
// the selection as one word for consumers that hold the keys themselves (fused.cu): position i is selected iff
// key > tau or (key == tau and i < tie_cut); word = 1<<63 | tie_cut<<16 | tau.  `holder` is true for the one thread
// that owns the need-th tie (need >= 1 always), which is the thread that knows tie_cut.
struct TopkNoHook {
  __device__ __forceinline__ void operator()(bool, unsigned long long) const {}
};

// Xc: how the CTAs of a thread-block cluster that share one row combine their bin counts and output offsets
// (topk.cu: TopkClusterXc over distributed shared memory); the default is a single CTA per row.
struct TopkNoXc {
  __device__ __forceinline__ uint32_t all_bins(uint32_t s, uint32_t, int) const { return s; }
  __device__ __forceinline__ uint32_t before_me(uint32_t, uint32_t) const { return 0u; }
};

// pos0: position of this CTA's first key in the row (0 unless a cluster splits the row); n is the row length.
template <int W, class Hook = TopkNoHook, int G = 1, class Xc = TopkNoXc>
__device__ __forceinline__ void topk_select_fast(const uint16_t* keys, uint32_t n, uint32_t k, uint32_t t,
                                                 uint32_t bar_id, TopkFastScratch<W, G>* sc,
                                                 const int32_t* __restrict__ in_idx, int32_t* __restrict__ out_idx,
                                                 uint16_t* __restrict__ out_val, Hook hook = Hook(), Xc xc = Xc(),
                                                 uint32_t pos0 = 0) {
  static_assert(W == 4 || W == 8 || W == 16 || W == 32, "W bytes per bin row are read as 32-bit words");
  static_assert(G == 1 || G == 2 || G == 4, "4, 8 or 16 keys per thread");
  constexpr uint32_t T = 32 * W;
  constexpr int NW = 2 * G, NK = 4 * G;
  const uint32_t lane = t & 31, warp = t >> 5;
  const uint32_t i0 = pos0 + t * NK;  // row position of this thread's first key
  uint32_t wd[NW];
#pragma unroll
  for (int x = 0; x < NW; ++x) wd[x] = 0;
  if (G == 1) {  // keys are zero-padded to a multiple of 8 past n
    if (i0 - pos0 < ((n + 7) & ~7u)) {
      const uint2 v = reinterpret_cast<const uint2*>(keys)[t];
      wd[0] = v.x, wd[1] = v.y;
    }
  } else {  // keys are zero-padded to the whole array
#pragma unroll
    for (int x = 0; x < G / 2; ++x) {
      const uint4 v = reinterpret_cast<const uint4*>(keys)[t * (G / 2) + x];
      wd[4 * x] = v.x, wd[4 * x + 1] = v.y, wd[4 * x + 2] = v.z, wd[4 * x + 3] = v.w;
    }
  }
  // payload ids fetched now so the compaction does not wait on them (G = 4: fetched at the end, registers)
  constexpr bool kPre = G <= 2;
  int32_t pay[kPre ? NK : 1];
  if (kPre) {
#pragma unroll
    for (int j = 0; j < NK; ++j) pay[j] = i0 + j < n ? (in_idx ? __ldg(in_idx + i0 + j) : (int32_t)(i0 + j)) : 0;
  }

  uint32_t prefix = 0, kk = k;
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) {
    constexpr uint32_t kOnes = 0x00010001u;
    const uint32_t shift = 12 - 4 * pass;
    const uint32_t hi_mask = (0xffffu >> (shift + 4)) * kOnes, pp = prefix * kOnes;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      uint32_t c[4] = {0, 0, 0, 0};
#pragma unroll
      for (int x = 2 * g; x < 2 * g + 2; ++x) {
        const uint32_t y = ((wd[x] >> (shift + 4)) & hi_mask) ^ pp;      // 0 in a half <=> that key matches the prefix
        const uint32_t d8lo = ((wd[x] << 3) >> shift) & 0x78u;           // 8 * digit of the low key
        const uint32_t d8hi = ((wd[x] >> 13) >> shift) & 0x78u;          // 8 * digit of the high key
        const uint32_t slo = (y & 0xffffu) == 0 ? d8lo : 255u, shi = (y >> 16) == 0 ? d8hi : 255u;
        const uint64_t a0 = shl64_clamped(slo), a1 = shl64_clamped(slo - 64u);
        const uint64_t b0 = shl64_clamped(shi), b1 = shl64_clamped(shi - 64u);
        c[0] += (uint32_t)a0 + (uint32_t)b0;
        c[1] += (uint32_t)(a0 >> 32) + (uint32_t)(b0 >> 32);
        c[2] += (uint32_t)a1 + (uint32_t)b1;
        c[3] += (uint32_t)(a1 >> 32) + (uint32_t)(b1 >> 32);
      }
      // zero-padded / out-of-range keys: key 0 never matches a non-zero prefix and sits in the lowest bin otherwise,
      // below every real key, so it cannot change which bin holds the k-th largest.
      uint32_t tot[4];
#pragma unroll
      for (int x = 0; x < 4; ++x) tot[x] = __reduce_add_sync(0xffffffffu, c[x]);  // warp-uniform
      if (lane < 16) {
        const uint32_t wsel = (lane >> 2) == 0 ? tot[0] : (lane >> 2) == 1 ? tot[1] : (lane >> 2) == 2 ? tot[2] : tot[3];
        sc->histT[pass & 1][lane][warp * G + g] = (uint8_t)(wsel >> ((lane & 3) * 8));
      }
    }
    group_bar(bar_id, T);
    uint32_t s = 0;
    if (lane < 16) {
      const uint32_t* row = reinterpret_cast<const uint32_t*>(sc->histT[pass & 1][lane]);
#pragma unroll
      for (int x = 0; x < W * G / 4; ++x) s = __dp4a(row[x], 0x01010101u, s);
    }
    s = xc.all_bins(s, lane, pass);  // the row's bin totals when several CTAs share it
    uint32_t suf = s;  // suffix sums over the 16 bin lanes (lanes 16..31 carry zeros)
#pragma unroll
    for (int off = 1; off < 16; off <<= 1) {
      const uint32_t x = __shfl_down_sync(0xffffffffu, suf, off);
      if (lane + off < 16) suf += x;
    }
    const uint32_t above = suf - s;
    const bool hit = lane < 16 && above < kk && above + s >= kk;
    const uint32_t b = __ffs(__ballot_sync(0xffffffffu, hit)) - 1;
    kk -= __shfl_sync(0xffffffffu, above, b);
    prefix = (prefix << 4) | b;
    TOPK_FAST_STAMP(pass);
  }
  const uint32_t tau = prefix, need = kk;

  uint32_t key[NK];
#pragma unroll
  for (int x = 0; x < NW; ++x) key[2 * x] = wd[x] & 0xffffu, key[2 * x + 1] = wd[x] >> 16;
  uint32_t ngt = 0, neq = 0;
#pragma unroll
  for (int j = 0; j < NK; ++j) {
    const bool in = i0 + j < n;
    ngt += in && key[j] > tau;
    neq += in && key[j] == tau;
  }
  uint32_t pk = (ngt << 16) | neq, spk = pk;  // both counts < 65536: scan them as one word
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, spk, off);
    if (lane >= (uint32_t)off) spk += a;
  }
  if (lane == 31) sc->wgt[warp] = spk;
  group_bar(bar_id, T);
  uint32_t before = spk - pk, cta_total = 0;
#pragma unroll
  for (int w2 = 0; w2 < W; ++w2) {
    const uint32_t x = sc->wgt[w2];
    cta_total += x;
    if ((uint32_t)w2 < warp) before += x;
  }
  before += xc.before_me(cta_total, t);  // (greater, equal) counts of the CTAs that hold earlier positions
  uint32_t gt_before = before >> 16, eq_before = before & 0xffffu;
  {
    uint32_t e = eq_before, cut = 0;
#pragma unroll
    for (int j = 0; j < NK; ++j)
      if (i0 + j < n && key[j] == tau && ++e == need) cut = i0 + j + 1;
    hook(cut != 0, (1ull << 63) | ((unsigned long long)cut << 16) | tau);
  }
#pragma unroll
  for (int j = 0; j < NK; ++j) {
    const bool in = i0 + j < n;
    const bool gt = in && key[j] > tau, eq = in && key[j] == tau;
    if (gt || (eq && eq_before < need)) {
      const uint32_t pos = gt_before + min(eq_before, need);
      out_idx[pos] = kPre ? pay[kPre ? j : 0] : in_idx ? __ldg(in_idx + i0 + j) : (int32_t)(i0 + j);
      if (out_val) out_val[pos] = half_of((uint16_t)key[j]);
    }
    gt_before += gt;
    eq_before += eq;
  }
}

// keys: shared, 16-byte aligned, readable (and zero-padded) up to a multiple of 8 elements past n.
// t: thread index inside the group [0, 32*W); bar_id: a named barrier reserved for this group.
// in_idx: payload ids (global; nullptr = position); out_idx/out_val: global rows of k entries.
// KPC: keys per ownership chunk (4 or 8).  The whole selection is a chain of dependent short instructions, so
// its time is (instructions per warp) x (issue latency): small chunks on more warps shorten the chain.
template <int W, int KPC>
__device__ __forceinline__ void topk_select(const uint16_t* keys, uint32_t n, uint32_t k, uint32_t t, uint32_t bar_id,
                                            TopkScratch<W>* sc, const int32_t* __restrict__ in_idx,
                                            int32_t* __restrict__ out_idx, uint16_t* __restrict__ out_val) {
  static_assert(KPC == 4 || KPC == 8, "chunk of 4 or 8 keys");
  constexpr uint32_t T = 32 * W;
  constexpr int WPC = KPC / 2;  // 32-bit words per chunk
  const uint32_t lane = t & 31, warp = t >> 5;
  const uint32_t nc = (n + KPC - 1) / KPC;          // chunks
  const uint32_t cpt = (nc + T - 1) / T;            // chunks per thread (blocked ownership)
  const uint32_t c0 = min(t * cpt, nc), c1 = min(c0 + cpt, nc);
  const uint32_t* kw = reinterpret_cast<const uint32_t*>(keys);
  auto load_chunk = [&](uint32_t c, uint32_t (&w)[WPC]) {
    if (KPC == 8) {
      const uint4 v = reinterpret_cast<const uint4*>(kw)[c];
      w[0] = v.x, w[1] = v.y, w[WPC - 2] = v.z, w[WPC - 1] = v.w;
    } else {
      const uint2 v = reinterpret_cast<const uint2*>(kw)[c];
      w[0] = v.x, w[1] = v.y;
    }
  };

  // payload ids of this thread's slice, fetched now so the compaction does not wait on them
  const bool pre = in_idx != nullptr && cpt == 1;
  int32_t pay[KPC];
  if (pre) {
#pragma unroll
    for (int j = 0; j < KPC; ++j) pay[j] = c0 * KPC + j < n ? __ldg(in_idx + c0 * KPC + j) : 0;
  }

  uint32_t prefix = 0, kk = k;
#pragma unroll 1
  for (int pass = 0; pass < 4; ++pass) {
    const uint32_t shift = 12 - 4 * pass;
    TOPK_STAMP(pass * 8 + 0);
    // 16 bins as 8-bit fields of 4 words, flushed to 16-bit fields (8 words) every 31 chunks
    uint32_t acc16[8];
#pragma unroll
    for (int x = 0; x < 8; ++x) acc16[x] = 0;
    for (uint32_t cb = c0; cb < c1; cb += 31) {
      uint32_t c8[4] = {0, 0, 0, 0};
      const uint32_t ce = min(cb + 31, c1);
      for (uint32_t c = cb; c < ce; ++c) {
        uint32_t w[WPC];
        load_chunk(c, w);
#pragma unroll
        for (int j = 0; j < KPC; ++j) {
          const uint32_t key = (j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xffffu);
          const bool match = pass == 0 || (key >> (shift + 4)) == prefix;
          const uint32_t d = (key >> shift) & 15u;
          const uint32_t inc = match ? (1u << ((d & 3u) * 8)) : 0u;
          const uint32_t q = d >> 2;
          c8[0] += q == 0 ? inc : 0u;
          c8[1] += q == 1 ? inc : 0u;
          c8[2] += q == 2 ? inc : 0u;
          c8[3] += q == 3 ? inc : 0u;
        }
      }
#pragma unroll
      for (int x = 0; x < 4; ++x) {  // word x holds bins 4x..4x+3
        acc16[2 * x] += c8[x] & 0x00ff00ffu;             // bins 4x (low half), 4x+2 (high half)
        acc16[2 * x + 1] += (c8[x] >> 8) & 0x00ff00ffu;  // bins 4x+1, 4x+3
      }
    }
    TOPK_STAMP(pass * 8 + 1);
    // warp totals: the packed 16-bit fields add without carries (a warp sees < 65536 keys), so 8 REDUX do 16 bins
    uint32_t mine = 0;
#pragma unroll
    for (int x = 0; x < 8; ++x) {
      const uint32_t tot = __reduce_add_sync(0xffffffffu, acc16[x]);
      if (lane == (uint32_t)x) mine = tot;
    }
    TOPK_STAMP(pass * 8 + 2);
    if (lane < 8) sc->hist[pass & 1][warp][lane] = mine;
    group_bar(bar_id, T);
    TOPK_STAMP(pass * 8 + 3);
    // lane b < 16 gathers bin b: word 2*(b/4) + (b&1), half (b>>1)&1
    uint32_t s = 0;
    if (lane < 16) {
      const uint32_t word = 2 * (lane >> 2) + (lane & 1), sh = ((lane >> 1) & 1) * 16;
#pragma unroll
      for (int w2 = 0; w2 < W; ++w2) s += (sc->hist[pass & 1][w2][word] >> sh) & 0xffffu;
    }
    TOPK_STAMP(pass * 8 + 4);
    // above = sum of bins higher than mine (lanes 16..31 carry zeros)
    uint32_t suf = s;
#pragma unroll
    for (int off = 1; off < 16; off <<= 1) {
      const uint32_t x = __shfl_down_sync(0xffffffffu, suf, off);
      if (lane + off < 16) suf += x;
    }
    const uint32_t above = suf - s;
    const bool hit = lane < 16 && above < kk && above + s >= kk;
    const uint32_t b = __ffs(__ballot_sync(0xffffffffu, hit)) - 1;  // exactly one lane hits while 1 <= kk <= candidates
    kk -= __shfl_sync(0xffffffffu, above, b);
    prefix = (prefix << 4) | b;
    TOPK_STAMP(pass * 8 + 5);
  }
  const uint32_t tau = prefix, need = kk;

  // stable compaction over the blocked ownership
  uint32_t ngt = 0, neq = 0;
  for (uint32_t c = c0; c < c1; ++c) {
    uint32_t w[WPC];
    load_chunk(c, w);
#pragma unroll
    for (int j = 0; j < KPC; ++j) {
      const uint32_t key = (j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xffffu);
      const bool in = c * KPC + j < n;
      ngt += in && key > tau;
      neq += in && key == tau;
    }
  }
  uint32_t sgt = ngt, seq = neq;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, sgt, off);
    const uint32_t b2 = __shfl_up_sync(0xffffffffu, seq, off);
    if (lane >= (uint32_t)off) {
      sgt += a;
      seq += b2;
    }
  }
  if (lane == 31) {
    sc->wgt[warp] = sgt;
    sc->weq[warp] = seq;
  }
  group_bar(bar_id, T);
  uint32_t gt_before = sgt - ngt, eq_before = seq - neq;
#pragma unroll
  for (int w2 = 0; w2 < W; ++w2) {
    if ((uint32_t)w2 < warp) {
      gt_before += sc->wgt[w2];
      eq_before += sc->weq[w2];
    }
  }
  for (uint32_t c = c0; c < c1; ++c) {
    uint32_t w[WPC];
    load_chunk(c, w);
#pragma unroll
    for (int j = 0; j < KPC; ++j) {
      const uint32_t key = (j & 1) ? (w[j >> 1] >> 16) : (w[j >> 1] & 0xffffu);
      const uint32_t i = c * KPC + j;
      const bool in = i < n;
      const bool gt = in && key > tau, eq = in && key == tau;
      if (gt || (eq && eq_before < need)) {
        const uint32_t pos = gt_before + min(eq_before, need);
        out_idx[pos] = pre ? pay[j] : in_idx ? __ldg(in_idx + i) : (int32_t)i;
        if (out_val) out_val[pos] = half_of((uint16_t)key);
      }
      gt_before += gt;
      eq_before += eq;
    }
  }
}

}  // namespace quest
