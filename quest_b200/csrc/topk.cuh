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

// keys: shared, 16-byte aligned, readable (and zero-padded) up to a multiple of 8 elements past n.
// t: thread index inside the group [0, 32*W); bar_id: a named barrier reserved for this group.
// in_idx: payload ids (global; nullptr = position); out_idx/out_val: global rows of k entries.
template <int W>
__device__ __forceinline__ void topk_select(const uint16_t* keys, uint32_t n, uint32_t k, uint32_t t, uint32_t bar_id,
                                            TopkScratch<W>* sc, const int32_t* __restrict__ in_idx,
                                            int32_t* __restrict__ out_idx, uint16_t* __restrict__ out_val) {
  constexpr uint32_t T = 32 * W;
  const uint32_t lane = t & 31, warp = t >> 5;
  const uint32_t n8 = (n + 7) >> 3;                 // 8-key chunks
  const uint32_t cpt = (n8 + T - 1) / T;            // chunks per thread (blocked ownership)
  const uint32_t c0 = min(t * cpt, n8), c1 = min(c0 + cpt, n8);
  const uint4* kv = reinterpret_cast<const uint4*>(keys);

  // payload ids of this thread's slice, fetched now so the compaction does not wait on them
  const bool pre = in_idx != nullptr && cpt == 1;
  int32_t pay[8];
  if (pre) {
#pragma unroll
    for (int j = 0; j < 8; ++j) pay[j] = c0 * 8 + j < n ? __ldg(in_idx + c0 * 8 + j) : 0;
  }

  uint32_t prefix = 0, kk = k;
#pragma unroll 1
  for (int pass = 0; pass < 4; ++pass) {
    const uint32_t shift = 12 - 4 * pass;
    uint64_t cw0 = 0, cw1 = 0, cw2 = 0, cw3 = 0;
    for (uint32_t c = c0; c < c1; ++c) {
      const uint4 v = kv[c];
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t key = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
        const bool match = pass == 0 || (key >> (shift + 4)) == prefix;
        const uint32_t d = (key >> shift) & 15u;
        const uint64_t inc = match ? (1ull << ((d & 3u) * 16)) : 0ull;
        const uint32_t q = d >> 2;
        cw0 += q == 0 ? inc : 0ull;
        cw1 += q == 1 ? inc : 0ull;
        cw2 += q == 2 ? inc : 0ull;
        cw3 += q == 3 ? inc : 0ull;
      }
    }
    // warp totals: the packed 16-bit fields add without carries (a warp sees < 65536 keys), so 8 REDUX do 16 bins
    uint32_t mine = 0;
#pragma unroll
    for (int x = 0; x < 8; ++x) {
      const uint64_t cw = (x >> 1) == 0 ? cw0 : (x >> 1) == 1 ? cw1 : (x >> 1) == 2 ? cw2 : cw3;
      const uint32_t tot = __reduce_add_sync(0xffffffffu, (uint32_t)(cw >> ((x & 1) * 32)));
      if (lane == (uint32_t)x) mine = tot;
    }
    if (lane < 8) sc->hist[pass & 1][warp][lane] = mine;  // word x = bins 2x (low half), 2x+1 (high half)
    group_bar(bar_id, T);
    uint32_t s = 0;
    if (lane < 16) {
#pragma unroll
      for (int w2 = 0; w2 < W; ++w2) s += (sc->hist[pass & 1][w2][lane >> 1] >> ((lane & 1) * 16)) & 0xffffu;
    }
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
  }
  const uint32_t tau = prefix, need = kk;

  // stable compaction over the blocked ownership
  uint32_t ngt = 0, neq = 0;
  for (uint32_t c = c0; c < c1; ++c) {
    const uint4 v = kv[c];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t key = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
      const bool in = c * 8 + j < n;
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
    const uint4 v = kv[c];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t key = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
      const uint32_t i = c * 8 + j;
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
