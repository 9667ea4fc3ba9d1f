// One persistent kernel per layer-step: estimate -> top-k -> sparse decode (+ LSE merge) for MHA, D = 128.
// Replaces the three launches behind quest/utils/__init__.py:170-276 (estimate.cu:6-84, topk.cu:7-46,
// approx_attn.cu:27-151 in the reference) when the shape allows it; quest_sparse_decode_step (capi.cu) falls back
// to the three stand-alone kernels otherwise.
//
// Why one kernel: the stages are each a few microseconds of HBM streaming or pure latency; as separate launches
// every boundary drains and refills the memory pipeline.  Here 2 CTAs of 16 warps per SM stay resident and the
// heads are processed in two groups so that the latency-bound selection of one group hides under the streaming
// of the other:
//
//   all CTAs        : E(g0) E(g1)                 stream the metadata, one head group after the other
//   Hg CTAs / group :       T(g0) T(g1)           per-head radix select; publishes (tau, tie cut) as one word
//   all CTAs        :             D(g0) D(g1)     each (head, chunk) CTA ranks the head's scores against the published
//                                                 word itself, gathers its share of the selected pages, merges per head
//
// Dependencies are words in the workspace (release/acquire at gpu scope, bounded spins); nothing waits unless its
// producer is actually late.  The last CTA to finish zeroes the sync words for the next launch.
//
// Arithmetic is the stand-alone kernels' (estimate.cu, topk.cuh, decode.cu): scores and the selected set are
// bit-identical to theirs, the attention output differs only by the chunking of the merge.
#include <stdlib.h>

#include "common.cuh"

namespace quest {
// -DQUEST_FUSED_CLOCKS: every CTA stamps %globaltimer at its phase boundaries (tools/fused_timeline.py)
#ifdef QUEST_FUSED_CLOCKS
constexpr int kClkEvents = 16;
__device__ unsigned long long g_fused_clk[1024 * kClkEvents];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define F_STAMP(i) do { if (threadIdx.x == 0) g_fused_clk[blockIdx.x * kClkEvents + (i)] = gtime(); } while (0)
#define TOPK_FAST_STAMP(i) F_STAMP(5 + (i))
#else
#define F_STAMP(i) do { } while (0)
#endif
}  // namespace quest

#include "topk.cuh"

namespace quest {

constexpr int kFThreads = 512;
constexpr int kFWarps = kFThreads / 32;
constexpr float kFNegBig = -1e30f;
constexpr uint32_t kFMaxGroups = 2;
constexpr uint32_t kFSpinLimit = 1u << 21;  // x ~64 ns: a lost producer costs ~0.15 s and an error code, not a hang

// sync words inside the workspace's first kWsSyncBytes (u32 index)
enum : uint32_t { kSyncDone = 0, kSyncErr = 1, kSyncEst = 8 /* [kFMaxGroups] */, kSyncSel = 32 /* u64 [kFMaxHeads] */ };
constexpr uint32_t kFMaxHeads = 256;
static_assert((kSyncSel + 2 * kFMaxHeads) * 4 <= kWsSyncBytes, "sync words exceed the reserved area");

struct FusedArgs {
  const __half* q;
  __half* o;
  const __half* kv;
  const int32_t* kv_indices;
  const __half* meta;
  const int32_t* meta_indices;
  __half* scores;
  int32_t* topk_idx;
  uint8_t* ws;
  uint32_t len;            // scored pages (all but the last)
  uint32_t k;              // selected pages per head
  uint32_t last_page_idx, last_page_len;
  uint32_t H, P, p_magic;
  uint32_t Hg, hg_shift;   // heads per group (power of two dividing 32), log2
  uint32_t n_groups;
  uint32_t nch;            // decode chunks per head
  float sm_scale_log2;
};

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// One thread waits until *p >= target; false (and the error word set) if the producer never shows up.
__device__ __noinline__ bool spin_until_ge(const uint32_t* p, uint32_t target, uint32_t* err) {
  for (uint32_t it = 0; it < kFSpinLimit; ++it) {
    if (ld_acquire_u32(p) >= target) return true;
    __nanosleep(40);
  }
  atomicExch(err, 1u);
  return false;
}
// One thread waits for a non-zero 64-bit word; 0 on timeout.
__device__ __noinline__ unsigned long long spin_until_set(const unsigned long long* p, uint32_t* err) {
  for (uint32_t it = 0; it < kFSpinLimit; ++it) {
    const unsigned long long v = ld_acquire_u64(p);
    if (v) return v;
    __nanosleep(40);
  }
  atomicExch(err, 1u);
  return 0ull;
}

// CTAs [base(g), base(g) + Hg) run the selection of group g's heads and sit out the estimate phases after g.
__device__ __forceinline__ uint32_t topk_base(const FusedArgs& a, uint32_t g) { return g * a.Hg * a.nch; }

// Page range of CTA b in estimate phase g (empty if b is busy with an earlier group's selection).
__device__ __forceinline__ void est_range(const FusedArgs& a, uint32_t b, uint32_t grid, uint32_t my_tk, uint32_t g,
                                          uint32_t& pb, uint32_t& pe) {
  if (my_tk < g) {
    pb = pe = 0;
    return;
  }
  uint32_t rank = b, parts = grid;
  for (uint32_t e = 0; e < g; ++e) {  // groups whose selection CTAs are excluded from phase g
    parts -= a.Hg;
    if (topk_base(a, e) < b) rank -= a.Hg;
  }
  pb = (uint32_t)(((uint64_t)rank * a.len) / parts);
  pe = (uint32_t)(((uint64_t)(rank + 1) * a.len) / parts);
}
__device__ __forceinline__ uint32_t est_parts(const FusedArgs& a, uint32_t grid, uint32_t g) { return grid - g * a.Hg; }

template <int LPR>
__device__ __forceinline__ float f_reduce4_transposed(const float a[4], uint32_t lir) {
  const bool b_hi = lir & (LPR / 2), b_lo = lir & (LPR / 4);
  const float s0 = b_hi ? a[0] : a[2], s1 = b_hi ? a[1] : a[3];
  const float k0 = b_hi ? a[2] : a[0], k1 = b_hi ? a[3] : a[1];
  const float r0 = __fadd_rn(k0, __shfl_xor_sync(0xffffffffu, s0, LPR / 2));
  const float r1 = __fadd_rn(k1, __shfl_xor_sync(0xffffffffu, s1, LPR / 2));
  const float sd = b_lo ? r0 : r1, kp = b_lo ? r1 : r0;
  float v = __fadd_rn(kp, __shfl_xor_sync(0xffffffffu, sd, LPR / 4));
#pragma unroll
  for (int off = LPR / 8; off > 0; off >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

struct FusedSmem {
  float o[2 * kFWarps][128];   // decode: lane-group states; merge: row-group sums
  float md[2 * kFWarps][2];
  uint16_t keys[4 * kFThreads];
  TopkFastScratch<kFWarps> tk;
  uint32_t wsum[kFWarps];
  unsigned long long sel;
  uint32_t flag;
};

// ---- estimate phases ------------------------------------------------------------------------------------------
// Same arithmetic as estimate_tile_kernel<128, kOrdered, 1> (estimate.cu) restricted to one head group per phase:
// a pass covers 32 (page, head) rows = 32/Hg pages; 4 passes are in flight per lane (register ring); the issue
// cursor runs one ring ahead of the consumer and rolls over to the next group, so the stream never drains at a
// phase boundary.  Rings are aligned to phases (a ring never mixes two groups' rows).  The first ring is issued
// before the programmatic-launch wait: the scored metadata entries are not written by the preceding kernel
// (the QUEST_EST_META_STABLE contract, which quest_sparse_decode_step states for its callers).
template <int kOrdered>
__device__ __forceinline__ void fused_estimate(const FusedArgs& a, const int32_t* idx_s, uint32_t b, uint32_t grid,
                                               uint32_t my_tk, uint32_t* sync) {
  constexpr int D = 128, LPR = 16, kDepth = 4;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lir = lane % LPR, rsub = lane / LPR;
  const uint32_t rt = warp * 2 + rsub;                  // this lane's row in a pass (0..31)
  const uint32_t pg_in = rt >> a.hg_shift, hin = rt & (a.Hg - 1);
  const uint32_t PPP = 32u >> a.hg_shift;               // pages per pass
  const uint32_t row_bytes = a.H * D * 2, slab_bytes = a.P * row_bytes;
  const uint8_t* base = reinterpret_cast<const uint8_t*>(a.meta) + (size_t)hin * (D * 2) + lir * 16;
  const uint32_t g_last = my_tk < a.n_groups ? my_tk : a.n_groups - 1;  // last phase this CTA takes part in

  // issue cursor
  uint32_t g_ld = 0, p_ld, pe_ld;
  est_range(a, b, grid, my_tk, 0, p_ld, pe_ld);
  H8 mx[kDepth], mn[kDepth];
#pragma unroll
  for (int s = 0; s < kDepth; ++s)
#pragma unroll
    for (int i = 0; i < 4; ++i) mx[s].h[i] = mn[s].h[i] = __half2half2(__ushort_as_half(0));
  auto issue = [&](int s) {
    if (s == 0) {  // ring boundary: move on to the next phase that still has pages for this CTA
      while (p_ld >= pe_ld && g_ld < g_last) {
        ++g_ld;
        est_range(a, b, grid, my_tk, g_ld, p_ld, pe_ld);
      }
    }
    const uint32_t pq = p_ld + pg_in;
    const bool ok = pq < pe_ld;
    const uint32_t pc = min(pq, a.len - 1);
    const uint32_t mp = a.p_magic ? __umulhi(pc, a.p_magic) : pc;
    const uint32_t e = pc - mp * a.P;
    const uint8_t* ad = base + (size_t)g_ld * a.Hg * (D * 2) + (uint64_t)(uint32_t)idx_s[mp] * (2u * slab_bytes) +
                        e * row_bytes;
    ld_stream_if(mx[s], ad, ok);
    ld_stream_if(mn[s], ad + slab_bytes, ok);
    p_ld += PPP;
  };
#pragma unroll
  for (int s = 0; s < kDepth; ++s) issue(s);
  pdl_wait();  // q, the sync words and the selected pages' contents belong to earlier work on the stream
  F_STAMP(1);

  for (uint32_t g = 0; g <= g_last; ++g) {
    uint32_t p, pe;
    est_range(a, b, grid, my_tk, g, p, pe);
    if (p < pe) {
      const uint32_t h = g * a.Hg + hin;
      float qv[8];
      uint32_t mk[4];
      {
        const uint4 raw = *reinterpret_cast<const uint4*>(a.q + (size_t)h * D + lir * 8);
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
          qv[2 * i] = f.x;
          qv[2 * i + 1] = f.y;
          // sign-bit test: for q = -0 either choice gives a +-0 product, which cannot change the sum
          mk[i] = ((w[i] & 0x8000u) ? 0u : 0xffffu) | ((w[i] & 0x80000000u) ? 0u : 0xffff0000u);
        }
      }
      __half* orow = a.scores + (size_t)h * a.len;
      for (; p < pe; p += kDepth * PPP) {
        float sa[kDepth];
#pragma unroll
        for (int s = 0; s < kDepth; ++s) {
          float sacc = 0.f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint32_t hi, lo;
            if (kOrdered) {
              hi = *reinterpret_cast<const uint32_t*>(&mx[s].h[i]);
              lo = *reinterpret_cast<const uint32_t*>(&mn[s].h[i]);
            } else {
              const __half2 x = __hmax2(mx[s].h[i], mn[s].h[i]);
              const __half2 y = __hmin2(mx[s].h[i], mn[s].h[i]);
              hi = *reinterpret_cast<const uint32_t*>(&x);
              lo = *reinterpret_cast<const uint32_t*>(&y);
            }
            const uint32_t sel = (hi & mk[i]) | (lo & ~mk[i]);
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&sel));
            // reference order (decode_attn.cuh:152-160): 8 sequential fp32 adds of exact products
            sacc = __fmaf_rn(qv[2 * i], f.x, sacc);
            sacc = __fmaf_rn(qv[2 * i + 1], f.y, sacc);
          }
          sa[s] = sacc;
          issue(s);  // refill this stage from the next ring (possibly the next phase)
        }
        const float v = f_reduce4_transposed<LPR>(sa, lir);
        const uint32_t pp = p + pg_in + (lir / (LPR / 4)) * PPP;  // page whose score this lane holds
        if ((lir % (LPR / 4)) == 0 && pp < pe) orow[pp] = __float2half_rn(v);
      }
    }
    // publish this CTA's share of phase g (every participant signs, also with an empty range)
    __syncthreads();
    if (tid == 0) red_release_add(sync + kSyncEst + g, 1u);
    F_STAMP(2 + g);  // 2, 3: estimate phase g signed
  }
}

// ---- selection of one head by the whole CTA -------------------------------------------------------------------
// Loads the head's scores (written by other CTAs of this launch: through L2, never the non-coherent path) as
// order-preserving keys; thread t ends up owning positions 4t .. 4t+3.
__device__ __forceinline__ void fused_stage_keys(const FusedArgs& a, FusedSmem* sm, uint32_t h) {
  const uint32_t tid = threadIdx.x;
  const uint16_t* s = reinterpret_cast<const uint16_t*>(a.scores) + (size_t)h * a.len;
  const uint32_t len8 = (a.len + 7) & ~7u;
  uint16_t v[4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
    v[j] = tid + j * kFThreads < a.len ? __ldcg(s + tid + j * kFThreads) : (uint16_t)0xffffu;  // 0xffff -> key 0
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (tid + j * kFThreads < len8) sm->keys[tid + j * kFThreads] = key_of(v[j]);
}

__device__ __forceinline__ void fused_topk(const FusedArgs& a, FusedSmem* sm, uint32_t h, unsigned long long* publish) {
  fused_stage_keys(a, sm, h);
  __syncthreads();
  F_STAMP(4);  // keys staged
  topk_select_fast<kFWarps>(sm->keys, a.len, a.k, threadIdx.x, 1, &sm->tk, a.kv_indices,
                            a.topk_idx + (size_t)h * a.k, (uint16_t*)nullptr, publish);
  F_STAMP(9);  // ids written (this thread)
}

// ---- decode of one (head, chunk) item by the whole CTA ------------------------------------------------------------
// The chunk is the rank range [pg0, pg1) of the head's selected pages in ascending position (+ the forced last
// page as rank k).  Instead of waiting for the selection CTA to write the id list and reading it back, every decode
// CTA ranks the head's 2K scores itself against the published word (key > tau, or key == tau below the tie cut):
// the scores and page ids are fetched while the selection is still running, and what is left after the word
// arrives is one block scan.  Then decode_kernel's scheme (decode.cu) with 16 warps: a warp owns whole pages of the
// chunk, a half-warp owns a token row, 4-stage register ring of (K row, V row) pairs, two tokens per softmax
// rescale.  Partial to the workspace, ticket per head, last arriver merges.
__device__ __forceinline__ void fused_decode(const FusedArgs& a, FusedSmem* sm, int32_t* s_pages, uint32_t h,
                                             uint32_t chunk, uint32_t* sync) {
  constexpr int D = 128, LPR = 16, TPL = 2, NG = kFWarps * TPL, kRing = 4;
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t lir = lane % LPR, sub = lane / LPR, gid = warp * TPL + sub;
  const uint32_t total_pages = a.k + 1, nch = a.nch;
  const uint32_t pg0 = (uint32_t)(((uint64_t)chunk * total_pages) / nch);
  const uint32_t pg1 = (uint32_t)(((uint64_t)(chunk + 1) * total_pages) / nch);
  const uint32_t npg = pg1 - pg0;

  // fetched before the selection word is known: this thread's 4 scores (positions 4*tid ..) and their page ids
  const uint32_t i0 = tid * 4;
  uint32_t key[4];
  int32_t pid[4];
  {
    const uint16_t* s = reinterpret_cast<const uint16_t*>(a.scores) + (size_t)h * a.len;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool in = i0 + j < a.len;
      key[j] = in ? (uint32_t)key_of(__ldcg(s + i0 + j)) : 0u;
      pid[j] = in ? __ldg(a.kv_indices + i0 + j) : 0;
    }
  }
  float qv[8];
  {
    const H8 qh = ld_plain(a.q + (size_t)h * D + lir * 8);
    h8_to_float(qh, qv);
#pragma unroll
    for (int i = 0; i < 8; ++i) qv[i] *= a.sm_scale_log2;
  }
  if (tid == 0)
    sm->sel = spin_until_set(reinterpret_cast<const unsigned long long*>(sync + kSyncSel) + h, sync + kSyncErr);
  __syncthreads();
  const unsigned long long word = sm->sel;
  F_STAMP(12);  // selection word seen
  if (!word) return;
  const uint32_t tau = (uint32_t)(word & 0xffffu), cut = (uint32_t)((word >> 16) & 0x7fffffffu);
  uint32_t selm = 0, cnt = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool in = i0 + j < a.len;
    const bool s = in && (key[j] > tau || (key[j] == tau && i0 + j < cut));
    selm |= (uint32_t)s << j;
    cnt += s;
  }
  uint32_t inc = cnt;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t x = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= (uint32_t)off) inc += x;
  }
  if (lane == 31) sm->wsum[warp] = inc;
  __syncthreads();
  uint32_t rank = inc - cnt;
#pragma unroll
  for (int w2 = 0; w2 < kFWarps; ++w2)
    if ((uint32_t)w2 < warp) rank += sm->wsum[w2];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if ((selm >> j) & 1u) {
      if (rank >= pg0 && rank < pg1) s_pages[rank - pg0] = pid[j];
      ++rank;
    }
  }
  if (tid == 0 && pg1 == total_pages) s_pages[a.k - pg0] = (int32_t)a.last_page_idx;
  __syncthreads();

  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  float m = kFNegBig, d = 0.f;
  H8 kk[kRing], vv[kRing];
#pragma unroll
  for (int st = 0; st < kRing; ++st)
#pragma unroll
    for (int i = 0; i < 4; ++i) kk[st].h[i] = vv[st].h[i] = __half2half2(__ushort_as_half(0));
  const uint32_t row_bytes = a.H * D * 2, slab_bytes = a.P * row_bytes;
  const uint64_t page_bytes = 2ull * slab_bytes;
  const uint8_t* kvb = reinterpret_cast<const uint8_t*>(a.kv) + (size_t)h * D * 2 + lir * 16 + (size_t)sub * row_bytes;
  const uint32_t n_tl = (a.P + TPL - 1) / TPL;
  const uint32_t n_sel_local = a.k > pg0 ? a.k - pg0 : 0;
  uint32_t pg_ld = warp, i_ld = 0, okm = 0, nv_ld = 0;
  const uint8_t* ptr = kvb;
  auto open_page = [&]() {
    if (pg_ld < npg) {
      ptr = kvb + (uint64_t)(uint32_t)s_pages[pg_ld] * page_bytes;
      nv_ld = pg_ld < n_sel_local ? a.P : a.last_page_len;
    } else {
      nv_ld = 0;
    }
  };
  open_page();
  auto issue = [&](int st) {
    const bool ok = i_ld * TPL + sub < nv_ld;
    ld_stream_if(kk[st], ptr, ok);
    ld_stream_if(vv[st], ptr + slab_bytes, ok);
    okm = (okm & ~(1u << st)) | ((uint32_t)ok << st);
    ptr += TPL * row_bytes;
    if (++i_ld == n_tl) {
      i_ld = 0;
      pg_ld += kFWarps;
      open_page();
    }
  };
  auto consume2 = [&](int st) {
    float sc[2];
    float mt = m;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float kf[8];
      h8_to_float(kk[st + j], kf);
      float x = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) x = fmaf(qv[i], kf[i], x);
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
      sc[j] = (okm >> (st + j)) & 1u ? x : kFNegBig;
      mt = fmaxf(mt, sc[j]);
    }
    const float scale = ex2(m - mt);
    m = mt;
    d *= scale;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= scale;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float p = (okm >> (st + j)) & 1u ? ex2(sc[j] - m) : 0.f;
      d += p;
      float vf[8];
      h8_to_float(vv[st + j], vf);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = fmaf(p, vf[i], acc[i]);
    }
  };
  const uint32_t my_pages = npg > warp ? (npg - warp + kFWarps - 1) / kFWarps : 0;
  const uint32_t my_tl = my_pages * n_tl;
#pragma unroll
  for (int st = 0; st < kRing; ++st) issue(st);
  for (uint32_t t = 0; t < my_tl; t += kRing) {
    consume2(0);
    issue(0);
    issue(1);
    consume2(2);
    issue(2);
    issue(3);
  }
  F_STAMP(13);  // this thread's pages are consumed

  // CTA merge of the NG lane-group states
#pragma unroll
  for (int i = 0; i < 8; ++i) sm->o[gid][lir * 8 + i] = acc[i];
  if (lir == 0) {
    sm->md[gid][0] = m;
    sm->md[gid][1] = d;
  }
  __syncthreads();
  float* ws_o = reinterpret_cast<float*>(a.ws + ws_o_off(a.H, nch));
  float* ws_md = reinterpret_cast<float*>(a.ws + ws_md_off(a.H));
  uint32_t* ticket = reinterpret_cast<uint32_t*>(a.ws + ws_ticket_off());
  if (tid < D) {
    float M = kFNegBig, dsum = 0.f, osum = 0.f;
#pragma unroll
    for (int g = 0; g < NG; ++g) M = fmaxf(M, sm->md[g][0]);
#pragma unroll
    for (int g = 0; g < NG; ++g) {
      const float w = ex2(sm->md[g][0] - M);
      dsum += sm->md[g][1] * w;
      osum += sm->o[g][tid] * w;
    }
    if (nch == 1) {
      a.o[(size_t)h * D + tid] = __float2half_rn(__fdividef(osum, dsum));
    } else {
      ws_o[((size_t)h * nch + chunk) * D + tid] = osum;
      if (tid == 0) {
        ws_md[((size_t)h * nch + chunk) * 2] = M;
        ws_md[((size_t)h * nch + chunk) * 2 + 1] = dsum;
      }
    }
  }
  if (nch == 1) return;
  __syncthreads();
  if (tid == 0) {
    uint32_t old;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(ticket + h) : "memory");
    sm->flag = old == nch - 1;
  }
  __syncthreads();
  if (!sm->flag) return;
  // last arriver for this head: all partial rows in one round trip (16 rows per sweep x 4 sweeps in flight)
  constexpr int NV = D / 4, RPS = kFThreads / NV, kSw = 4;
  const uint32_t cr = tid / NV, c4 = tid % NV;
  const float4* po = reinterpret_cast<const float4*>(ws_o + (size_t)h * nch * D);
  const float2* pmd = reinterpret_cast<const float2*>(ws_md + (size_t)h * nch * 2);
  float4 a4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float MM = kFNegBig, dd = 0.f;
  for (uint32_t cb = 0; cb < nch; cb += RPS * kSw) {
    float4 v[kSw];
    float2 md[kSw];
#pragma unroll
    for (int i = 0; i < kSw; ++i) {
      const uint32_t c = cb + i * RPS + cr;
      v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      md[i] = make_float2(kFNegBig, 0.f);
      if (c < nch) {
        v[i] = __ldcg(po + (size_t)c * NV + c4);
        md[i] = __ldcg(pmd + c);
      }
    }
    float mb = MM;
#pragma unroll
    for (int i = 0; i < kSw; ++i) mb = fmaxf(mb, md[i].x);
    const float sc0 = ex2(MM - mb);
    MM = mb;
    dd *= sc0;
    a4.x *= sc0, a4.y *= sc0, a4.z *= sc0, a4.w *= sc0;
#pragma unroll
    for (int i = 0; i < kSw; ++i) {
      const float w = ex2(md[i].x - MM);
      dd += md[i].y * w;
      a4.x += v[i].x * w, a4.y += v[i].y * w, a4.z += v[i].z * w, a4.w += v[i].w * w;
    }
  }
  *reinterpret_cast<float4*>(&sm->o[cr][c4 * 4]) = a4;
  if (c4 == 0) {
    sm->md[cr][0] = MM;
    sm->md[cr][1] = dd;
  }
  __syncthreads();
  if (tid < D) {
    float Mf = kFNegBig, df = 0.f, of = 0.f;
#pragma unroll
    for (int g = 0; g < RPS; ++g) Mf = fmaxf(Mf, sm->md[g][0]);
#pragma unroll
    for (int g = 0; g < RPS; ++g) {
      const float w = ex2(sm->md[g][0] - Mf);
      df += sm->md[g][1] * w;
      of += sm->o[g][tid] * w;
    }
    a.o[(size_t)h * D + tid] = __float2half_rn(__fdividef(of, df));
    if (tid == 0) ticket[h] = 0;  // ready for the next launch
  }
}

// ---- the kernel -------------------------------------------------------------------------------------------------
template <int kOrdered>
__global__ void __launch_bounds__(kFThreads, 2) fused_step_kernel(const FusedArgs a) {
  extern __shared__ __align__(16) uint8_t f_dyn[];  // [n_meta] metadata page ids | [max pages per chunk] page ids
  __shared__ FusedSmem sm;
  const uint32_t tid = threadIdx.x, b = blockIdx.x, grid = gridDim.x;
  const uint32_t n_meta = (a.len + a.P - 1) / a.P;
  int32_t* idx_s = reinterpret_cast<int32_t*>(f_dyn);
  int32_t* s_pages = idx_s + n_meta;
  uint32_t* sync = reinterpret_cast<uint32_t*>(a.ws);
  unsigned long long* sel_words = reinterpret_cast<unsigned long long*>(sync + kSyncSel);

  // role: CTA base(g) + j runs the selection of head g*Hg + j
  uint32_t my_tk = 0xffffffffu, tk_head = 0;
  for (uint32_t g = 0; g < a.n_groups; ++g) {
    const uint32_t base = topk_base(a, g);
    if (b >= base && b < base + a.Hg) {
      my_tk = g;
      tk_head = g * a.Hg + (b - base);
    }
  }

  F_STAMP(0);
  pdl_trigger();  // the next kernel on the stream may be scheduled as this one's CTAs retire
  for (uint32_t i = tid; i < n_meta; i += kFThreads) idx_s[i] = __ldg(a.meta_indices + i);
  __syncthreads();

  fused_estimate<kOrdered>(a, idx_s, b, grid, my_tk, sync);  // contains the programmatic-launch wait

  if (my_tk != 0xffffffffu) {
    if (tid == 0) sm.flag = spin_until_ge(sync + kSyncEst + my_tk, est_parts(a, grid, my_tk), sync + kSyncErr);
    __syncthreads();
    F_STAMP(10);  // all scores of the group are in
    if (sm.flag) fused_topk(a, &sm, tk_head, sel_words + tk_head);
    __syncthreads();
    F_STAMP(11);
  }

  if (b < a.H * a.nch) {
    const uint32_t h = b / a.nch, chunk = b - h * a.nch, g = h >> a.hg_shift;
    // the head's scores are complete once every participant of phase g has signed
    if (tid == 0) sm.flag = spin_until_ge(sync + kSyncEst + g, est_parts(a, grid, g), sync + kSyncErr);
    __syncthreads();
    const bool ok = sm.flag;
    __syncthreads();  // sm.flag is reused by the merge
    if (ok) fused_decode(a, &sm, s_pages, h, chunk, sync);
    F_STAMP(14);
  }

  // the last CTA out zeroes the sync words for the next launch (everyone is past every wait by now)
  __syncthreads();
  if (tid == 0) {
    uint32_t old;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(sync + kSyncDone) : "memory");
    sm.flag = old == grid - 1;
  }
  __syncthreads();
  if (sm.flag) {
    for (uint32_t i = tid; i < kWsSyncBytes / 4; i += kFThreads)
      if (i != kSyncErr) sync[i] = 0;  // the error word is sticky: the host reads and clears it
  }
  F_STAMP(15);
}

static int fused_grid() {
  static int grid = -1;
  if (grid < 0) {
    int dev = 0, sms = 0, occ = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fused_step_kernel<1>, kFThreads, 4096) != cudaSuccess) occ = 0;
    if (occ > 2) occ = 2;
    grid = sms > 0 && occ > 0 ? sms * occ : 0;
  }
  return grid;
}

// QUEST_OK when launched; QUEST_ERR_UNSUPPORTED when the shape is outside the fused kernel's domain (the caller
// then uses the three stand-alone kernels).
int launch_fused_step(const void* q, void* o, const void* kv_data, const int32_t* kv_indices, uint32_t len,
                      uint32_t last_page_idx, uint32_t last_page_len, const void* meta_data,
                      const int32_t* meta_indices, void* scores, int32_t* topk_idx, uint32_t k, uint32_t Hq,
                      uint32_t Hkv, uint32_t D, uint32_t P, void* workspace, size_t workspace_bytes, uint32_t flags,
                      cudaStream_t stream) {
  static int enabled = -1, n_groups_env = 0, coop = 0;
  if (enabled < 0) {
    const char* e = getenv("QUEST_FUSED");
    enabled = (e && e[0] == '0') ? 0 : 1;
    if (const char* g = getenv("QUEST_FUSED_GROUPS")) n_groups_env = atoi(g);
    if (const char* c = getenv("QUEST_FUSED_COOP")) coop = atoi(c);
  }
  if (!enabled) return QUEST_ERR_UNSUPPORTED;
  if (Hq != Hkv || D != 128 || len == 0 || len > 4u * kFThreads || k == 0 || k > len) return QUEST_ERR_UNSUPPORTED;
  if (Hq > kFMaxHeads) return QUEST_ERR_UNSUPPORTED;
  const int grid = fused_grid();
  if (grid <= 0 || (uint32_t)grid < Hq) return QUEST_ERR_UNSUPPORTED;
  // two head groups when the heads split into 32-row tiles, else one
  uint32_t n_groups = n_groups_env > 0 ? (uint32_t)n_groups_env : 2;
  if (n_groups > kFMaxGroups || Hq % n_groups || 32 % (Hq / n_groups)) n_groups = 1;
  const uint32_t Hg = Hq / n_groups;
  if (32 % Hg) return QUEST_ERR_UNSUPPORTED;
  uint32_t nch = (uint32_t)grid / Hq;
  if (nch > k + 1) nch = k + 1;
  if (!workspace || workspace_bytes < ws_bytes(Hq, D, nch)) return QUEST_ERR_UNSUPPORTED;
  const uint32_t n_meta = (len + P - 1) / P;
  const uint32_t max_ppc = (k + 1 + nch - 1) / nch + 1;
  const size_t smem = ((size_t)n_meta + max_ppc) * sizeof(int32_t);
  if (smem > 40 * 1024) return QUEST_ERR_UNSUPPORTED;

  FusedArgs a;
  a.q = (const __half*)q, a.o = (__half*)o, a.kv = (const __half*)kv_data, a.kv_indices = kv_indices;
  a.meta = (const __half*)meta_data, a.meta_indices = meta_indices, a.scores = (__half*)scores, a.topk_idx = topk_idx;
  a.ws = (uint8_t*)workspace;
  a.len = len, a.k = k, a.last_page_idx = last_page_idx, a.last_page_len = last_page_len;
  a.H = Hq, a.P = P, a.p_magic = P <= 1 ? 0u : (uint32_t)((0x100000000ull + P - 1) / P);
  a.Hg = Hg, a.hg_shift = 31 - __builtin_clz(Hg), a.n_groups = n_groups, a.nch = nch;
  a.sm_scale_log2 = kLog2e / sqrtf((float)D);

  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  // The CTAs wait on one another, so all of them must become resident: the grid is exactly the occupancy limit
  // (2 x SMs), the kernels before it on the stream never wait on it, and every spin is bounded.  QUEST_FUSED_COOP=1
  // asks the driver to enforce co-residency (cooperative launch) instead of programmatic dependent launch.
  cudaLaunchAttribute attr[1];
  if (coop) {
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
  } else {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled();
  }
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = (flags & QUEST_EST_META_ORDERED) ? cudaLaunchKernelEx(&cfg, fused_step_kernel<1>, a)
                                                   : cudaLaunchKernelEx(&cfg, fused_step_kernel<0>, a);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

}  // namespace quest

#ifdef QUEST_FUSED_CLOCKS
extern "C" int quest_debug_fused_clocks(unsigned long long* host_out, uint32_t n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, quest::g_fused_clk, (size_t)n_ctas * quest::kClkEvents * 8);
}
#endif
