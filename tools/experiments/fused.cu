// One persistent kernel per layer-step: estimate -> top-k -> sparse decode (+ LSE merge) for MHA, D = 128.
// Replaces the three launches behind quest/utils/__init__.py:170-276 (estimate.cu:6-84, topk.cu:7-46,
// approx_attn.cu:27-151 in the reference) when the shape allows it; quest_sparse_decode_step (capi.cu) falls back
// to the three stand-alone kernels otherwise.
//
// Why one kernel: the stages are each a few microseconds of HBM streaming or pure latency; as separate launches
// every boundary drains and refills the memory pipeline.  Here every q head owns one thread-block cluster of 8 CTAs
// (16 warps each, 2 CTAs per SM) that stays resident for the whole step, and the heads are processed in two groups
// so that the latency-bound selection of one group hides under the streaming of the other:
//
//   all CTAs          : E(g0) E(g1)               stream the metadata, one head group after the other
//   rank 0 of a head  :       T(g0) T(g1)         radix select of the head; the (tau, tie cut) word goes to the 7 peers
//                                                 through distributed shared memory + one cluster barrier
//   the head's cluster:             D(g0) D(g1)   every CTA ranks the head's scores against the word itself, gathers its
//                                                 eighth of the selected pages; partials meet in rank 0's shared memory
//
// The only global synchronisation is one arrival counter per head group (estimate done); bounded spins; the last
// CTA to finish zeroes the sync words for the next launch.
//
// Arithmetic is the stand-alone kernels' (estimate.cu, topk.cuh, decode.cu): scores and the selected set are
// bit-identical to theirs, the attention output differs only by the chunking of the merge.
#include <cooperative_groups.h>
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
constexpr int kFCluster = 8;                // CTAs per head
constexpr uint32_t kFMaxGroups = 2;
constexpr uint32_t kFSpinLimit = 1u << 21;  // x ~64 ns: a lost producer costs ~0.15 s and an error code, not a hang

// sync words inside the workspace's first kWsSyncBytes (u32 index)
enum : uint32_t { kSyncDone = 0, kSyncErr = 1, kSyncEst = 8 /* [kFMaxGroups] */ };

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
  uint32_t scores_given;   // 1: the scores were written by an earlier kernel (estimate.cu); only select + decode here
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
#ifdef QUEST_FUSED_CLOCKS
  if (atomicExch(err, 1u) == 0u) {  // first timeout: who, what was seen, what was expected
    g_fused_clk[1000 * kClkEvents + 0] = blockIdx.x;
    g_fused_clk[1000 * kClkEvents + 1] = ld_acquire_u32(p);
    g_fused_clk[1000 * kClkEvents + 2] = target;
    g_fused_clk[1000 * kClkEvents + 3] = gtime();
  }
#else
  atomicExch(err, 1u);
#endif
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
  float o[2 * kFWarps][128];   // decode: lane-group states
  float md[2 * kFWarps][2];
  float part_o[kFCluster][128];  // rank 0: the cluster's partial outputs, written by the peers (DSMEM)
  float part_md[kFCluster][2];
  uint16_t keys[4 * kFThreads];
  TopkFastScratch<kFWarps> tk;
  uint32_t wsum[kFWarps];
  unsigned long long sel;      // the head's selection word, written by rank 0 (DSMEM)
  uint32_t flag;
};

// ---- estimate phases ------------------------------------------------------------------------------------------
// Same arithmetic as estimate_tile_kernel<128, kOrdered, 1> (estimate.cu).  The work list is phase-major: phase g =
// the Hg heads of group g over all pages, cut into passes of 32 (page, head) rows = 32/Hg consecutive pages; the
// passes of a phase are dealt round-robin to the phase's participants, so every CTA works on phase 0 first, then
// phase 1, ...: the groups' scores complete in order across the whole GPU.  A lane keeps 4 passes in flight
// (register ring) and the ring rolls over phase boundaries, so the stream never drains; because one ring can hold
// passes of different groups, q and its sign masks are read per pass from shared memory.
// The first ring is issued before the programmatic-launch wait: the scored metadata entries are not written by
// the preceding kernel (the QUEST_EST_META_STABLE contract, which quest_sparse_decode_step states for its callers).
struct EstCursor {
  uint32_t g, i;  // next pass: phase g, index i within the phase; g > g_last = exhausted
};

template <int kOrdered>
__device__ __forceinline__ void fused_estimate(const FusedArgs& a, const int32_t* idx_s, float* q_s, uint32_t* m_s,
                                               uint32_t b, uint32_t grid, uint32_t my_tk, uint32_t* sync) {
  constexpr int D = 128, LPR = 16, kDepth = 4;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lir = lane % LPR, rsub = lane / LPR;
  const uint32_t rt = warp * 2 + rsub;                  // this lane's row in a pass (0..31)
  const uint32_t pg_in = rt >> a.hg_shift, hin = rt & (a.Hg - 1);
  const uint32_t PPP = 32u >> a.hg_shift;               // pages per pass
  const uint32_t ppg = (a.len + PPP - 1) / PPP;         // passes per phase
  const uint32_t row_bytes = a.H * D * 2, slab_bytes = a.P * row_bytes;
  const uint8_t* base = reinterpret_cast<const uint8_t*>(a.meta) + (size_t)hin * (D * 2) + lir * 16;
  const uint32_t g_last = my_tk < a.n_groups ? my_tk : a.n_groups - 1;  // last phase this CTA takes part in

  // phase g is shared by the CTAs that are not selecting for an earlier group (rank 0 of the heads below g*Hg sit
  // out); the deal continues where the previous phase stopped so that the odd pass does not always land on the same CTAs
  auto parts = [&](uint32_t g) { return grid - g * a.Hg; };
  auto first_pass = [&](uint32_t g) {
    const uint32_t n = parts(g), r = b - min(g * a.Hg, (b + kFCluster - 1) / kFCluster);
    return (r + n - (g * ppg) % n) % n;
  };
  auto settle = [&](EstCursor& c) {  // move to the next phase that still has a pass for this CTA
    while (c.g <= g_last && c.i >= ppg) {
      ++c.g;
      if (c.g <= g_last) c.i = first_pass(c.g);
    }
  };

  EstCursor ld = {0u, first_pass(0)};
  settle(ld);
  uint32_t tag[kDepth];  // pass held by each stage: g << 24 | i, or ~0 when empty
  H8 mx[kDepth], mn[kDepth];
#pragma unroll
  for (int s = 0; s < kDepth; ++s)
#pragma unroll
    for (int i = 0; i < 4; ++i) mx[s].h[i] = mn[s].h[i] = __half2half2(__ushort_as_half(0));
  auto issue = [&](int s) {
    const bool live = ld.g <= g_last;
    const uint32_t pq = ld.i * PPP + pg_in;
    const bool ok = live && pq < a.len;
    const uint32_t pc = min(pq, a.len - 1);
    const uint32_t mp = a.p_magic ? __umulhi(pc, a.p_magic) : pc;
    const uint32_t e = pc - mp * a.P;
    const uint8_t* ad = base + (size_t)(live ? ld.g : 0u) * a.Hg * (D * 2) +
                        (uint64_t)(uint32_t)idx_s[mp] * (2u * slab_bytes) + e * row_bytes;
    ld_stream_if(mx[s], ad, ok);
    ld_stream_if(mn[s], ad + slab_bytes, ok);
    tag[s] = live ? (ld.g << 24) | ld.i : 0xffffffffu;
    if (live) {
      ld.i += parts(ld.g);
      settle(ld);
    }
  };
#pragma unroll
  for (int s = 0; s < kDepth; ++s) issue(s);
  pdl_wait();  // q, the sync words and the selected pages' contents belong to earlier work on the stream
  F_STAMP(1);
  for (uint32_t i = tid; i < a.H * D / 2; i += kFThreads) {
    const __half2 h2 = reinterpret_cast<const __half2*>(a.q)[i];
    const float2 f = __half22float2(h2);
    q_s[2 * i] = f.x;
    q_s[2 * i + 1] = f.y;
    const uint32_t bits = *reinterpret_cast<const uint32_t*>(&h2);
    // sign-bit test: for q = -0 either choice gives a +-0 product, which cannot change the sum
    m_s[i] = ((bits & 0x8000u) ? 0u : 0xffffu) | ((bits & 0x80000000u) ? 0u : 0xffff0000u);
  }
  __syncthreads();

  uint32_t signed_upto = 0;  // phases [0, signed_upto) are signed by this CTA
  while (tag[0] != 0xffffffffu) {  // uniform over the CTA
    float sa[kDepth];
    uint32_t ctag[kDepth];
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      ctag[s] = tag[s];
      const uint32_t g = tag[s] == 0xffffffffu ? 0u : tag[s] >> 24;
      const uint32_t h = g * a.Hg + hin;
      const float4* qp = reinterpret_cast<const float4*>(q_s + h * D + lir * 8);
      const float4 q0 = qp[0], q1 = qp[1];
      const uint4 msk = *reinterpret_cast<const uint4*>(m_s + h * (D / 2) + lir * 4);
      const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      const uint32_t mk[4] = {msk.x, msk.y, msk.z, msk.w};
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
      issue(s);  // refill this stage with the CTA's next pass (possibly of the next phase)
    }
    const float v = f_reduce4_transposed<LPR>(sa, lir);
    // after the transposed reduce this lane holds the row of stage lir / 4
    const uint32_t js = lir / (LPR / 4);
    const uint32_t t = js == 0 ? ctag[0] : js == 1 ? ctag[1] : js == 2 ? ctag[2] : ctag[3];
    if ((lir % (LPR / 4)) == 0 && t != 0xffffffffu) {
      const uint32_t pp = (t & 0xffffffu) * PPP + pg_in;
      if (pp < a.len) a.scores[(size_t)((t >> 24) * a.Hg + hin) * a.len + pp] = __float2half_rn(v);
    }
    // sign the phases this CTA has no pass left in: everything below the phase of the oldest pass in the ring
    const uint32_t next_g = tag[0] == 0xffffffffu ? g_last + 1 : tag[0] >> 24;
    if (next_g > signed_upto) {  // uniform
      __syncthreads();
      if (tid == 0)
        for (uint32_t g = signed_upto; g < next_g; ++g) red_release_add(sync + kSyncEst + g, 1u);
      F_STAMP(2 + min(next_g - 1, 1u));
      signed_upto = next_g;
    }
  }
  if (signed_upto <= g_last) {  // phases in which this CTA had no pass at all
    __syncthreads();
    if (tid == 0)
      for (uint32_t g = signed_upto; g <= g_last; ++g) red_release_add(sync + kSyncEst + g, 1u);
  }
}

// ---- cluster plumbing ----------------------------------------------------------------------------------------
namespace cg = cooperative_groups;

__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire;" ::: "memory"); }

// ---- selection of one head by the whole CTA (rank 0 of the head's cluster) -------------------------------------
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

// Called by every thread of the selecting CTA as soon as the selection is decided: the thread that knows the tie
// cut stores the word into all 8 CTAs' shared memory, then everyone arrives at the cluster barrier (release), so
// the peers see it one barrier latency later while this CTA goes on to write the id list.
struct PublishToCluster {
  FusedSmem* sm;
  __device__ __forceinline__ void operator()(bool holder, unsigned long long word) const {
    if (holder) {
      cg::cluster_group cl = cg::this_cluster();
#pragma unroll
      for (int r = 0; r < kFCluster; ++r) *cl.map_shared_rank(&sm->sel, r) = word;
    }
    cluster_arrive();
    F_STAMP(11);  // selection published
  }
};

__device__ __forceinline__ void fused_topk(const FusedArgs& a, FusedSmem* sm, uint32_t h) {
  fused_stage_keys(a, sm, h);
  __syncthreads();
  F_STAMP(4);  // keys staged
  topk_select_fast<kFWarps>(sm->keys, a.len, a.k, threadIdx.x, 1, &sm->tk, a.kv_indices,
                            a.topk_idx + (size_t)h * a.k, (uint16_t*)nullptr, PublishToCluster{sm});
  F_STAMP(9);  // ids written (this thread)
}

// ---- decode of one eighth of a head's pages by one CTA ---------------------------------------------------------
// The CTA's share is the rank range [pg0, pg1) of the head's selected pages in ascending position (+ the forced
// last page as rank k).  Instead of waiting for an id list, every CTA ranks the head's 2K scores itself against the
// selection word (key > tau, or key == tau below the tie cut): scores and page ids are in registers before the word
// arrives, and what is left afterwards is one block scan.  The share's token rows are then dealt to the 16 warps as
// one flat list (a warp may finish one page and start the next), a half-warp owns a token row, 4-stage register ring
// of (K row, V row) pairs, two tokens per softmax rescale (decode.cu's arithmetic).
struct DecodePrefetch {
  uint32_t key[4];
  int32_t pid[4];
};

__device__ __forceinline__ void fused_decode_prefetch(const FusedArgs& a, uint32_t h, DecodePrefetch& pf) {
  const uint32_t i0 = threadIdx.x * 4;
  const uint16_t* s = reinterpret_cast<const uint16_t*>(a.scores) + (size_t)h * a.len;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool in = i0 + j < a.len;
    pf.key[j] = in ? (uint32_t)key_of(__ldcg(s + i0 + j)) : 0u;
    pf.pid[j] = in ? __ldg(a.kv_indices + i0 + j) : 0;
  }
}

// Returns false when the selection word is missing (the selecting CTA timed out).
__device__ __forceinline__ bool fused_decode(const FusedArgs& a, FusedSmem* sm, int32_t* s_pages, uint32_t h,
                                             uint32_t chunk, const DecodePrefetch& pf) {
  constexpr int D = 128, LPR = 16, TPL = 2, NG = kFWarps * TPL, kRing = 4;
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t lir = lane % LPR, sub = lane / LPR, gid = warp * TPL + sub;
  const uint32_t total_pages = a.k + 1;
  const uint32_t pg0 = (uint32_t)(((uint64_t)chunk * total_pages) / kFCluster);
  const uint32_t pg1 = (uint32_t)(((uint64_t)(chunk + 1) * total_pages) / kFCluster);
  const uint32_t npg = pg1 - pg0;
  const uint32_t i0 = tid * 4;

  const unsigned long long word = sm->sel;
  F_STAMP(12);  // selection word seen
  if (!word) return false;
  const uint32_t tau = (uint32_t)(word & 0xffffu), cut = (uint32_t)((word >> 16) & 0x7fffffffu);
  uint32_t selm = 0, cnt = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool in = i0 + j < a.len;
    const bool s = in && (pf.key[j] > tau || (pf.key[j] == tau && i0 + j < cut));
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
      if (rank >= pg0 && rank < pg1) s_pages[rank - pg0] = pf.pid[j];
      ++rank;
    }
  }
  if (tid == 0 && pg1 == total_pages) s_pages[a.k - pg0] = (int32_t)a.last_page_idx;
  float qv[8];
  {
    const H8 qh = ld_plain(a.q + (size_t)h * D + lir * 8);
    h8_to_float(qh, qv);
#pragma unroll
    for (int i = 0; i < 8; ++i) qv[i] *= a.sm_scale_log2;
  }
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
  const uint32_t n_tl = (a.P + TPL - 1) / TPL;                 // token-loads (stages) per page
  const uint32_t n_sel_local = a.k > pg0 ? a.k - pg0 : 0;      // local pages below this are selected (full) pages
  // this warp's slice of the share's npg * n_tl stages
  const uint32_t st_all = npg * n_tl;
  const uint32_t st0 = (warp * st_all) / kFWarps, st1 = ((warp + 1) * st_all) / kFWarps;
  uint32_t pg_ld = st0 / n_tl, i_ld = st0 - pg_ld * n_tl, rem_ld = st1 - st0, okm = 0, nv_ld = 0;
  const uint8_t* ptr = kvb;
  auto open_page = [&]() {
    if (pg_ld < npg) {
      ptr = kvb + (uint64_t)(uint32_t)s_pages[pg_ld] * page_bytes + (uint64_t)i_ld * (TPL * row_bytes);
      nv_ld = pg_ld < n_sel_local ? a.P : a.last_page_len;
    } else {
      nv_ld = 0;
    }
  };
  open_page();
  auto issue = [&](int st) {
    const bool ok = rem_ld != 0 && i_ld * TPL + sub < nv_ld;
    ld_stream_if(kk[st], ptr, ok);
    ld_stream_if(vv[st], ptr + slab_bytes, ok);
    okm = (okm & ~(1u << st)) | ((uint32_t)ok << st);
    if (rem_ld) {
      --rem_ld;
      ptr += TPL * row_bytes;
      if (++i_ld == n_tl) {  // warp-uniform
        i_ld = 0;
        ++pg_ld;
        open_page();
      }
    }
  };
  auto consume2 = [&](int st) {  // stages st, st+1: two token rows per lane group, one softmax rescale
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
  const uint32_t my_tl = st1 - st0;
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
  F_STAMP(13);  // this thread's rows are consumed

  // CTA merge of the NG lane-group states, then the partial goes to rank 0's shared memory
#pragma unroll
  for (int i = 0; i < 8; ++i) sm->o[gid][lir * 8 + i] = acc[i];
  if (lir == 0) {
    sm->md[gid][0] = m;
    sm->md[gid][1] = d;
  }
  __syncthreads();
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
    cg::cluster_group cl = cg::this_cluster();
    FusedSmem* lead = cl.map_shared_rank(sm, 0);
    lead->part_o[chunk][tid] = osum;
    if (tid == 0) {
      lead->part_md[chunk][0] = M;
      lead->part_md[chunk][1] = dsum;
    }
  }
  return true;
}

// rank 0: merge the cluster's 8 partials (state.cuh:53-62 / cascade.cuh:279-340 in the reference) and write o
__device__ __forceinline__ void fused_merge(const FusedArgs& a, FusedSmem* sm, uint32_t h) {
  const uint32_t tid = threadIdx.x;
  if (tid < 128) {
    float M = kFNegBig, dsum = 0.f, osum = 0.f;
#pragma unroll
    for (int c = 0; c < kFCluster; ++c) M = fmaxf(M, sm->part_md[c][0]);
#pragma unroll
    for (int c = 0; c < kFCluster; ++c) {
      const float w = ex2(sm->part_md[c][0] - M);
      dsum += sm->part_md[c][1] * w;
      osum += sm->part_o[c][tid] * w;
    }
    a.o[(size_t)h * 128 + tid] = __float2half_rn(__fdividef(osum, dsum));
  }
}

// ---- the kernel -------------------------------------------------------------------------------------------------
template <int kOrdered>
__global__ void __cluster_dims__(kFCluster, 1, 1) __launch_bounds__(kFThreads, 2) fused_step_kernel(const FusedArgs a) {
  extern __shared__ __align__(16) uint8_t f_dyn[];
  __shared__ FusedSmem sm;
  const uint32_t tid = threadIdx.x, b = blockIdx.x, grid = gridDim.x;
  const uint32_t n_meta = (a.len + a.P - 1) / a.P;
  float* q_s = reinterpret_cast<float*>(f_dyn);                       // [H][128] q as fp32
  uint32_t* m_s = reinterpret_cast<uint32_t*>(q_s + a.H * 128);       // [H][64] sign masks of q (packed halves)
  int32_t* idx_s = reinterpret_cast<int32_t*>(m_s + a.H * 64);        // [n_meta] metadata page ids
  int32_t* s_pages = idx_s + n_meta;                                  // [pages of this CTA's share] selected page ids
  uint32_t* sync = reinterpret_cast<uint32_t*>(a.ws);

  // cluster h = head h; rank 0 also runs the head's selection and sits out the estimate phases after its group
  const uint32_t h = b / kFCluster, r = b - h * kFCluster, g = h >> a.hg_shift;
  const uint32_t my_tk = r == 0 ? g : 0xffffffffu;
  const uint32_t est_parts = grid - g * a.Hg;  // participants of estimate phase g

  F_STAMP(0);
  bool scores_ok = true;
  if (a.scores_given) {
    pdl_wait();  // scores, q and the new token's row come from earlier kernels on the stream
  } else {
    for (uint32_t i = tid; i < n_meta; i += kFThreads) idx_s[i] = __ldg(a.meta_indices + i);
    __syncthreads();
    fused_estimate<kOrdered>(a, idx_s, q_s, m_s, b, grid, my_tk, sync);  // contains the programmatic-launch wait
    // the head's scores are complete once every participant of phase g has signed
    if (tid == 0) sm.flag = spin_until_ge(sync + kSyncEst + g, est_parts, sync + kSyncErr);
    __syncthreads();
    scores_ok = sm.flag;
  }
  F_STAMP(10);  // all scores of the group are in
  // No early programmatic-launch trigger in the all-in-one mode: a later grid's clusters placed into the slots of
  // retiring CTAs fragment the GPCs, and a cluster of this grid that is not resident yet may then never fit --
  // a deadlock until the spins time out (seen on B200: 1 of 32 clusters missing, tools/fused_stress.py).
  DecodePrefetch pf;
  if (r == 0) {
    if (scores_ok) {
      fused_topk(a, &sm, h);  // every thread arrives at the cluster barrier inside (PublishToCluster)
    } else {
      if (tid == 0) {
        cg::cluster_group cl = cg::this_cluster();
        for (int p = 0; p < kFCluster; ++p) *cl.map_shared_rank(&sm.sel, p) = 0ull;
      }
      cluster_arrive();
    }
    // this CTA's keys are already staged in shared memory (position order)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t i = tid * 4 + j;
      pf.key[j] = i < a.len ? (uint32_t)sm.keys[i] : 0u;
      pf.pid[j] = i < a.len ? __ldg(a.kv_indices + i) : 0;
    }
  } else {
    if (scores_ok) fused_decode_prefetch(a, h, pf);
    cluster_arrive();
  }
  cluster_wait();  // the selection word is in sm.sel

  const bool ok = scores_ok && fused_decode(a, &sm, s_pages, h, r, pf);
  if (!ok && tid < 128) {  // keep the merge well-defined on the error path
    cg::cluster_group cl = cg::this_cluster();
    FusedSmem* lead = cl.map_shared_rank(&sm, 0);
    lead->part_o[r][tid] = 0.f;
    if (tid == 0) lead->part_md[r][0] = kFNegBig, lead->part_md[r][1] = 0.f;
  }
  cluster_arrive();
  cluster_wait();  // all 8 partials are in rank 0's shared memory
  if (r == 0) fused_merge(a, &sm, h);
  F_STAMP(14);

  if (a.scores_given) return;  // no global sync words were used
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

// 1 when H clusters of 8 CTAs can be co-resident (they wait on one another), else 0
static int fused_clusters_fit(uint32_t H, size_t smem) {
  static int max_clusters = -1;
  if (max_clusters < 0) {
    // dynamic shared memory is capped at 40 KB by the launcher; static + dynamic may pass the 48 KB default
    cudaFuncSetAttribute(fused_step_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
    cudaFuncSetAttribute(fused_step_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kFCluster * 32);
    cfg.blockDim = dim3(kFThreads);
    cfg.dynamicSmemBytes = 40 * 1024;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, fused_step_kernel<1>, &cfg) != cudaSuccess) n = 0;
    cudaGetLastError();
    max_clusters = n;
  }
  (void)smem;
  return (uint32_t)max_clusters >= H;
}

// QUEST_OK when launched; QUEST_ERR_UNSUPPORTED when the shape is outside the fused kernel's domain (the caller
// then uses the three stand-alone kernels).
int launch_fused_step(const void* q, void* o, const void* kv_data, const int32_t* kv_indices, uint32_t len,
                      uint32_t last_page_idx, uint32_t last_page_len, const void* meta_data,
                      const int32_t* meta_indices, void* scores, int32_t* topk_idx, uint32_t k, uint32_t Hq,
                      uint32_t Hkv, uint32_t D, uint32_t P, void* workspace, size_t workspace_bytes, uint32_t flags,
                      cudaStream_t stream) {
  // QUEST_FUSED: 0 (default) = three stand-alone kernels; 1 = estimate kernel + this kernel for select/decode/merge
  // (clusters never wait on one another: safe next to other work on the GPU); 2 = everything in this kernel (needs all
  // clusters resident together: exclusive use of the GPU, no programmatic early launch).
  // Measured on B200 at 32K/2048 (profiles/r01_fused_modes.txt): 21.5 / 23.2 / 26.5 us per layer -- the stand-alone
  // chain wins, so the cluster kernel stays an experiment behind this switch.
  static int enabled = -1, n_groups_env = 0;
  if (enabled < 0) {
    const char* e = getenv("QUEST_FUSED");
    enabled = e ? atoi(e) : 0;
    if (const char* g = getenv("QUEST_FUSED_GROUPS")) n_groups_env = atoi(g);
  }
  if (!enabled) return QUEST_ERR_UNSUPPORTED;
  if (Hq != Hkv || D != 128 || len == 0 || len > 4u * kFThreads || k == 0 || k > len) return QUEST_ERR_UNSUPPORTED;
  if (Hq > 32) return QUEST_ERR_UNSUPPORTED;
  // two head groups when the heads split that way into 32-row tiles, else one
  uint32_t n_groups = n_groups_env > 0 ? (uint32_t)n_groups_env : 2;
  if (n_groups > kFMaxGroups) n_groups = kFMaxGroups;
  while (n_groups > 1 && (Hq % n_groups || 32 % (Hq / n_groups))) --n_groups;
  const uint32_t Hg = Hq / n_groups;
  if (32 % Hg) return QUEST_ERR_UNSUPPORTED;
  if (!workspace || workspace_bytes < kWsSyncBytes) return QUEST_ERR_UNSUPPORTED;
  const uint32_t n_meta = (len + P - 1) / P;
  const uint32_t max_ppc = (k + 1 + kFCluster - 1) / kFCluster + 1;
  const size_t smem = (size_t)Hq * 768 + ((size_t)n_meta + max_ppc) * sizeof(int32_t);
  if (smem > 40 * 1024 || !fused_clusters_fit(Hq, smem)) return QUEST_ERR_UNSUPPORTED;

  FusedArgs a;
  a.q = (const __half*)q, a.o = (__half*)o, a.kv = (const __half*)kv_data, a.kv_indices = kv_indices;
  a.meta = (const __half*)meta_data, a.meta_indices = meta_indices, a.scores = (__half*)scores, a.topk_idx = topk_idx;
  a.ws = (uint8_t*)workspace;
  a.len = len, a.k = k, a.last_page_idx = last_page_idx, a.last_page_len = last_page_len;
  a.H = Hq, a.P = P, a.p_magic = P <= 1 ? 0u : (uint32_t)((0x100000000ull + P - 1) / P);
  a.Hg = Hg, a.hg_shift = 31 - __builtin_clz(Hg), a.n_groups = n_groups;
  a.scores_given = enabled == 1;
  if (a.scores_given) {
    const int rc = quest_estimate(q, scores, len, meta_data, meta_indices, Hq, Hkv, D, P, QUEST_EST_META_STABLE | flags, stream);
    if (rc) return rc;
  }
  a.sm_scale_log2 = kLog2e / sqrtf((float)D);

  // The CTAs wait on one another, so all of them must be resident together: fused_clusters_fit() checked that the
  // Hq clusters fit, the kernels before this one on the stream never wait on it, and every spin is bounded.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(Hq * kFCluster);
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = a.scores_given ? pdl_enabled() : 0;
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
