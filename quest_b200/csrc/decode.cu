// Sparse paged decode attention over per-head selected pages, split-KV with an in-kernel LSE combine.
// Replaces BatchDecodeWithPagedKVCacheKernel<partition_kv=true> + VariableLengthMergeStates + the
// BatchDecodeHandler planner (reference kernels/include/decode/decode_attn.cuh:440-646, 753-893, 962-1002;
// decode_handler.cuh:73-121; flashinfer cascade.cuh:279-340, state.cuh:53-62).
//
// Unit of work = (q head, chunk of pages).  Each group of D/8 lanes owns whole 16-byte-per-lane token
// rows; every warp keeps two items (8 independent 128-bit loads per lane) in flight while it consumes the
// previous one.  Partials go to a caller-owned workspace; the last CTA to arrive for a head (atomic
// ticket) merges them -- no separate merge launch.
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace quest {

// -DQUEST_DEC_CLOCKS: every CTA stamps %globaltimer at its phase boundaries (tools/decode_timeline.py)
#ifdef QUEST_DEC_CLOCKS
__device__ unsigned long long g_dec_clk[4096 * 8];
#define D_STAMP(i)                                                                             \
  do {                                                                                         \
    if (threadIdx.x == 0) {                                                                    \
      unsigned long long t_;                                                                   \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                   \
      g_dec_clk[((blockIdx.y * gridDim.x + blockIdx.x) & 4095) * 8 + (i)] = t_;                \
    }                                                                                          \
  } while (0)
#else
#define D_STAMP(i) do { } while (0)
#endif

#ifndef QUEST_DEC_MINBLOCKS
#define QUEST_DEC_MINBLOCKS 7  // resident CTAs per SM the kernel is compiled for (72 registers); 8 = 64 registers, A/B in profiles/
#endif
#ifndef QUEST_DEC_THREADS
#define QUEST_DEC_THREADS 128  // threads per decode CTA; A/B builds: -DQUEST_DEC_THREADS=256 -DQUEST_DEC_MINBLOCKS=3
#endif
constexpr int kDecThreads = QUEST_DEC_THREADS;
constexpr int kDecWarps = kDecThreads / 32;
constexpr float kNegBig = -1e30f;

// workspace layout: common.cuh (ws_ticket_off / ws_md_off / ws_o_off)

// t / P for any page size: p_magic = ceil(2^32 / P) (0 for P == 1); exact while t < 2^32 / P.
__host__ __device__ inline uint32_t div_magic(uint32_t P) { return P <= 1 ? 0u : (uint32_t)((0x100000000ull + P - 1) / P); }
__device__ __forceinline__ uint32_t div_p(uint32_t t, uint32_t p_magic) { return p_magic ? __umulhi(t, p_magic) : t; }

// One CTA = (q head, chunk of pages).  Each of the 4 warps walks its share of the chunk's tokens in items of
// 2 warp-wide K loads + 2 V loads (a lane group of D/8 lanes owns one 16-byte-per-lane token row), with
// the next item's loads issued before the current item is consumed: 8 independent 128-bit loads per lane
// stay in flight, ~128 KB per SM at 8 CTAs/SM, and the softmax arithmetic hides under them.
// kRowCompact (the sequence-sharded step's attend phase, where most list entries are -1 = pages another shard holds): every CTA of a
// head scans the head's WHOLE list, counts the entries with rows and takes an equal share of them -- entry j of the squeezed
// list goes to chunk j / ceil(n / chunks) -- instead of the entries at its positions in the sparse list, whose number of
// rows varies from chunk to chunk (Poisson around total / shards per chunk) with the slowest chunk setting the head's time.
template <int D, int kThreads, int kMinBlocks, bool kRowCompact = false>
__global__ void __launch_bounds__(kThreads, kMinBlocks)
decode_kernel(const __half* __restrict__ q_, __half* __restrict__ o_, float* __restrict__ lse_,
              const __half* __restrict__ kv_, const int32_t* __restrict__ indices_, uint32_t idx_stride,
              uint32_t n_sel, uint32_t last_page_idx_, uint32_t last_page_len_, uint32_t Hkv, uint32_t G,
              uint32_t P, uint32_t p_magic, uint32_t ppc, float sm_scale_log2, uint8_t* __restrict__ ws_,
              const BatchArgs ba, float* __restrict__ o32, const uint32_t* __restrict__ ho_epoch, unsigned int* __restrict__ err_word) {
  const __half* __restrict__ q = q_;
  __half* __restrict__ o = o_;
  float* __restrict__ lse = lse_;
  const __half* __restrict__ kv = kv_;
  const int32_t* __restrict__ indices = indices_;
  uint32_t last_page_idx = last_page_idx_, last_page_len = last_page_len_;
  uint8_t* __restrict__ ws = ws_;
  if (ba.descs != nullptr) {  // descriptor mode: blockIdx.z = sequence (its record is constant during the step)
    const SeqDesc* d = ba.descs + blockIdx.z;
    const size_t b = blockIdx.z, Hq_ = gridDim.y;
    q += b * Hq_ * D;
    o += b * Hq_ * D;
    if (lse) lse += b * Hq_;
    kv = reinterpret_cast<const __half*>(d->kv_data + ba.kv_layer_off);
    indices += b * Hq_ * idx_stride;
    last_page_len = d->kv_last_page_len;
    last_page_idx = (uint32_t)d->kv_indices[d->num_kv_pages - 1];
    ws += b * ba.ws_stride;
  }
  constexpr int kWarps = kThreads / 32;
  constexpr int LPR = D / 8;
  constexpr int TPL = 32 / LPR;           // token rows per warp-wide load
  constexpr int IT = 2 * TPL;             // tokens per item
  constexpr int NG = kWarps * TPL;     // lane groups per CTA
  extern __shared__ int32_t s_pages[];    // [pages of this chunk]
  __shared__ float s_o[NG][D];
  __shared__ float s_md[NG][2];
  __shared__ uint32_t s_last;
  const uint32_t chunk = blockIdx.x, nchunks = gridDim.x, h = blockIdx.y, Hq = gridDim.y;
  const uint32_t kvh = h / G;
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t lir = lane % LPR, sub = lane / LPR, gid = warp * TPL + sub;
  const uint32_t total_pages = n_sel + 1;  // + forced last page
  const uint32_t pg0 = chunk * ppc, pg1 = min(pg0 + ppc, total_pages);
  const uint32_t npg = pg1 - pg0, ntok = npg * P;
  const uint32_t n_items = (ntok + IT - 1) / IT;
  const size_t slab = (size_t)P * Hkv * D;  // K -> V offset inside a page
  const size_t row = (size_t)Hkv * D;

  D_STAMP(0);
  const int32_t* idx_row = indices + (size_t)h * idx_stride;
  uint32_t qh[4];  // this lane's 8 q dims as packed halves: q.k runs on mixed-precision FMAs (fp16 x fp16 + fp32, exact products)
  uint32_t n_rows = npg, last_pos = 0xffffffffu;  // kRowCompact: this chunk's share of the squeezed list, and where the last page sits in it
  if (kRowCompact) {
    __shared__ uint32_t s_scan[kWarps + 1];
    pdl_wait();
    D_STAMP(1);
    const uint32_t per = (total_pages + kThreads - 1) / kThreads;  // consecutive entries per thread
    const uint32_t i0 = min(tid * per, total_pages), i1 = min(i0 + per, total_pages);
    auto entry = [&](uint32_t i) -> int32_t {
      return i < n_sel ? (int32_t)ld_ordered32(idx_row + i) : (last_page_len ? (int32_t)last_page_idx : -1);
    };
    uint32_t cnt = 0;
    for (uint32_t i = i0; i < i1; ++i) cnt += entry(i) >= 0 ? 1u : 0u;
    uint32_t incl = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t a = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= (uint32_t)off) incl += a;
    }
    if (lane == 31) s_scan[warp] = incl;
    if (tid == 0) s_scan[kWarps] = 0xffffffffu;
    __syncthreads();
    uint32_t ci = incl - cnt, tot = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      ci += (uint32_t)w < warp ? s_scan[w] : 0u;
      tot += s_scan[w];
    }
    const uint32_t share = (tot + nchunks - 1) / nchunks;
    const uint32_t lo = min(chunk * share, tot), hi = min(lo + share, tot);
    for (uint32_t i = i0; i < i1; ++i) {  // second pass: the list is in L1 / L2 now
      const int32_t pid = entry(i);
      if (pid >= 0) {
        if (ci >= lo && ci < hi) {
          s_pages[ci - lo] = pid;
          if (i == n_sel) s_scan[kWarps] = ci - lo;  // the forced last page: last_page_len tokens
        }
        ++ci;
      }
    }
    const uint4 raw = ld_ordered16(q + (size_t)h * D + lir * 8);
    qh[0] = raw.x, qh[1] = raw.y, qh[2] = raw.z, qh[3] = raw.w;
    __syncthreads();
    n_rows = hi - lo;
    last_pos = s_scan[kWarps];
  } else if (ho_epoch == nullptr) {
    pdl_wait();  // q, the page ids (top-k kernel) and the last page's new row (append) come from earlier kernels
    D_STAMP(1);
    for (uint32_t i = tid; i < npg; i += kThreads)
      s_pages[i] = pg0 + i < n_sel ? (int32_t)ld_ordered32(idx_row + pg0 + i) : (int32_t)last_page_idx;
  } else {
    // Page-id hand-over (common.cuh, Handoff): no grid dependency.  This launch was released by the top-k kernel once that had
    // waited for the estimate launch, so q, the appended K/V row and the step counter are complete and visible; the chunk's
    // page ids arrive as tagged 8-byte entries in this CTA's own partial-output slot, which it overwrites at its end.
    D_STAMP(1);
    {
      const uint4 raw = ld_ordered16(q + (size_t)h * D + lir * 8);  // in flight while the ids are awaited
      qh[0] = raw.x, qh[1] = raw.y, qh[2] = raw.z, qh[3] = raw.w;
    }
    const uint32_t tag = handoff_tag(ld_ordered32(ho_epoch));
    const uint8_t* slot = ws + ws_o_off(Hq, nchunks) + (size_t)(h * nchunks + chunk) * (D * sizeof(float));
    for (uint32_t i = tid; i < npg; i += kThreads) {
      int32_t pid = (int32_t)last_page_idx;
      if (pg0 + i < n_sel) {
        unsigned long long v = ld_relaxed_u64(slot + i * 8u);
        for (uint32_t spins = 0; (uint32_t)(v >> 32) != tag; ++spins) {
          if (spins > (1u << 22)) {  // seconds: the top-k launch never delivered -- fail loudly, attend to nothing for this id
            device_error_raise(err_word, QUEST_DEVERR_STEP);
            v = 0xffffffffull;  // id -1
            break;
          }
          __nanosleep(20);
          v = ld_relaxed_u64(slot + i * 8u);
        }
        pid = (int32_t)(uint32_t)v;
      }
      s_pages[i] = pid;
    }
  }
  if (!kRowCompact && ho_epoch == nullptr) {
    const uint4 raw = ld_ordered16(q + (size_t)h * D + lir * 8);
    qh[0] = raw.x, qh[1] = raw.y, qh[2] = raw.z, qh[3] = raw.w;
  }
  __syncthreads();
  D_STAMP(2);  // page ids and q are in
#if QUEST_TRIG_DEC
  pdl_trigger();
#endif

  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  float m = kNegBig, d = 0.f;

  // Each warp owns whole pages of the chunk (local pages warp, warp+4, ...) and walks their token rows with a
  // 4-deep register ring of (K row, V row) pairs: a stage is one warp-wide load of TPL token rows, so addresses
  // advance by a constant inside a page and are recomputed (one shared-memory lookup) only at page boundaries.
  // Invalid rows (past the last page's length, past the chunk) are not loaded (predicated) and enter the
  // softmax with weight 0; their stale registers always hold finite halves.
  constexpr int kRing = 4;
  H8 kk[kRing], vv[kRing];
#pragma unroll
  for (int st = 0; st < kRing; ++st)
#pragma unroll
    for (int i = 0; i < 4; ++i) kk[st].h[i] = vv[st].h[i] = __half2half2(__ushort_as_half(0));
  const uint32_t row_bytes = Hkv * D * 2, slab_bytes = P * row_bytes;
  const uint64_t page_bytes = 2ull * slab_bytes;
  const uint8_t* kvb = reinterpret_cast<const uint8_t*>(kv) + (size_t)kvh * D * 2 + lir * 16 + (size_t)sub * row_bytes;
  const uint32_t n_tl = (P + TPL - 1) / TPL;                   // token-loads per page
  const uint32_t n_sel_local = n_sel > pg0 ? n_sel - pg0 : 0;  // local pages below this are selected (full) pages
  // issue-side cursor
  uint32_t pg_ld = warp, i_ld = 0, okm = 0;
  const uint8_t* ptr = kvb;
  uint32_t nv_ld = 0;  // valid tokens of the page under the cursor
  // A list entry with nothing to read -- id < 0: a selected page another sequence shard holds; the forced last page on a shard
  // that does not hold it (last_page_len 0) -- is stepped over, not walked with its loads switched off: on one of 8 shards 7 of
  // 8 entries are such pages, and walking them cost the softmax arithmetic of a full list (18.6 -> 12.6 us per layer there;
  // squeezing the chunk's list in shared memory first, so that its warps share the entries evenly, measured no better: 14.4 us --
  // what is left is the spread of the entries with rows over the chunks, profiles/r02_b_experiments.md).
  const uint32_t n_list = kRowCompact ? n_rows : npg;  // entries of s_pages this CTA walks
  auto has_rows = [&](uint32_t pg) { return kRowCompact || (s_pages[pg] >= 0 && (pg < n_sel_local || last_page_len > 0)); };
  auto open_page = [&]() {
    while (pg_ld < n_list && !has_rows(pg_ld)) pg_ld += kWarps;
    if (pg_ld < n_list) {
      ptr = kvb + (uint64_t)(uint32_t)s_pages[pg_ld] * page_bytes;
      nv_ld = kRowCompact ? (pg_ld == last_pos ? last_page_len : P) : (pg_ld < n_sel_local ? P : last_page_len);
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
    if (++i_ld == n_tl) {  // warp-uniform
      i_ld = 0;
      pg_ld += kWarps;
      open_page();
    }
  };
  auto consume2 = [&](int st) {  // stages st, st+1: two token rows per lane group, one softmax rescale
    float sc[2];
    float mt = m;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float a = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) a = fma_h2_seq(qh[i], *reinterpret_cast<const uint32_t*>(&kk[st + j].h[i]), a);
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
      sc[j] = (okm >> (st + j)) & 1u ? a * sm_scale_log2 : kNegBig;  // softmax scale and log2(e) applied to the fp32 dot product
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
  // token-loads this warp must consume: its pages x n_tl (warp-uniform; the shuffles use the full mask)
  uint32_t my_pages = 0;
  for (uint32_t pg = warp; pg < n_list; pg += kWarps) my_pages += has_rows(pg) ? 1u : 0u;
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

  D_STAMP(3);  // warp 0's rows are consumed
  // CTA merge of the NG group states.
#pragma unroll
  for (int i = 0; i < 8; ++i) s_o[gid][lir * 8 + i] = acc[i];
  if (lir == 0) {
    s_md[gid][0] = m;
    s_md[gid][1] = d;
  }
  __syncthreads();
  constexpr int CPT = (D + kThreads - 1) / kThreads;  // output columns per thread
  float M = kNegBig, dsum = 0.f, osum[CPT];
#pragma unroll
  for (int g = 0; g < NG; ++g) M = fmaxf(M, s_md[g][0]);
#pragma unroll
  for (int c = 0; c < CPT; ++c) osum[c] = 0.f;
#pragma unroll
  for (int g = 0; g < NG; ++g) {
    const float w = ex2(s_md[g][0] - M);
    dsum += s_md[g][1] * w;
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const uint32_t col = tid + c * kThreads;
      if (col < D) osum[c] += s_o[g][col] * w;
    }
  }
  if (nchunks == 1) {
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const uint32_t col = tid + c * kThreads;
      if (col < D) {
        const float val = dsum > 0.f ? __fdividef(osum[c], dsum) : 0.f;
        if (o32 != nullptr) o32[(size_t)h * D + col] = val;  // a shard's partial for the cross-GPU merge: not rounded to fp16
        else o[(size_t)h * D + col] = __float2half_rn(val);
      }
    }
    if (lse && tid == 0) lse[h] = dsum > 0.f ? M + lg2(dsum) : -INFINITY;
    return;
  }
  float* ws_o = reinterpret_cast<float*>(ws + ws_o_off(Hq, nchunks));
  float* ws_md = reinterpret_cast<float*>(ws + ws_md_off(Hq));
  uint32_t* ticket = reinterpret_cast<uint32_t*>(ws + ws_ticket_off());
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    const uint32_t col = tid + c * kThreads;
    if (col < D) ws_o[((size_t)h * nchunks + chunk) * D + col] = osum[c];
  }
  if (tid == 0) {
    ws_md[((size_t)h * nchunks + chunk) * 2] = M;
    ws_md[((size_t)h * nchunks + chunk) * 2 + 1] = dsum;
  }
  // Release/acquire through the ticket: the CTA's partial is ordered before tid 0's acq_rel atomic by the
  // barrier, and the last arriver's acquire orders every other CTA's partial before its loads below.
  __syncthreads();
  if (tid == 0) {
    uint32_t old;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(ticket + h) : "memory");
    s_last = old == nchunks - 1;
  }
  __syncthreads();
  D_STAMP(4);  // partial published, ticket taken
  if (!s_last) return;
  // Last arriver: one round trip.  All partial rows are requested as independent 128-bit loads
  // (kThreads / (D/4) rows per sweep, 8 sweeps in flight) before any weight is needed.
  constexpr int NV = D / 4;               // float4 per partial row
  constexpr int RPS = kThreads / NV;   // rows per sweep
  static_assert(RPS >= 1 && RPS <= NG, "partial-row sweep must fit the group-state buffer");
  const uint32_t cr = tid / NV, c4 = tid % NV;
  const float4* po = reinterpret_cast<const float4*>(ws_o + (size_t)h * nchunks * D);
  const float2* pmd = reinterpret_cast<const float2*>(ws_md + (size_t)h * nchunks * 2);
  float4 a4 = make_float4(0.f, 0.f, 0.f, 0.f);
  float MM = kNegBig, dd = 0.f;
  for (uint32_t cb = 0; cb < nchunks; cb += RPS * 8) {
    float4 v[8];
    float2 md[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t c = cb + i * RPS + cr;
      v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      md[i] = make_float2(kNegBig, 0.f);
      if (c < nchunks) {
        v[i] = __ldcg(po + (size_t)c * NV + c4);
        md[i] = __ldcg(pmd + c);
      }
    }
    float mb = MM;
#pragma unroll
    for (int i = 0; i < 8; ++i) mb = fmaxf(mb, md[i].x);
    const float sc0 = ex2(MM - mb);
    MM = mb;
    dd *= sc0;
    a4.x *= sc0, a4.y *= sc0, a4.z *= sc0, a4.w *= sc0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float w = ex2(md[i].x - MM);
      dd += md[i].y * w;
      a4.x += v[i].x * w, a4.y += v[i].y * w, a4.z += v[i].z * w, a4.w += v[i].w * w;
    }
  }
  // combine the RPS row groups (each saw a disjoint subset of the chunks) through shared memory
  *reinterpret_cast<float4*>(&s_o[cr][c4 * 4]) = a4;
  if (c4 == 0) {
    s_md[cr][0] = MM;
    s_md[cr][1] = dd;
  }
  __syncthreads();
  float Mf = kNegBig;
#pragma unroll
  for (int g = 0; g < RPS; ++g) Mf = fmaxf(Mf, s_md[g][0]);
  float df = 0.f, of[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) of[c] = 0.f;
#pragma unroll
  for (int g = 0; g < RPS; ++g) {
    const float w = ex2(s_md[g][0] - Mf);
    df += s_md[g][1] * w;
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const uint32_t col = tid + c * kThreads;
      if (col < D) of[c] += s_o[g][col] * w;
    }
  }
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    const uint32_t col = tid + c * kThreads;
    if (col < D) {
      const float val = df > 0.f ? __fdividef(of[c], df) : 0.f;
      if (o32 != nullptr) o32[(size_t)h * D + col] = val;
      else o[(size_t)h * D + col] = __float2half_rn(val);
    }
  }
  if (tid == 0) {
    if (lse) lse[h] = df > 0.f ? Mf + lg2(df) : -INFINITY;
    ticket[h] = 0;  // ready for the next launch
  }
  D_STAMP(5);  // head merged
}

template <int D>
static int launch_decode(const void* q, void* o, float* lse, const void* kv, const int32_t* indices,
                         uint32_t idx_stride, uint32_t n_sel, uint32_t last_page_idx, uint32_t last_page_len,
                         uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t ppc, void* ws, cudaStream_t stream,
                         const BatchArgs& ba = BatchArgs{}, uint32_t batch = 1, float* o32 = nullptr,
                         const uint32_t* ho_epoch = nullptr) {
  const uint32_t nchunks = (n_sel + 1 + ppc - 1) / ppc;
  const uint32_t npg_max = ppc < n_sel + 1 ? ppc : n_sel + 1;
  const size_t smem = (size_t)npg_max * sizeof(int32_t);
  if (smem > 32 * 1024) return QUEST_ERR_UNSUPPORTED;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nchunks, Hq, batch);
  cfg.blockDim = dim3(kDecThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const float sm_scale_log2 = kLog2e / sqrtf((float)D);
  // partial (fp32) output = a sequence shard's attend phase: its list is mostly -1 entries
  const bool row_compact = o32 != nullptr && ba.descs == nullptr && ho_epoch == nullptr && nchunks > 1 && debug_option(QUEST_DBG_NO_ROW_COMPACT) == 0;
  auto kern = row_compact ? decode_kernel<D, kDecThreads, QUEST_DEC_MINBLOCKS, true> : decode_kernel<D, kDecThreads, QUEST_DEC_MINBLOCKS, false>;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, (const __half*)q, (__half*)o, lse,
                                     (const __half*)kv, indices, idx_stride, n_sel, last_page_idx, last_page_len, Hkv,
                                     Hq / Hkv, P, div_magic(P), ppc, sm_scale_log2, (uint8_t*)ws, ba, o32, ho_epoch,
                                     ho_epoch ? device_error_word() : nullptr);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

// Layer step with the page-id hand-over (capi.cu): the decode launch takes its page ids from the slots the top-k kernel fills.
// Usable when the chunk's ids fit its partial-output slot and the plan has more than one chunk.
bool decode_handoff_fits(uint32_t n_sel, uint32_t D, uint32_t ppc) { return ppc * 8u <= D * sizeof(float) && n_sel + 1 > ppc; }
int decode_run_handoff(const void* q, void* o, const void* kv, const int32_t* indices, uint32_t n_sel, uint32_t last_page_idx,
                       uint32_t last_page_len, uint32_t Hq, uint32_t Hkv, uint32_t D, uint32_t P, uint32_t ppc, void* ws,
                       cudaStream_t s) {
  const uint32_t* epoch = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(ws) + kWsEpochOff);
  switch (D) {
    case 64: return launch_decode<64>(q, o, nullptr, kv, indices, n_sel, n_sel, last_page_idx, last_page_len, Hq, Hkv, P, ppc, ws, s, BatchArgs{}, 1, nullptr, epoch);
    case 128: return launch_decode<128>(q, o, nullptr, kv, indices, n_sel, n_sel, last_page_idx, last_page_len, Hq, Hkv, P, ppc, ws, s, BatchArgs{}, 1, nullptr, epoch);
    case 256: return launch_decode<256>(q, o, nullptr, kv, indices, n_sel, n_sel, last_page_idx, last_page_len, Hq, Hkv, P, ppc, ws, s, BatchArgs{}, 1, nullptr, epoch);
    default: return QUEST_ERR_UNSUPPORTED;
  }
}

// descriptor mode (quest_decode_layer_step_batch)
int decode_batch(const void* q, void* o, const int32_t* indices, uint32_t n_sel, uint32_t Hq, uint32_t Hkv, uint32_t D,
                 uint32_t P, uint32_t ppc, void* ws, const BatchArgs& ba, uint32_t batch, cudaStream_t s) {
  switch (D) {
    case 64: return launch_decode<64>(q, o, nullptr, nullptr, indices, n_sel, n_sel, 0, 0, Hq, Hkv, P, ppc, ws, s, ba, batch);
    case 128: return launch_decode<128>(q, o, nullptr, nullptr, indices, n_sel, n_sel, 0, 0, Hq, Hkv, P, ppc, ws, s, ba, batch);
    case 256: return launch_decode<256>(q, o, nullptr, nullptr, indices, n_sel, n_sel, 0, 0, Hq, Hkv, P, ppc, ws, s, ba, batch);
    default: return QUEST_ERR_UNSUPPORTED;
  }
}

}  // namespace quest

extern "C" int quest_decode_plan(uint32_t num_sel_pages, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                 uint32_t head_dim, uint32_t page_size, uint32_t num_sms,
                                 uint32_t* pages_per_chunk, uint32_t* num_chunks) {
  if (!pages_per_chunk || !num_chunks || !num_qo_heads || !num_kv_heads || !page_size) return QUEST_ERR_BAD_ARG;
  if (head_dim != 64 && head_dim != 128 && head_dim != 256) return QUEST_ERR_UNSUPPORTED;
  if (num_sms == 0) num_sms = 148;
  const uint32_t total = num_sel_pages + 1;
  const uint32_t slots = num_sms * QUEST_DEC_MINBLOCKS;  // resident 128-thread CTAs per SM
  // Smallest chunk that keeps every CTA resident in one wave, but at least two items per warp.
  const uint32_t item_tokens = 2 * (32 / (head_dim / 8));
  uint32_t min_ppc = (quest::kDecWarps * item_tokens * 2 + page_size - 1) / page_size;
  if (min_ppc == 0) min_ppc = 1;
  uint32_t ppc = min_ppc;
  while (ppc < total && (uint64_t)num_qo_heads * ((total + ppc - 1) / ppc) > slots) ++ppc;
  if (ppc > total) ppc = total;
  *pages_per_chunk = ppc;
  *num_chunks = (total + ppc - 1) / ppc;
  return QUEST_OK;
}

extern "C" size_t quest_decode_workspace_bytes(uint32_t num_qo_heads, uint32_t head_dim, uint32_t num_chunks) {
  return quest::ws_bytes(num_qo_heads, head_dim, num_chunks);
}

static int decode_run_checked(const void* q, void* o, float* o32, float* lse, const void* kv_data, const int32_t* indices,
                              uint32_t idx_row_stride, uint32_t num_sel_pages, uint32_t last_page_idx, uint32_t last_page_len,
                              uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                              uint32_t pages_per_chunk, void* workspace, size_t workspace_bytes, void* stream) {
  if (!q || (!o && !o32) || !kv_data || !num_qo_heads || !num_kv_heads || !page_size || !pages_per_chunk)
    return QUEST_ERR_BAD_ARG;
  if (num_sel_pages && !indices) return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (last_page_len > page_size) return QUEST_ERR_BAD_ARG;  // 0: this shard does not hold the last page
  const uint32_t nchunks = (num_sel_pages + 1 + pages_per_chunk - 1) / pages_per_chunk;
  if (nchunks > 1) {
    if (!workspace || workspace_bytes < quest_decode_workspace_bytes(num_qo_heads, head_dim, nchunks))
      return QUEST_ERR_WORKSPACE;
  }
  cudaStream_t s = (cudaStream_t)stream;
#define QUEST_DEC(DD)                                                                                         \
  return quest::launch_decode<DD>(q, o, lse, kv_data, indices, idx_row_stride, num_sel_pages, last_page_idx, \
                                  last_page_len, num_qo_heads, num_kv_heads, page_size, pages_per_chunk,     \
                                  workspace, s, quest::BatchArgs{}, 1, o32)
  switch (head_dim) {
    case 64: QUEST_DEC(64);
    case 128: QUEST_DEC(128);
    case 256: QUEST_DEC(256);
    default: return QUEST_ERR_UNSUPPORTED;
  }
#undef QUEST_DEC
}

extern "C" int quest_decode_run(const void* q, void* o, float* lse, const void* kv_data, const int32_t* indices,
                                uint32_t idx_row_stride, uint32_t num_sel_pages, uint32_t last_page_idx,
                                uint32_t last_page_len, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                uint32_t head_dim, uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                size_t workspace_bytes, void* stream) {
  if (!o) return QUEST_ERR_BAD_ARG;
  return decode_run_checked(q, o, nullptr, lse, kv_data, indices, idx_row_stride, num_sel_pages, last_page_idx, last_page_len,
                            num_qo_heads, num_kv_heads, head_dim, page_size, pages_per_chunk, workspace, workspace_bytes, stream);
}

extern "C" int quest_decode_run_partial(const void* q, float* o_partial, float* lse, const void* kv_data, const int32_t* indices,
                                        uint32_t idx_row_stride, uint32_t num_sel_pages, uint32_t last_page_idx,
                                        uint32_t last_page_len, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                        uint32_t head_dim, uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  if (!o_partial || !lse) return QUEST_ERR_BAD_ARG;
  return decode_run_checked(q, nullptr, o_partial, lse, kv_data, indices, idx_row_stride, num_sel_pages, last_page_idx,
                            last_page_len, num_qo_heads, num_kv_heads, head_dim, page_size, pages_per_chunk, workspace,
                            workspace_bytes, stream);
}

// ---- quest_decode_run_static: the same launch with its split-KV scratch in a file-scope device array ----------------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): nothing
// here allocates -- no cudaMalloc / cudaFree, no allocating library routine.  Every GPU that loads the library holds the
// array until exit.  In host code the array's NAME is only a host-side stand-in, not a device address, so the address is
// resolved once per device with cudaGetSymbolAddress (that device current) and the kernel is handed that pointer.
namespace quest {
constexpr size_t kStaticScratchBytes = (size_t)4 << 20;  // 4 MiB: e.g. 32 q heads x 128 dims x up to ~230 chunks
static __device__ __align__(256) uint8_t g_decode_static_scratch[kStaticScratchBytes];
constexpr int kStaticMaxDevices = 64;
static void* g_static_scratch_addr[kStaticMaxDevices];  // per device, resolved on first use
static std::mutex g_static_scratch_mu;

static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kStaticMaxDevices) return QUEST_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_decode_static_scratch);  // this device's copy of the array
    if (e != cudaSuccess) return (int)e;
    if (ptr == nullptr) return QUEST_ERR_UNSUPPORTED;
    g_static_scratch_addr[dev] = ptr;
  }
  *out = g_static_scratch_addr[dev];
  return QUEST_OK;
}
}  // namespace quest

extern "C" size_t quest_decode_static_scratch_bytes(void) { return quest::kStaticScratchBytes; }

extern "C" int quest_decode_run_static(const void* q, void* o, float* lse, const void* kv_data, const int32_t* indices,
                                       uint32_t idx_row_stride, uint32_t num_sel_pages, uint32_t last_page_idx,
                                       uint32_t last_page_len, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                       uint32_t head_dim, uint32_t page_size, uint32_t pages_per_chunk, void* stream) {
  if (!o || !num_qo_heads || !pages_per_chunk) return QUEST_ERR_BAD_ARG;
  const uint32_t nchunks = (num_sel_pages + 1 + pages_per_chunk - 1) / pages_per_chunk;
  const size_t need = quest_decode_workspace_bytes(num_qo_heads, head_dim, nchunks);
  if (need > quest::kStaticScratchBytes) return QUEST_ERR_SCRATCH_LIMIT;  // plan with larger chunks, or use quest_decode_run
  void* scratch = nullptr;
  int rc = quest::static_scratch_for_current_device(&scratch);
  if (rc != QUEST_OK) return rc;
  // The array keeps the previous call's contents.  What must start at zero are the reserved words and the merge tickets
  // (one per q head; they reset themselves only when a launch runs to completion): cleared on the stream, ahead of the kernel.
  cudaError_t e = cudaMemsetAsync(scratch, 0, quest::ws_hist_off(num_qo_heads), (cudaStream_t)stream);
  if (e != cudaSuccess) return (int)e;
  return decode_run_checked(q, o, nullptr, lse, kv_data, indices, idx_row_stride, num_sel_pages, last_page_idx, last_page_len,
                            num_qo_heads, num_kv_heads, head_dim, page_size, pages_per_chunk, scratch,
                            quest::kStaticScratchBytes, stream);
}

#ifdef QUEST_DEC_CLOCKS
extern "C" int quest_debug_decode_clocks(unsigned long long* host_out, uint32_t n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, quest::g_dec_clk, (size_t)n_ctas * 8 * 8);
}
#endif
