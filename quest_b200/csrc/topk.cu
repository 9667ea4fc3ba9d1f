// Per-head top-k page selector: one CTA per head, scores resident in shared memory as order-preserving
// 16-bit keys, four 4-bit MSB radix-select passes (topk.cuh), then a stable (ascending position) compaction.
// Replaces decode_select_k -> raft radix_topk_one_block_kernel<half,int,8,512>
// (reference kernels/include/topk/decode_select_k.cuh:25-62; raft select_radix.cuh:82-104 key twiddle,
// :380-399 choose_bucket, :404-440 last_filter).  Unlike raft's atomicAdd race among ties, the tie rule
// here is deterministic: lowest position first.
#include <stdlib.h>

#include <cooperative_groups.h>

#include "topk.cuh"

namespace quest {

constexpr uint32_t kTopkMaxLen = 96 * 1024;  // 192 KB of u16 keys

template <int W, int KPC>
__global__ void __launch_bounds__(32 * W)
topk_kernel(const uint16_t* __restrict__ scores, const int32_t* __restrict__ in_idx, uint32_t in_stride,
            uint16_t* __restrict__ out_val, int32_t* __restrict__ out_idx, uint32_t len, uint32_t k) {
  extern __shared__ __align__(16) uint16_t keys[];  // [len rounded up to 8]
  __shared__ TopkScratch<W> sc;
  constexpr uint32_t T = 32 * W;
  const uint32_t head = blockIdx.x, tid = threadIdx.x;
  const uint16_t* s = scores + (size_t)head * len;
  const uint32_t len8 = (len + 7) & ~7u;

#if QUEST_TRIG_TOPK
  pdl_trigger();
#endif
  pdl_wait();  // scores come from the estimate kernel
  // rows of an odd `len` are only 2-byte aligned: scalar loads, all independent
  for (uint32_t i = tid; i < len8; i += T) keys[i] = i < len ? key_of(s[i]) : (uint16_t)0;
  __syncthreads();
  topk_select<W, KPC>(keys, len, k, tid, 1, &sc, in_idx ? in_idx + (size_t)head * in_stride : nullptr,
                 out_idx + (size_t)head * k, out_val ? out_val + (size_t)head * k : nullptr);
}

template <int W, int G>
__global__ void __launch_bounds__(32 * W)
topk_fast_kernel(const uint16_t* __restrict__ scores, const int32_t* __restrict__ in_idx_, uint32_t in_stride_,
                 uint16_t* __restrict__ out_val, int32_t* __restrict__ out_idx_, uint32_t len_, uint32_t k_,
                 const BatchArgs ba) {
  constexpr int NK = 4 * G;  // keys per thread
  __shared__ __align__(16) uint16_t keys[NK * 32 * W];
  __shared__ TopkFastScratch<W, G> sc;
  constexpr uint32_t T = 32 * W;
  const uint32_t head = blockIdx.x, tid = threadIdx.x;
  const int32_t* __restrict__ in_idx = in_idx_;
  int32_t* __restrict__ out_idx = out_idx_;
  uint32_t in_stride = in_stride_, len = len_, k = k_;
#if QUEST_TRIG_TOPK
  pdl_trigger();
#endif
  pdl_wait();  // scores come from the estimate kernel
  if (ba.descs != nullptr) {
    // descriptor mode: blockIdx.y = sequence; row length and payload (the sequence's page table) from its record.
    // A sequence with fewer scored pages than k selects all of them; the rest of its row is -1 (no page).
    const SeqDesc* d = ba.descs + blockIdx.y;
    len = d->num_kv_pages - 1;
    in_idx = d->kv_indices;
    in_stride = 0;
    scores += (size_t)blockIdx.y * ba.score_stride;
    out_idx += (size_t)blockIdx.y * gridDim.x * k_;
    if (len > (uint32_t)(NK * T)) len = NK * T;  // never: the host sizes the variant by its bound
    if (k > len) {
      for (uint32_t i = len + tid; i < k_; i += T) out_idx[(size_t)head * k_ + i] = -1;
      k = len;
    }
    if (k == 0) return;
  }
  const uint16_t* s = scores + (size_t)head * len;
  // rows of an odd `len` are only 2-byte aligned: scalar loads, all independent (NK per thread); the whole key
  // array is written (0xffff -> key 0 past the row's end)
  uint16_t v[NK];
#pragma unroll
  for (int j = 0; j < NK; ++j) v[j] = tid + j * T < len ? s[tid + j * T] : (uint16_t)0xffffu;
#pragma unroll
  for (int j = 0; j < NK; ++j) keys[tid + j * T] = key_of(v[j]);
  __syncthreads();
  topk_select_fast<W, TopkNoHook, G>(keys, len, k, tid, 1, &sc, in_idx ? in_idx + (size_t)head * in_stride : nullptr,
                                     out_idx + (size_t)head * k_, out_val ? out_val + (size_t)head * k_ : nullptr);
}

template <int W, int G>
static int launch_topk_fast(const void* scores, const int32_t* in_idx, uint32_t in_stride, void* out_val, int32_t* out_idx,
                            uint32_t num_heads, uint32_t len, uint32_t k, cudaStream_t stream,
                            const BatchArgs& ba = BatchArgs{}, uint32_t batch = 1) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_heads, batch);
  cfg.blockDim = dim3(32 * W);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, topk_fast_kernel<W, G>, (const uint16_t*)scores, in_idx, in_stride,
                                     (uint16_t*)out_val, out_idx, len, k, ba);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

// ---- one row shared by a thread-block cluster --------------------------------------------------------------------
// The selection is bound by the integer pipe of the ONE SM that holds a row (a 8191-page row: 11.3 us, a 2047-page
// row: 3.5 us), and only `heads` of the 148 SMs are busy.  Rows longer than 2048 pages are therefore split over a
// cluster of C CTAs (C x 2048 keys, 4 per thread): every CTA counts its slice, the 16 bin totals of each pass and the
// (greater, equal) totals of the final compaction are exchanged through distributed shared memory with one cluster
// barrier each (slots double-buffered on the pass parity), and every CTA takes the same decisions.
template <int C>
struct TopkClusterXc {
  uint32_t (*bins)[C][16];   // [parity][rank][bin] in this CTA's shared memory
  uint32_t* tot;             // [rank]
  uint32_t rank;
  __device__ __forceinline__ uint32_t all_bins(uint32_t s, uint32_t lane, int pass) const {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    if (threadIdx.x < 16) {
#pragma unroll
      for (int r = 0; r < C; ++r) cl.map_shared_rank(&bins[pass & 1][rank][0], r)[threadIdx.x] = s;
    }
    cl.sync();
    uint32_t a = 0;
    if (lane < 16) {
#pragma unroll
      for (int r = 0; r < C; ++r) a += bins[pass & 1][r][lane];
    }
    return a;
  }
  __device__ __forceinline__ uint32_t before_me(uint32_t cta_total, uint32_t t) const {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    if (t == 0) {
#pragma unroll
      for (int r = 0; r < C; ++r) *cl.map_shared_rank(&tot[rank], r) = cta_total;
    }
    cl.sync();  // also the last remote access to anybody's shared memory: CTAs may exit after it
    uint32_t a = 0;
#pragma unroll
    for (int r = 0; r < C; ++r)
      if ((uint32_t)r < rank) a += tot[r];
    return a;
  }
};

template <int W, int C>
__global__ void __launch_bounds__(32 * W)
topk_cluster_kernel(const uint16_t* __restrict__ scores, const int32_t* __restrict__ in_idx_, uint32_t in_stride_,
                    uint16_t* __restrict__ out_val, int32_t* __restrict__ out_idx_, uint32_t len_, uint32_t k_,
                    const BatchArgs ba) {
  namespace cg = cooperative_groups;
  constexpr uint32_t T = 32 * W, NK = 4;
  __shared__ __align__(16) uint16_t keys[NK * T];
  __shared__ TopkFastScratch<W, 1> sc;
  __shared__ uint32_t x_bins[2][C][16];
  __shared__ uint32_t x_tot[C];
  const uint32_t rank = cg::this_cluster().block_rank();
  const uint32_t head = blockIdx.x / C, n_heads = gridDim.x / C, tid = threadIdx.x;
  const int32_t* __restrict__ in_idx = in_idx_;
  int32_t* __restrict__ out_idx = out_idx_;
  uint32_t in_stride = in_stride_, len = len_, k = k_;
  pdl_wait();  // scores come from the estimate kernel
  if (ba.descs != nullptr) {  // descriptor mode, as in topk_fast_kernel (uniform over the cluster)
    const SeqDesc* d = ba.descs + blockIdx.y;
    len = d->num_kv_pages - 1;
    in_idx = d->kv_indices;
    in_stride = 0;
    scores += (size_t)blockIdx.y * ba.score_stride;
    out_idx += (size_t)blockIdx.y * n_heads * k_;
    if (len > C * NK * T) len = C * NK * T;
    if (k > len) {
      if (rank == 0)
        for (uint32_t i = len + tid; i < k_; i += T) out_idx[(size_t)head * k_ + i] = -1;
      k = len;
    }
    if (k == 0) return;  // the whole cluster returns: nobody waits at a cluster barrier
  }
  const uint16_t* s = scores + (size_t)head * len;
  const uint32_t pos0 = rank * NK * T;
  uint16_t v[NK];
#pragma unroll
  for (uint32_t j = 0; j < NK; ++j) v[j] = pos0 + tid + j * T < len ? s[pos0 + tid + j * T] : (uint16_t)0xffffu;
#pragma unroll
  for (uint32_t j = 0; j < NK; ++j) keys[tid + j * T] = key_of(v[j]);
  cg::this_cluster().sync();  // keys visible to the CTA; every CTA of the cluster is running before anyone writes into it
  TopkClusterXc<C> xc = {x_bins, x_tot, rank};
  topk_select_fast<W, TopkNoHook, 1, TopkClusterXc<C>>(
      keys, len, k, tid, 1, &sc, in_idx ? in_idx + (size_t)head * in_stride : nullptr, out_idx + (size_t)head * k_,
      out_val ? out_val + (size_t)head * k_ : nullptr, TopkNoHook(), xc, pos0);
}

template <int W, int C>
static int launch_topk_cluster(const void* scores, const int32_t* in_idx, uint32_t in_stride, void* out_val, int32_t* out_idx,
                               uint32_t num_heads, uint32_t len, uint32_t k, cudaStream_t stream,
                               const BatchArgs& ba = BatchArgs{}, uint32_t batch = 1) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_heads * C, batch);
  cfg.blockDim = dim3(32 * W);
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = C;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  cudaError_t e = cudaLaunchKernelEx(&cfg, topk_cluster_kernel<W, C>, (const uint16_t*)scores, in_idx, in_stride,
                                     (uint16_t*)out_val, out_idx, len, k, ba);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

// ---- selection from published bin totals (layer step) -------------------------------------------------------------
// In the layer step the estimate kernel counts every score into its head's 2048 bin totals (key >> 5; common.cuh
// hist_count) while it writes it, so the selection does not have to histogram the row: the threshold BIN is one block
// scan over the totals (raft spends a histogram pass over the data per 8 key bits, select_radix.cuh:930-1090; the
// register radix select above four 4-bit passes), the low 5 key bits are resolved among the few keys of that bin, and
// the ranks are one more block scan.  Same selection rule and output order as topk_select_fast.  Every thread owns
// 4*G consecutive positions (blocked, so ranks follow positions) and 2048/512 = 4 bins, which it zeroes again once
// read: the totals are back at rest when the kernel ends.  Rows of up to 4*G*512 pages (G = 4: 8192).
constexpr int kHistThreads = 512;

// -DQUEST_TOPK_HIST_CLOCKS: every CTA stamps %globaltimer at its phase boundaries (tools/step_timeline.py)
#ifdef QUEST_TOPK_HIST_CLOCKS
__device__ unsigned long long g_topk_hist_clk[256 * 4];
#define TH_STAMP(i)                                                                \
  do {                                                                             \
    if (threadIdx.x == 0) {                                                        \
      unsigned long long t_;                                                       \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                       \
      g_topk_hist_clk[(blockIdx.x & 255) * 4 + (i)] = t_;                          \
    }                                                                              \
  } while (0)
#else
#define TH_STAMP(i) do { } while (0)
#endif

// kOwn: nobody has published bin totals (stand-alone quest_topk, the global selection of the sequence-sharded step): the CTA
// counts its register-resident keys into 2048 shared-memory bins first -- one pass of shared-memory atomics instead of the four
// radix passes of the register select -- and goes on as above.
// C > 1: the row is split over a thread-block cluster of C CTAs (2048*G positions each, in rank order).  Every CTA reads the same
// bin totals and takes the same decisions; what depends on the other slices -- the threshold bin's low-bit counts, the (greater,
// equal) totals of the lower ranks, with kOwn the bin counts themselves -- is read from their shared memory after a cluster
// barrier (three in all, four with kOwn).  The integer work per row is what bounds a selection on one SM (8191 scores: 7.7 us in
// one CTA), and a batch-1 step keeps only `heads` of the 148 SMs busy.
template <int G, bool kOwn = false, int C = 1>
__global__ void __launch_bounds__(kHistThreads, (G >= 4 && G <= 6) ? 2 : 0)  // two rows per SM while 64 registers hold the keys (batched steps: rows > SMs)
topk_hist_kernel(const uint16_t* __restrict__ scores, uint32_t* __restrict__ hist_, const int32_t* __restrict__ in_idx_,
                 uint32_t in_stride_, int32_t* __restrict__ out_idx_, uint32_t len_, uint32_t k_, const BatchArgs ba,
                 const Handoff ho, const HistOut xo) {
  constexpr int NK = 4 * G, W = kHistThreads / 32;
  constexpr uint32_t BPT = kHistBins / kHistThreads;  // 4 bins = two packed words per thread
  static_assert(BPT == 4, "one 8-byte load of bin totals per thread");
  __shared__ uint32_t s_wtot[W], s_wtot2[W], s_sub[32], s_bstar, s_need1;
  __shared__ __align__(16) uint32_t s_bins[kOwn ? kHistBins : 4];
  __shared__ uint32_t s_tot;
  namespace cg = cooperative_groups;
  uint32_t rank = 0;
  if constexpr (C > 1) rank = cg::this_cluster().block_rank();
  const uint32_t head = blockIdx.x / C, n_heads = gridDim.x / C, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int32_t* __restrict__ in_idx = in_idx_;
  int32_t* __restrict__ out_idx = out_idx_;
  uint32_t* __restrict__ hist = hist_;
  uint32_t in_stride = in_stride_, len = len_, k = k_;
  // hand-over mode: the decode launch may start now; its CTAs poll for the ids this kernel deposits (common.cuh, Handoff)
  // (releasing it later, once the threshold bin is known, measured the same: 18.15 against 18.13 us per layer step)
  // With QUEST_TRIG_EST the estimate launch releases this kernel before it has completed: the decode CTAs of the hand-over, which
  // wait for nothing, may then only be released from behind the wait below.
  constexpr bool kReleaseLate = QUEST_TRIG_EST != 0;
  if (!kReleaseLate && (QUEST_TRIG_TOPK || ho.epoch != nullptr)) pdl_trigger();
  if (t < 32) s_sub[t] = 0;
  TH_STAMP(0);
  pdl_wait();  // scores and bin totals come from the estimate kernel
  if (kReleaseLate && (QUEST_TRIG_TOPK || ho.epoch != nullptr)) pdl_trigger();
  TH_STAMP(1);
  const unsigned long long tag = ho.epoch != nullptr ? (unsigned long long)handoff_tag(ld_ordered32(ho.epoch)) << 32 : 0ull;
  if (ba.descs != nullptr) {  // descriptor mode, as in topk_fast_kernel
    const SeqDesc* d = ba.descs + blockIdx.y;
    len = d->num_kv_pages - 1;
    in_idx = d->kv_indices;
    in_stride = 0;
    scores += (size_t)blockIdx.y * ba.score_stride;
    out_idx += (size_t)blockIdx.y * n_heads * k_;
    hist += (size_t)blockIdx.y * (ba.ws_stride / 4);
    if (len > (uint32_t)(C * NK * kHistThreads)) len = C * NK * kHistThreads;  // never: the host sizes the variant by its bound
  }
  // this thread's bins [2048 - 4(t+1), 2048 - 4t): read, and put back to zero for the next step
  uint2 hv = make_uint2(0u, 0u);
  uint2* hp = nullptr;
  if (!kOwn) {
    hp = reinterpret_cast<uint2*>(hist + (size_t)head * (kHistBins / 2) + (kHistBins - BPT * (t + 1)) / 2);
    hv = *hp;
    if (C == 1) *hp = make_uint2(0u, 0u);  // C > 1: rank 0, once every CTA of the cluster has read them
  } else {
    *reinterpret_cast<uint4*>(s_bins + BPT * t) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (k > len) {
    for (uint32_t i = len + t; i < k_ && rank == 0; i += kHistThreads) {
      out_idx[(size_t)head * (xo.idx_stride ? xo.idx_stride : k_) + i] = -1;
      if (xo.val != nullptr) xo.val[(size_t)head * (xo.val_stride ? xo.val_stride : k_) + i] = 0xfc00u;  // -inf: sorts last
    }
    k = len;
  }
  if (k == 0) {  // uniform over the cluster: nobody waits at a barrier
    if (C > 1 && !kOwn && rank == 0) *hp = make_uint2(0u, 0u);
    return;
  }
  const uint16_t* s = scores + (size_t)head * len;
  const uint32_t p0 = (rank * kHistThreads + t) * NK;
  // rows of an odd `len` are only 2-byte aligned: scalar loads, all independent
  uint32_t key[NK];
  int32_t pay[NK];
#pragma unroll
  for (int j = 0; j < NK; ++j) key[j] = p0 + j < len ? (uint32_t)score_key(s[p0 + j]) : 0u;
  const int32_t* pt = in_idx ? in_idx + (size_t)head * in_stride : nullptr;
#pragma unroll
  for (int j = 0; j < NK; ++j) pay[j] = p0 + j < len ? (pt ? __ldg(pt + p0 + j) : (int32_t)(p0 + j)) : 0;
  const uint32_t nvalid = len > p0 ? min(len - p0, (uint32_t)NK) : 0u;

  uint32_t tau = 0, need2 = 0;
  const bool all = k >= len;  // every page of the row
  if (!all) {
    // ---- threshold bin: the inclusive scan over threads counts the keys in a thread's bins and all higher ones
    uint32_t c[BPT] = {hv.x & 0xffffu, hv.x >> 16, hv.y & 0xffffu, hv.y >> 16};  // ascending bins
    if (kOwn) {
      __syncthreads();  // bins zeroed
#pragma unroll
      for (int j = 0; j < NK; ++j)
        if ((uint32_t)j < nvalid) atomicAdd(&s_bins[key[j] >> kHistBinShift], 1u);
      __syncthreads();
      if constexpr (C > 1) {
        cg::cluster_group cl = cg::this_cluster();
        cl.sync();  // every slice counted
        c[0] = c[1] = c[2] = c[3] = 0;
#pragma unroll
        for (int r = 0; r < C; ++r) {
          const uint4 cv = *reinterpret_cast<const uint4*>(cl.map_shared_rank(s_bins, r) + kHistBins - BPT * (t + 1));
          c[0] += cv.x, c[1] += cv.y, c[2] += cv.z, c[3] += cv.w;
        }
      } else {
        const uint4 cv = *reinterpret_cast<const uint4*>(s_bins + kHistBins - BPT * (t + 1));
        c[0] = cv.x, c[1] = cv.y, c[2] = cv.z, c[3] = cv.w;
      }
    }
    const uint32_t mine = c[0] + c[1] + c[2] + c[3];
    uint32_t incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t a = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= (uint32_t)off) incl += a;
    }
    if (lane == 31) s_wtot[warp] = incl;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < W; ++w)
      if ((uint32_t)w < warp) incl += s_wtot[w];
    uint32_t above = incl - mine;
    if (above < k && k <= incl) {  // exactly one thread: the k-th largest key lies in one of its bins
      uint32_t b = 0, n1 = 0;
#pragma unroll
      for (int x = (int)BPT - 1; x >= 0; --x) {
        const bool hit = above < k && k <= above + c[x];
        b = hit ? (uint32_t)x : b;
        n1 = hit ? k - above : n1;
        above += c[x];
      }
      s_bstar = kHistBins - BPT * (t + 1) + b;
      s_need1 = n1;
    }
    __syncthreads();
    const uint32_t bstar = s_bstar, need1 = s_need1;
    // ---- the low key bits among the keys of the threshold bin
#pragma unroll
    for (int j = 0; j < NK; ++j)
      if ((uint32_t)j < nvalid && (key[j] >> kHistBinShift) == bstar) atomicAdd(&s_sub[key[j] & 31u], 1u);
    __syncthreads();
    if constexpr (C > 1) cg::this_cluster().sync();  // every slice's low-bit counts are in place; every CTA has read the bin totals
    {  // every warp resolves the sub-bin by itself: lane i looks at sub-bin 31 - i
      uint32_t cs = s_sub[31 - lane];
      if constexpr (C > 1) {
        cs = 0;
#pragma unroll
        for (int r = 0; r < C; ++r) cs += cg::this_cluster().map_shared_rank(s_sub, r)[31 - lane];
      }
      uint32_t inc = cs;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= (uint32_t)off) inc += a;
      }
      const uint32_t ab = inc - cs;
      const uint32_t hit = __ballot_sync(0xffffffffu, ab < need1 && need1 <= inc);
      const uint32_t l = (__ffs(hit) - 1) & 31u;
      tau = (bstar << kHistBinShift) | (31 - l);
      need2 = need1 - __shfl_sync(0xffffffffu, ab, l);
    }
  }
  // ---- ranks: (greater, equal) counts of the positions before this thread's, as one packed block scan
  const uint32_t bump = all ? 1u : 0u;  // all: tau = 0 and every in-row key counts as greater
  uint32_t ngt = 0, neq = 0;
#pragma unroll
  for (int j = 0; j < NK; ++j) {
    const bool in = (uint32_t)j < nvalid;
    ngt += in && key[j] + bump > tau;
    neq += in && key[j] + bump == tau;
  }
  const uint32_t pk = (ngt << 16) | neq;
  uint32_t spk = pk;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, spk, off);
    if (lane >= (uint32_t)off) spk += a;
  }
  if (lane == 31) s_wtot2[warp] = spk;
  __syncthreads();
  uint32_t before = spk - pk;
#pragma unroll
  for (int w = 0; w < W; ++w)
    if ((uint32_t)w < warp) before += s_wtot2[w];
  if constexpr (C > 1) {  // + the (greater, equal) totals of the slices before this one
    cg::cluster_group cl = cg::this_cluster();
    if (t == kHistThreads - 1) s_tot = before + pk;
    if (!kOwn && rank == 0) *hp = make_uint2(0u, 0u);  // all: no barrier has been passed yet, but nobody reads the totals either
    cl.sync();
#pragma unroll
    for (int r = 0; r < C; ++r)
      if ((uint32_t)r < rank) before += *cl.map_shared_rank(&s_tot, r);
  }
  uint32_t gt_before = before >> 16, eq_before = before & 0xffffu;
  int32_t* out = out_idx + (size_t)head * (xo.idx_stride ? xo.idx_stride : k_);
  uint16_t* outv = xo.val != nullptr ? xo.val + (size_t)head * (xo.val_stride ? xo.val_stride : k_) : nullptr;
#pragma unroll
  for (int j = 0; j < NK; ++j) {
    const bool in = (uint32_t)j < nvalid;
    const bool gt = in && key[j] + bump > tau, eq = in && key[j] + bump == tau;
    const uint32_t pos = gt_before + min(eq_before, need2);
    if ((gt || (eq && eq_before < need2)) && pos < k) {  // pos < k always, unless the totals were not at rest
      out[pos] = pay[j];
      if (outv != nullptr) outv[pos] = (key[j] & 0x8000u) ? (uint16_t)(key[j] & 0x7fffu) : (uint16_t)~key[j];  // score_key undone
      if (ho.epoch != nullptr) {
        const uint32_t chunk = pos / ho.ppc;
        st_relaxed_u64(ho.slots + (size_t)(head * ho.nchunks + chunk) * ho.slot_bytes + (pos - chunk * ho.ppc) * 8u,
                       tag | (uint32_t)pay[j]);
      }
    }
    gt_before += gt;
    eq_before += eq;
  }
  if constexpr (C > 1) cg::this_cluster().sync();  // nobody leaves while its shared memory may still be read
  TH_STAMP(2);
}

#ifdef QUEST_TOPK_HIST_CLOCKS
extern "C" int quest_debug_topk_hist_clocks(unsigned long long* host_out, uint32_t n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, g_topk_hist_clk, (size_t)n_ctas * 4 * 8);
}
#endif

template <int G, bool kOwn = false, int C = 1>
static int launch_topk_hist(const void* scores, uint32_t* hist, const int32_t* in_idx, uint32_t in_stride, int32_t* out_idx,
                            uint32_t num_heads, uint32_t len, uint32_t k, cudaStream_t stream,
                            const BatchArgs& ba = BatchArgs{}, uint32_t batch = 1, const Handoff& ho = Handoff{},
                            const HistOut& xo = HistOut{}) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_heads * C, batch);
  cfg.blockDim = dim3(kHistThreads);
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = C;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = C > 1 ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, topk_hist_kernel<G, kOwn, C>, (const uint16_t*)scores, hist, in_idx, in_stride, out_idx, len, k, ba, ho, xo);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

static bool topk_cluster_fits(uint32_t rows, uint32_t c);
static bool topk_use_cluster();

// Layer step: rows whose bin totals the estimate kernel has published in `hist` (max_len bounds every row).
bool topk_hist_covers(uint32_t max_len) { return max_len <= 32u * kHistThreads && debug_option(QUEST_DBG_TOPK_NO_HIST) == 0; }
int topk_from_hist(const void* scores, uint32_t* hist, const int32_t* in_idx, uint32_t in_stride, int32_t* out_idx,
                   uint32_t num_heads, uint32_t max_len, uint32_t k, const BatchArgs& ba, uint32_t batch, cudaStream_t s,
                   const Handoff& ho, const HistOut& xo) {
  if (max_len <= 4u * kHistThreads) return launch_topk_hist<1>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  // (a cluster of two CTAs at 2049 ... 4096 pages is slower than one CTA with 8 keys per thread: 5.4 against 4.7 us)
  if (max_len <= 8u * kHistThreads) return launch_topk_hist<2>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  // rows of 4097 ... 16384 pages: one row over a cluster of 4 CTAs (2048 / 4096 positions each) while all clusters run in one wave
  if (topk_use_cluster() && max_len <= 16u * kHistThreads && topk_cluster_fits(num_heads * batch, 4))
    return launch_topk_hist<1, false, 4>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  if (topk_use_cluster() && max_len <= 32u * kHistThreads && topk_cluster_fits(num_heads * batch, 4))
    return launch_topk_hist<2, false, 4>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  if (max_len <= 16u * kHistThreads) return launch_topk_hist<4>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  // in-between bounds: a batch just past 8192 pages (BASELINE configs[3]: 8193) pays for the keys per thread it is compiled for
  if (max_len <= 20u * kHistThreads) return launch_topk_hist<5>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  if (max_len <= 24u * kHistThreads) return launch_topk_hist<6>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  if (max_len <= 32u * kHistThreads) return launch_topk_hist<8>(scores, hist, in_idx, in_stride, out_idx, num_heads, max_len, k, s, ba, batch, ho, xo);
  return QUEST_ERR_UNSUPPORTED;
}

// Stand-alone rows of up to 4096 scores: the same one-pass selection over bin totals the CTA counts itself (32 heads, graph:
// 3.8 -> 3.1 us at 2047 scores, 6.2 -> 4.6 us at 4095; at 8191 the 4-CTA cluster kernel stays ahead, 6.7 against 7.7 us).
static bool topk_own_hist_covers(uint32_t len, uint32_t rows) {
  if (!topk_hist_covers(len)) return false;
  return len <= 8u * kHistThreads || (topk_use_cluster() && topk_cluster_fits(rows, 4));
}
static int topk_own_hist(const void* scores, const int32_t* in_idx, uint32_t in_stride, void* out_val, int32_t* out_idx,
                         uint32_t num_heads, uint32_t len, uint32_t k, cudaStream_t s) {
  const HistOut xo = {(uint16_t*)out_val, 0u, 0u};
  const BatchArgs ba = {};
  const Handoff ho = {};
#define QUEST_OWN(G) launch_topk_hist<G, true>(scores, nullptr, in_idx, in_stride, out_idx, num_heads, len, k, s, ba, 1, ho, xo)
  if (len <= 4u * kHistThreads) return QUEST_OWN(1);
  if (len <= 8u * kHistThreads) return QUEST_OWN(2);
#undef QUEST_OWN
  if (len <= 16u * kHistThreads) return launch_topk_hist<1, true, 4>(scores, nullptr, in_idx, in_stride, out_idx, num_heads, len, k, s, ba, 1, ho, xo);
  return launch_topk_hist<2, true, 4>(scores, nullptr, in_idx, in_stride, out_idx, num_heads, len, k, s, ba, 1, ho, xo);
}

// A cluster pays off while all its CTAs (512 threads at 40 registers: 3 per SM) run in one wave; beyond that the single-CTA rows
// (1024 threads, 2 per SM) finish sooner (configs[3] on one GPU: 8 x 32 rows).
static bool topk_cluster_fits(uint32_t rows, uint32_t c) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return (uint64_t)rows * c <= (uint64_t)3 * sms;
}

static bool topk_use_cluster() { return debug_option(QUEST_DBG_TOPK_NO_CLUSTER) == 0; }  // A/B hook

// descriptor mode (quest_decode_layer_step_batch): max_len bounds every sequence's row; fast variants only
int topk_batch(const void* scores, int32_t* out_idx, uint32_t num_heads, uint32_t max_len, uint32_t k, const BatchArgs& ba,
               uint32_t batch, cudaStream_t s) {
  if (max_len <= 1024) return launch_topk_fast<8, 1>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
  if (max_len <= 2048) return launch_topk_fast<16, 1>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
  if (topk_use_cluster()) {
    if (max_len <= 4096) return launch_topk_fast<32, 1>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);  // measured equal
    if (max_len > 4096 && max_len <= 8192 && topk_cluster_fits(num_heads * batch, 4))
      return launch_topk_cluster<16, 4>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
    if (max_len > 8192 && max_len <= 16384 && topk_cluster_fits(num_heads * batch, 8))
      return launch_topk_cluster<16, 8>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
  }
  if (max_len <= 4096) return launch_topk_fast<32, 1>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
  if (max_len <= 8192) return launch_topk_fast<32, 2>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
  if (max_len <= 16384) return launch_topk_fast<32, 4>(scores, nullptr, 0, nullptr, out_idx, num_heads, max_len, k, s, ba, batch);
  return QUEST_ERR_UNSUPPORTED;
}

template <int W, int KPC>
static int launch_topk(const void* scores, const int32_t* in_idx, uint32_t in_stride, void* out_val, int32_t* out_idx,
                       uint32_t num_heads, uint32_t len, uint32_t k, cudaStream_t stream) {
  const size_t smem = (((size_t)len + 7) & ~(size_t)7) * sizeof(uint16_t);
  auto kern = topk_kernel<W, KPC>;
  if (smem > 40 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_heads);
  cfg.blockDim = dim3(32 * W);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, (const uint16_t*)scores, in_idx, in_stride, (uint16_t*)out_val,
                                     out_idx, len, k);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

}  // namespace quest

namespace quest {
// test helper: the bin totals the estimate kernel would have published for these scores
__global__ void hist_build_kernel(const __half* __restrict__ scores, uint32_t* __restrict__ hist, uint32_t len) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x)
    hist_count(hist, blockIdx.y, scores[(size_t)blockIdx.y * len + i]);
}
}  // namespace quest

// Test hook (not part of the operator surface): topk_hist_kernel on arbitrary scores.  hist: num_heads x 4 KB, zero on
// entry and zero again when the kernel has run.
extern "C" int quest_debug_topk_hist(const void* scores, const int32_t* in_idx, uint32_t in_idx_row_stride, int32_t* out_idx,
                                     uint32_t num_heads, uint32_t len, uint32_t k, void* hist, void* stream) {
  if (!scores || !out_idx || !hist || !num_heads || k > len || !quest::topk_hist_covers(len)) return QUEST_ERR_BAD_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  quest::hist_build_kernel<<<dim3(8, num_heads), 256, 0, s>>>((const __half*)scores, (uint32_t*)hist, len);
  return quest::topk_from_hist(scores, (uint32_t*)hist, in_idx, in_idx_row_stride, out_idx, num_heads, len, k, quest::BatchArgs{}, 1, s, quest::Handoff{}, quest::HistOut{});
}

extern "C" uint32_t quest_topk_max_len(void) { return quest::kTopkMaxLen; }

extern "C" int quest_topk(const void* scores, const int32_t* in_idx, uint32_t in_idx_row_stride, void* out_val,
                          int32_t* out_idx, uint32_t num_heads, uint32_t len, uint32_t k, void* stream) {
  if (!scores || !out_idx || !num_heads) return QUEST_ERR_BAD_ARG;
  if (k == 0) return QUEST_OK;
  if (k > len) return QUEST_ERR_BAD_ARG;
  if (len > quest::kTopkMaxLen) return QUEST_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  const bool no_fast = quest::debug_option(QUEST_DBG_TOPK_GENERIC) != 0;  // test hook: force the generic kernels
  // one pass over self-counted bin totals (QUEST_DBG_TOPK_NO_HIST = 1: the register radix / cluster kernels below)
  if (!no_fast && quest::topk_own_hist_covers(len, num_heads))
    return quest::topk_own_hist(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (!no_fast && len <= 1024) return quest::launch_topk_fast<8, 1>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (!no_fast && len <= 2048) return quest::launch_topk_fast<16, 1>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (!no_fast && quest::topk_use_cluster()) {
    if (len <= 4096) return quest::launch_topk_fast<32, 1>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);  // measured equal
    if (len > 4096 && len <= 8192 && quest::topk_cluster_fits(num_heads, 4))
      return quest::launch_topk_cluster<16, 4>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
    if (len > 8192 && len <= 16384 && quest::topk_cluster_fits(num_heads, 8))
      return quest::launch_topk_cluster<16, 8>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  }
  if (!no_fast && len <= 4096) return quest::launch_topk_fast<32, 1>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (!no_fast && len <= 8192) return quest::launch_topk_fast<32, 2>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (!no_fast && len <= 16384) return quest::launch_topk_fast<32, 4>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (len <= 2048) return quest::launch_topk<16, 4>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  if (len <= 8192) return quest::launch_topk<32, 8>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
  return quest::launch_topk<32, 8>(scores, in_idx, in_idx_row_stride, out_val, out_idx, num_heads, len, k, s);
}
