// Page selection inside the consumer: every (head, chunk) CTA of the sparse decode kernel derives its own page ids
// from what the estimate kernel left behind for the head, so the layer step has no top-k launch and no [H, k] id
// round trip between two kernels.
//
// Replaces, for the fused layer step, decode_select_k -> raft radix_topk_one_block_kernel<half,int,8,512>
// (reference kernels/include/topk/decode_select_k.cuh:25-62; raft select_radix.cuh:82-104 key twiddle, :380-399
// choose_bucket, :404-440 last_filter) with the SAME selection rule as topk.cuh: every key > tau
// (the k-th largest) is taken, keys == tau lowest position first, output in ascending position.
//
// The estimate kernel's epilogue publishes, per q head:
//   keys[h][pos]   the order-preserving 16-bit key of score[h][pos] (aligned row of key_stride entries)
//   hist[h][b]     how many of the head's keys have key >> 5 == b (2048 bins, 16-bit counts packed in pairs;
//                  red.global.add, so the counts exist the moment the estimate grid has drained)
// From the bin totals the threshold bin is one block scan (raft needs a histogram pass over the data per 8 bits);
// the low 5 bits are resolved among the handful of keys in that bin; the ranks are one more block scan.  Every
// chunk CTA of a head does this redundantly (4 + 4 KB from L2 at 2047 pages) -- redundant work is free while the
// SMs wait for the selection anyway, and nobody waits on anybody else.
#pragma once
#include "common.cuh"

namespace quest {

constexpr uint32_t kSelBinShift = 5;
constexpr uint32_t kSelBins = 1u << (16 - kSelBinShift);          // 2048
constexpr uint32_t kSelHistBytes = kSelBins * 2;                  // per head
constexpr uint32_t kSelMaxLen = 65535;                            // 16-bit counts

__host__ __device__ inline uint32_t sel_keys_per_thread(uint32_t len, uint32_t threads) {
  const uint32_t kpt = (len + threads - 1) / threads;
  return (kpt + 7) & ~7u;  // whole 16-byte vectors
}
__host__ __device__ inline uint32_t sel_key_stride(uint32_t max_len, uint32_t threads) {
  return sel_keys_per_thread(max_len ? max_len : 1, threads) * threads;
}

// What the estimate kernel publishes for the selection (ws == nullptr: nothing): the step's workspace and where the
// bin totals and the key rows sit in it (a batch: one workspace block per sequence, BatchArgs::ws_stride apart).
struct SelOut {
  uint8_t* ws;
  size_t hist_off, keys_off;
  uint32_t key_stride;
};
struct SelRows {
  uint16_t* keys;  // [Hq][key_stride]
  uint32_t* hist;  // [Hq][kSelBins / 2]
  uint32_t key_stride;
};
__device__ __forceinline__ SelRows sel_rows(const SelOut& so, size_t seq_off) {
  SelRows r;
  r.keys = so.ws ? reinterpret_cast<uint16_t*>(so.ws + seq_off + so.keys_off) : nullptr;
  r.hist = so.ws ? reinterpret_cast<uint32_t*>(so.ws + seq_off + so.hist_off) : nullptr;
  r.key_stride = so.key_stride;
  return r;
}

__device__ __forceinline__ uint16_t sel_key_of(uint16_t h) { return (h & 0x8000u) ? (uint16_t)~h : (uint16_t)(h | 0x8000u); }

// estimate epilogue: one score of head h at page position pos
__device__ __forceinline__ void sel_publish(const SelRows& so, uint32_t h, uint32_t pos, __half score) {
  const uint16_t key = sel_key_of(__half_as_ushort(score));
  so.keys[(size_t)h * so.key_stride + pos] = key;
  const uint32_t bin = key >> kSelBinShift;
  asm volatile("red.global.add.u32 [%0], %1;" ::"l"(so.hist + (size_t)h * (kSelBins / 2) + (bin >> 1)), "r"(1u << ((bin & 1u) * 16))
               : "memory");
}

static_assert(kSelHistBytes == kWsHistBytesPerHead, "workspace layout (common.cuh)");

// Back to zero for the next estimate launch; called by the last CTA that needed head h's bin totals.
template <int kThreads>
__device__ __forceinline__ void sel_hist_clear(uint8_t* ws, uint32_t Hq, uint32_t h) {
  uint4* p = reinterpret_cast<uint4*>(ws + ws_hist_off(Hq) + (size_t)h * kSelHistBytes);
  for (uint32_t i = threadIdx.x; i < kSelHistBytes / 16; i += kThreads) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

template <int kThreads>
struct SelScratch {
  uint32_t wtot[kThreads / 32];
  uint32_t wtot2[kThreads / 32];
  uint32_t sub[32];
  uint32_t bstar, need1;
};

// Selects k of the head's `len` keys and hands the CTA the pages whose rank (ascending position among the selected)
// falls in [r0, r1): s_pages[rank - r0] = page_table[pos]; the same ids go to ids_out[rank] (the operator's
// topk_idx row) when ids_out != nullptr.  All kThreads threads call it; the caller synchronises before reading
// s_pages.  keys / hist: this head's rows.  k >= len selects every page.
// kRegVecs > 0: the thread's kpt == 8 * kRegVecs keys are loaded once, together with the bin totals, and stay in
// registers (rows of up to 16 * kThreads pages); 0: any row length, the keys are re-read (L1) in each phase.
template <int kThreads, int kRegVecs>
__device__ __forceinline__ void select_window_impl(const uint16_t* __restrict__ keys, const uint32_t* __restrict__ hist,
                                                   uint32_t len, uint32_t k, uint32_t r0, uint32_t r1,
                                                   const int32_t* __restrict__ page_table, int32_t* __restrict__ s_pages,
                                                   int32_t* __restrict__ ids_out, SelScratch<kThreads>* sc) {
  constexpr uint32_t W = kThreads / 32;
  constexpr uint32_t BPT = kSelBins / kThreads;  // histogram bins per thread
  static_assert(BPT % 8 == 0, "bins per thread are read as 16-byte vectors");
  const uint32_t t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const uint32_t kpt = kRegVecs ? 8u * kRegVecs : sel_keys_per_thread(len, kThreads);
  const uint32_t nvec = kpt / 8;
  const uint32_t p0 = t * kpt;  // first position this thread owns (blocked: ranks follow positions)
  const uint4* kv4 = reinterpret_cast<const uint4*>(keys + p0);
  uint4 kreg[kRegVecs ? kRegVecs : 1];
  if (kRegVecs) {
#pragma unroll
    for (int j = 0; j < kRegVecs; ++j) kreg[j] = kv4[j];
  }
  auto vec = [&](uint32_t j) -> uint4 { return kRegVecs ? kreg[kRegVecs ? j : 0] : kv4[j]; };
  uint32_t tau = 0, need2 = 0;

  if (k < len) {
    // ---- A: threshold bin from the published bin totals.  Thread t owns bins [2048 - BPT*(t+1), 2048 - BPT*t): the
    // inclusive scan over threads counts the keys in its bins and all higher ones.
    uint32_t cnt[BPT / 2];
    {
      const uint4* hp = reinterpret_cast<const uint4*>(hist + (kSelBins - BPT * (t + 1)) / 2);
#pragma unroll
      for (uint32_t x = 0; x < BPT / 8; ++x) {
        const uint4 v = hp[x];
        cnt[4 * x] = v.x, cnt[4 * x + 1] = v.y, cnt[4 * x + 2] = v.z, cnt[4 * x + 3] = v.w;
      }
    }
    if (t < 32) sc->sub[t] = 0;
    uint32_t mine = 0;
#pragma unroll
    for (uint32_t x = 0; x < BPT / 2; ++x) mine += (cnt[x] & 0xffffu) + (cnt[x] >> 16);
    uint32_t incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t a = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= (uint32_t)off) incl += a;
    }
    if (lane == 31) sc->wtot[warp] = incl;
    __syncthreads();
#pragma unroll
    for (uint32_t w = 0; w < W; ++w)
      if (w < warp) incl += sc->wtot[w];
    uint32_t above = incl - mine;
    if (above < k && k <= incl) {  // exactly one thread: the k-th largest key lies in one of its bins
#pragma unroll
      for (int x = (int)BPT - 1; x >= 0; --x) {
        const uint32_t c = (cnt[x >> 1] >> ((x & 1) * 16)) & 0xffffu;
        if (above < k && k <= above + c) {
          sc->bstar = kSelBins - BPT * (t + 1) + x;
          sc->need1 = k - above;
        }
        above += c;
      }
    }
    __syncthreads();
    const uint32_t bstar = sc->bstar, need1 = sc->need1;

    // ---- B: the low bits among the keys of the threshold bin (a handful: 2047 keys spread over ~50 bins)
#pragma unroll
    for (uint32_t j = 0; j < (kRegVecs ? kRegVecs : nvec); ++j) {
      const uint4 v = vec(j);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int x = 0; x < 8; ++x) {
        const uint32_t key = (x & 1) ? (w[x >> 1] >> 16) : (w[x >> 1] & 0xffffu);
        if (p0 + j * 8 + x < len && (key >> kSelBinShift) == bstar) atomicAdd(&sc->sub[key & 31u], 1u);
      }
    }
    __syncthreads();
    {  // every warp resolves the sub-bin by itself (no broadcast round): lane i looks at sub-bin 31 - i
      const uint32_t c = sc->sub[31 - lane];
      uint32_t inc = c;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= (uint32_t)off) inc += a;
      }
      const uint32_t ab = inc - c;
      const uint32_t hit = __ballot_sync(0xffffffffu, ab < need1 && need1 <= inc);
      const uint32_t l = (__ffs(hit) - 1) & 31u;
      tau = (bstar << kSelBinShift) | (31 - l);
      need2 = need1 - __shfl_sync(0xffffffffu, ab, l);
    }
  }
  // k >= len: tau = 0, need2 = 0 and every in-row key counts as "greater"

  // ---- C: ranks.  (greater, equal) counts of the positions before this thread's, as one packed block scan.
  const bool all = k >= len;
  uint32_t ngt = 0, neq = 0;
#pragma unroll
  for (uint32_t j = 0; j < (kRegVecs ? kRegVecs : nvec); ++j) {
    const uint4 v = vec(j);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int x = 0; x < 8; ++x) {
      const uint32_t key = (x & 1) ? (w[x >> 1] >> 16) : (w[x >> 1] & 0xffffu);
      const bool in = p0 + j * 8 + x < len;
      ngt += in && (all || key > tau);
      neq += in && !all && key == tau;
    }
  }
  const uint32_t pk = (ngt << 16) | neq;
  uint32_t spk = pk;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, spk, off);
    if (lane >= (uint32_t)off) spk += a;
  }
  if (lane == 31) sc->wtot2[warp] = spk;
  __syncthreads();
  uint32_t before = spk - pk;
#pragma unroll
  for (uint32_t w = 0; w < W; ++w)
    if (w < warp) before += sc->wtot2[w];
  uint32_t gt_before = before >> 16, eq_before = before & 0xffffu;
  // this thread's selected keys occupy ranks [lo, hi): skip the id loop when the window misses them
  const uint32_t lo = gt_before + min(eq_before, need2);
  const uint32_t hi = gt_before + ngt + min(eq_before + neq, need2);
  if (hi > r0 && lo < r1) {
#pragma unroll
    for (uint32_t j = 0; j < (kRegVecs ? kRegVecs : nvec); ++j) {
      const uint4 v = vec(j);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int x = 0; x < 8; ++x) {
        const uint32_t key = (x & 1) ? (w[x >> 1] >> 16) : (w[x >> 1] & 0xffffu);
        const uint32_t pos = p0 + j * 8 + x;
        const bool in = pos < len;
        const bool gt = in && (all || key > tau), eq = in && !all && key == tau;
        if (gt || (eq && eq_before < need2)) {
          const uint32_t rank = gt_before + min(eq_before, need2);
          if (rank >= r0 && rank < r1) {
            const int32_t pid = __ldg(page_table + pos);
            s_pages[rank - r0] = pid;
            if (ids_out) ids_out[rank] = pid;
          }
        }
        gt_before += gt;
        eq_before += eq;
      }
    }
  }
}

template <int kThreads>
__device__ __forceinline__ void select_window(const uint16_t* __restrict__ keys, const uint32_t* __restrict__ hist,
                                              uint32_t len, uint32_t k, uint32_t r0, uint32_t r1,
                                              const int32_t* __restrict__ page_table, int32_t* __restrict__ s_pages,
                                              int32_t* __restrict__ ids_out, SelScratch<kThreads>* sc) {
  const uint32_t kpt = sel_keys_per_thread(len, kThreads);  // uniform over the CTA
  if (kpt == 16)
    select_window_impl<kThreads, 2>(keys, hist, len, k, r0, r1, page_table, s_pages, ids_out, sc);
  else
    select_window_impl<kThreads, 0>(keys, hist, len, k, r0, r1, page_table, s_pages, ids_out, sc);
}

}  // namespace quest
