// Paged causal prefill attention, warp-specialised tcgen05 + TMA version (sm_100a).
// Replaces BatchPrefillWithPagedKVCacheKernel (reference kernels/include/prefill/prefill.cuh:674-869) for page sizes
// that are a power of two >= 8 (the reference accepts 1, 8, 16, 32: prefill.cuh:1031-1035); other shapes stay on the
// cp.async kernel in prefill.cu.
//
// One CTA = 128 queries of one head, eight warps in five roles.  Warps 0-3 (one query row per thread = one TMEM
// lane) run the online softmax; lane 0 of warp 7 issues the S = Q K^T MMAs, lane 0 of warp 4 the O += P V MMAs; warp 5
// loads Q and the K tiles, warp 6 the V tiles (TMA: lane l sends the boxes of the tile's l-th page).  Each role runs its own in-order loop and they meet
// only on mbarriers, so TMA issue latency, MMA issue latency and the softmax arithmetic overlap.
//   * K/V pages come straight from the paged pool through ONE 5-d tensor map [page][k|v][slot][head][dim] with
//     128-byte swizzle: a box is (64 dims, 1 head, min(P,64) slots) of one page, so a 64-token tile of K (or V) is
//     64/P x D/64 boxes that land in the layout tcgen05 reads, no thread touches the data or computes an address.
//   * S = Q K^T is double-buffered in TMEM (2 x 64 columns) and issued two tiles ahead, O += P V accumulates in TMEM
//     (D columns); the tensor pipe runs S(j+2) and PV(j) while the softmax warps work on tile j+1.
//   * K stages are released by the completion of S, V stages and the P tile by the completion of P V (tcgen05.commit
//     -> mbarrier), the softmax warps hand P over with a 128-arrival mbarrier: no CTA-wide barrier in the loop.
//   * softmax: lazy maximum (O in TMEM is rescaled only when a row maximum grows by more than 2^8), FMNMX3 maximum,
//     packed FFMA2 / FADD2 for the scale-subtract and the row sum.
#include <cuda.h>

#include "common.cuh"
#include "umma.cuh"

namespace quest {

constexpr int kTmBQ = 128;
constexpr int kTmBK = 64;
constexpr int kTmThreads = 256;        // 4 softmax warps + PV-MMA warp + K-loader + V-loader + S-MMA warp
constexpr uint32_t kTmTmemCols = 256;  // S0 [0,64), S1 [64,128), O [128, 128+D)
constexpr float kTmLazyLog2 = 8.f;

__device__ unsigned int g_pf_tma_timeouts;
__device__ unsigned int g_pf_tma_first[4];  // first wait that ran out: barrier slot, parity, thread, block (x | y << 16)

#ifdef QUEST_PF_CLOCKS  // per-iteration timeline of one CTA (tools/prefill_timeline.py); not in the product build
__device__ unsigned long long g_pf_clk[64][16];
#define PF_STAMP(j, slot)                                                             \
  do {                                                                                \
    if (blockIdx.x == gridDim.x / 2 && blockIdx.y == 0 && (j) < 64) g_pf_clk[(j)][(slot)] = clock64(); \
  } while (0)
#else
#define PF_STAMP(j, slot) do {} while (0)
#endif

template <int D>
struct TmSmem {
  static constexpr uint32_t kQ = (D / 64) * kTmBQ * 128;
  static constexpr uint32_t kTile = (D / 64) * kTmBK * 128;
  static constexpr uint32_t kP = kTmBQ * 128;
  static constexpr uint32_t off_q = 0, off_k = kQ, off_v = off_k + 2 * kTile, off_p = off_v + 2 * kTile;
  static constexpr uint32_t off_bar = off_p + kP, total = off_bar + 192;  // 16 barriers, TMEM slot, abort flag
};
// barrier slots (8 bytes each) from off_bar.  kBarPV completes once per tile and may only be waited on by a thread
// that is at most ONE phase behind it (the parity test cannot tell phase c from phase c+2); kBarDone is the one-shot
// "every PV has completed" signal the epilogue waits on, whatever the skew.
enum : uint32_t { kBarQ = 0, kBarK = 1, kBarV = 3, kBarS = 5, kBarPV = 7, kBarVFree = 9, kBarSFree = 11, kBarP = 13, kBarDone = 15, kNumBars = 16 };

// Page id of this lane's box of the 64-token tile at t0 (lane l < boxes per tile = 64 / min(P,64)), or -1 when the
// tile has no such page.  Loaded one or more iterations before it is used, so the latency is off the issue path.
__device__ __forceinline__ int tm_box_page(const int32_t* __restrict__ kv_indices, uint32_t num_pages, uint32_t t0,
                                           uint32_t p_shift, uint32_t lane, uint32_t kv_end) {
  const uint32_t nbox = p_shift < 6 ? (64u >> p_shift) : 1u;
  const uint32_t page_i = (t0 >> p_shift) + (p_shift < 6 ? lane : 0u);
  return (lane < nbox && t0 < kv_end && page_i < num_pages) ? __ldg(kv_indices + page_i) : -1;
}

// One 64-token tile of K (slab 0) or V (slab 1) -> one stage, issued by the whole producer warp: lane l sends the
// boxes of page `page` (its box of the tile), lane 0 posts the byte count.  Pages past the table are skipped (their
// rows keep whatever the stage held: K rows are masked by the softmax, V rows are zeroed by it).
template <int D>
__device__ __forceinline__ void tm_issue_tile(const CUtensorMap* kvmap, uint32_t dst, uint32_t bar, int page, uint32_t t0,
                                              uint32_t kvh, uint32_t slab, uint32_t p_shift, uint32_t lane) {
  const uint32_t box_shift = p_shift < 6 ? p_shift : 6;  // tokens per box = min(P, 64)
  const uint32_t have = __popc(__ballot_sync(0xffffffffu, page >= 0));
  if (lane == 0) umma::mbar_arrive_expect_tx(bar, (have << box_shift) * 128u * (D / 64));
  if (page >= 0) {
    const int entry = (int)(t0 & ((1u << p_shift) - 1));  // non-zero only when a page holds several tiles
#pragma unroll
    for (uint32_t nb = 0; nb < D / 64; ++nb)
      umma::tma_load_5d(dst + nb * (kTmBK * 128) + ((lane << box_shift) * 128u), kvmap, bar, (int)(nb * 64), (int)kvh, entry,
                        (int)slab, page);
  }
}

template <int D, bool kPTmem>
__global__ void __launch_bounds__(kTmThreads, 2)
prefill_tma_kernel(const __grid_constant__ CUtensorMap qmap, const __grid_constant__ CUtensorMap kvmap,
                   __half* __restrict__ o, uint32_t qo_len, const int32_t* __restrict__ kv_indices, uint32_t num_pages,
                   uint32_t kv_len, int causal, uint32_t Hq, uint32_t Hkv, uint32_t p_shift, float sm_scale_log2,
                   unsigned int* __restrict__ err_word) {
  using namespace umma;
  using L = TmSmem<D>;
  extern __shared__ __align__(1024) uint8_t tm_smem[];
  const uint32_t sbase = smem_u32(tm_smem);
  const uint32_t q_s = sbase + L::off_q, k_s = sbase + L::off_k, v_s = sbase + L::off_v, p_s = sbase + L::off_p;
  const uint32_t bars = sbase + L::off_bar;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tm_smem + L::off_bar + kNumBars * 8);
  volatile uint32_t* abort_flag = reinterpret_cast<volatile uint32_t*>(tm_smem + L::off_bar + kNumBars * 8 + 4);
  auto bar = [&](uint32_t i) { return bars + i * 8; };
  auto wait = [&](uint32_t slot, uint32_t parity) {
    if (!mbar_wait_abortable(bars + slot * 8, parity, abort_flag) && atomicCAS(&g_pf_tma_first[3], 0u, 0x80000000u | blockIdx.x | (blockIdx.y << 16)) == 0u) {
      g_pf_tma_first[0] = slot;
      g_pf_tma_first[1] = parity;
      g_pf_tma_first[2] = threadIdx.x;
    }
  };

  const uint32_t tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
  const uint32_t qtile = gridDim.x - 1 - blockIdx.x;  // longest (latest) query tiles first
  const uint32_t q0 = qtile * kTmBQ, h = blockIdx.y, kvh = h / (Hq / Hkv);
  const uint32_t pos0 = kv_len - qo_len;
  const uint32_t last_row = min(q0 + kTmBQ, qo_len) - 1;
  const uint32_t kv_end = causal ? pos0 + last_row + 1 : kv_len;
  const uint32_t ntiles = (kv_end + kTmBK - 1) / kTmBK;

  if (warp == 4) tmem_alloc<kTmTmemCols>(smem_u32(tmem_slot));
  if (tid == 0) {
    for (uint32_t i = 0; i < kNumBars; ++i) mbar_init(bar(i), (i >= kBarSFree && i < kBarDone) ? 128u : 1u);
    *abort_flag = (sbase & 1023u) ? 1u : 0u;
    mbar_init_fence();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = *tmem_slot;
  constexpr uint32_t kIdescS = idesc_f16(kTmBQ, kTmBK, 0), kIdescO = idesc_f16(kTmBQ, D, 1);
  pdl_wait();

  if (warp == 4) {
    // ------------------------------------------------------------------ PV warp: lane 0 issues O += P(j) V(j)
    if (lane == 0) {
      for (uint32_t j = 0; j < ntiles; ++j) {
        const uint32_t b = j & 1, ph = (j >> 1) & 1;
        PF_STAMP(j, 2);
        wait(kBarP + b, ph);    // P(j) has been handed over
        PF_STAMP(j, 3);
        wait(kBarV + b, ph);    // V(j) has landed
        PF_STAMP(j, 4);
        fence_after_sync();
#pragma unroll
        for (uint32_t kk = 0; kk < kTmBK / 16; ++kk) {
          const uint64_t bd = desc_mn_major(v_s + b * L::kTile + kk * 2048, kTmBK * 128);
          if constexpr (kPTmem)
            mma_f16_ts(tbase + 128, tbase + b * 64 + kk * 8, bd, kIdescO, (j | kk) > 0);
          else
            mma_f16_ss(tbase + 128, desc_k_major(p_s + kk * 32), bd, kIdescO, (j | kk) > 0);
        }
        mma_commit(bar(kBarPV));         // softmax warps: the P tile is free, O is up to date
        mma_commit(bar(kBarVFree + b));  // V loader: the stage is free
        PF_STAMP(j, 5);
      }
      mma_commit(bar(kBarDone));  // one-shot: O holds every tile
    }
    __syncwarp();
  } else if (warp == 7) {
    // ------------------------------------------------------------------ S warp: lane 0 issues S(j) = Q K(j)^T, two tiles ahead
    if (lane == 0) {
      wait(kBarQ, 0);
      for (uint32_t j = 0; j < ntiles; ++j) {
        const uint32_t b = j & 1, ph = (j >> 1) & 1;
        // the buffer is free once S(j-2) has been read out of it, or, with P(j-2) stored over it, once PV(j-2) is done
        if (j >= 2) wait((kPTmem ? kBarVFree : kBarSFree) + b, ph ^ 1);
        wait(kBarK + b, ph);                      // K(j) has landed
        PF_STAMP(j, 6);
        fence_after_sync();
#pragma unroll
        for (uint32_t kk = 0; kk < D / 16; ++kk) {
          const uint64_t a = desc_k_major(q_s + (kk / 4) * (kTmBQ * 128) + (kk % 4) * 32);
          const uint64_t bd = desc_k_major(k_s + b * L::kTile + (kk / 4) * (kTmBK * 128) + (kk % 4) * 32);
          mma_f16_ss(tbase + b * 64, a, bd, kIdescS, kk > 0);
        }
        mma_commit(bar(kBarS + b));  // S(j) is ready for the softmax warps, and its K stage is free for the loader
        PF_STAMP(j, 7);
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // ------------------------------------------------------------------ K loader warp (and Q): lane l sends box l
    int pg = tm_box_page(kv_indices, num_pages, 0, p_shift, lane, kv_end);
    int pg1 = tm_box_page(kv_indices, num_pages, kTmBK, p_shift, lane, kv_end);
    int pg2 = tm_box_page(kv_indices, num_pages, 2 * kTmBK, p_shift, lane, kv_end);
    if (lane == 0) {
      tma_prefetch_desc(&qmap);
      tma_prefetch_desc(&kvmap);
      mbar_arrive_expect_tx(bar(kBarQ), L::kQ);
#pragma unroll
      for (uint32_t nb = 0; nb < D / 64; ++nb)
        tma_load_3d(q_s + nb * (kTmBQ * 128), &qmap, bar(kBarQ), (int)(nb * 64), (int)h, (int)q0);
    }
    tm_issue_tile<D>(&kvmap, k_s, bar(kBarK), pg, 0, kvh, 0, p_shift, lane);
    if (ntiles > 1) tm_issue_tile<D>(&kvmap, k_s + L::kTile, bar(kBarK + 1), pg1, kTmBK, kvh, 0, p_shift, lane);
    for (uint32_t j = 0; j + 2 < ntiles; ++j) {  // K(j+2) takes the stage of K(j) once S(j) is done
      const int pg3 = tm_box_page(kv_indices, num_pages, (j + 3) * kTmBK, p_shift, lane, kv_end);
      wait(kBarS + (j & 1), (j >> 1) & 1);
      if (lane == 0) PF_STAMP(j, 0);
      tm_issue_tile<D>(&kvmap, k_s + (j & 1) * L::kTile, bar(kBarK + (j & 1)), pg2, (j + 2) * kTmBK, kvh, 0, p_shift, lane);
      pg2 = pg3;
    }
  } else if (warp == 6) {
    // ------------------------------------------------------------------ V loader warp
    int pg = tm_box_page(kv_indices, num_pages, 0, p_shift, lane, kv_end);
    int pg1 = tm_box_page(kv_indices, num_pages, kTmBK, p_shift, lane, kv_end);
    int pg2 = tm_box_page(kv_indices, num_pages, 2 * kTmBK, p_shift, lane, kv_end);
    tm_issue_tile<D>(&kvmap, v_s, bar(kBarV), pg, 0, kvh, 1, p_shift, lane);
    if (ntiles > 1) tm_issue_tile<D>(&kvmap, v_s + L::kTile, bar(kBarV + 1), pg1, kTmBK, kvh, 1, p_shift, lane);
    for (uint32_t j = 0; j + 2 < ntiles; ++j) {  // V(j+2) takes the stage of V(j) once PV(j) is done
      const int pg3 = tm_box_page(kv_indices, num_pages, (j + 3) * kTmBK, p_shift, lane, kv_end);
      wait(kBarVFree + (j & 1), (j >> 1) & 1);
      if (lane == 0) PF_STAMP(j, 1);
      tm_issue_tile<D>(&kvmap, v_s + (j & 1) * L::kTile, bar(kBarV + (j & 1)), pg2, (j + 2) * kTmBK, kvh, 1, p_shift, lane);
      pg2 = pg3;
    }
  } else {
    // ------------------------------------------------------------------ softmax warps: row tid of the tile
    const uint32_t row = q0 + tid;
    const bool row_ok = row < qo_len;
    const uint32_t lim = (causal && row_ok) ? pos0 + row : kv_end - 1;  // last kv position this row sees
    const uint32_t mask_from = causal ? pos0 + q0 : kv_end - 1;         // tiles reaching past this need the mask
    const uint32_t lane_addr = (warp * 32u) << 16;
    const uint32_t t_o = tbase + 128 + lane_addr;
    const uint64_t scale2 = pack_f32x2(sm_scale_log2, sm_scale_log2);
    uint32_t sreg[kTmBK];
    float m_used = -INFINITY, dsum = 0.f;

    for (uint32_t j = 0; j < ntiles; ++j) {
      const uint32_t b = j & 1, t0 = j * kTmBK;
      if (tid == 0) PF_STAMP(j, 8);
      wait(kBarS + b, (j >> 1) & 1);
      if (tid == 0) PF_STAMP(j, 9);
      fence_after_sync();
      __syncwarp();
      tmem_ld32(tbase + b * 64 + lane_addr, sreg);
      tmem_ld32(tbase + b * 64 + 32 + lane_addr, sreg + 32);
      tmem_ld_wait();
      if constexpr (!kPTmem) {
        fence_before_sync();
        mbar_arrive(bar(kBarSFree + b));  // the S warp may overwrite this buffer with S(j+2)
      }

      if (t0 + kTmBK - 1 > mask_from) {
#pragma unroll
        for (int i = 0; i < kTmBK; ++i)
          if (t0 + i > lim) sreg[i] = 0xff800000u;  // -inf
      }
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < kTmBK; i += 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c) mx[c] = max3(mx[c], __uint_as_float(sreg[i + 2 * c]), __uint_as_float(sreg[i + 2 * c + 1]));
      }
      const float tmax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      // lazy maximum: p <= 2^8 stays exact enough in fp16, d and O share the same (stale) maximum so O/d is exact
      float sc = 1.f;
      const bool grow = (tmax - m_used) * sm_scale_log2 > kTmLazyLog2;
      if (grow) {
        sc = ex2((m_used - tmax) * sm_scale_log2);
        m_used = tmax;
        dsum *= sc;
      }
      const float mneg = -m_used * sm_scale_log2;
      const uint64_t mneg2 = pack_f32x2(mneg, mneg);
      uint64_t ps[4] = {0ull, 0ull, 0ull, 0ull};
      uint32_t pw[kTmBK / 2];
#pragma unroll
      for (int i = 0; i < kTmBK / 2; ++i) {
        const uint64_t x = fma_f32x2(pack_f32x2(__uint_as_float(sreg[2 * i]), __uint_as_float(sreg[2 * i + 1])), scale2, mneg2);
        float x0, x1;
        unpack_f32x2(x, x0, x1);
        const float p0 = ex2(x0), p1 = ex2(x1);
        ps[i & 3] = add_f32x2(ps[i & 3], pack_f32x2(p0, p1));
        const __half2 hp = __floats2half2_rn(p0, p1);
        pw[i] = *reinterpret_cast<const uint32_t*>(&hp);
      }
      {
        float a0, a1;
        unpack_f32x2(add_f32x2(add_f32x2(ps[0], ps[1]), add_f32x2(ps[2], ps[3])), a0, a1);
        dsum += a0 + a1;
      }
      if (tid == 0) PF_STAMP(j, 10);
      if (j > 0) {
        // O may be rescaled (and, with P in shared memory, the P tile rewritten) once PV(j-1) is done; with P in
        // tensor memory only the rare rescale needs that wait
        const bool any_grow = __any_sync(0xffffffffu, grow);
        if (!kPTmem || any_grow) {
          wait(kBarPV, (j - 1) & 1);
          fence_after_sync();
          __syncwarp();
        }
        if (tid == 0) PF_STAMP(j, 11);
        if (any_grow) {
#pragma unroll 1
          for (uint32_t c0 = 0; c0 < D; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(t_o + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * sc);
            tmem_st32(t_o + c0, r);
          }
          tmem_st_wait();
        }
      }
      if constexpr (kPTmem) {  // P(j) over the first half of S(j)'s own columns, this thread's lane
        tmem_st32(tbase + b * 64 + lane_addr, pw);
        tmem_st_wait();
      } else {
#pragma unroll
        for (int c = 0; c < kTmBK / 8; ++c)
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" :: "r"(p_s + sw128_off(tid, c)), "r"(pw[4 * c]), "r"(pw[4 * c + 1]),
                       "r"(pw[4 * c + 2]), "r"(pw[4 * c + 3]) : "memory");
      }
      if (j + 1 == ntiles && kv_end < t0 + kTmBK) {  // rows of V past the end of the sequence must be zero, not stale
        wait(kBarV + b, (j >> 1) & 1);
        if (tid < kTmBK && t0 + tid >= kv_end) {
#pragma unroll
          for (uint32_t nb = 0; nb < D / 64; ++nb)
#pragma unroll
            for (uint32_t c = 0; c < 8; ++c)
              asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" :: "r"(v_s + b * L::kTile + nb * (kTmBK * 128) + tid * 128 + c * 16),
                           "r"(0u) : "memory");
        }
      }
      if (!kPTmem || (j + 1 == ntiles && kv_end < t0 + kTmBK)) fence_smem_to_async();
      fence_before_sync();
      mbar_arrive(bar(kBarP + b));  // one barrier per S/P buffer: a warp that runs a tile ahead must not add to tile j's count
      if (tid == 0) PF_STAMP(j, 12);
    }

    // NOT kBarPV: with P in tensor memory this warp may not have waited on it for many tiles, the barrier can be
    // two phases behind (only PV(ntiles-3) is implied by S(ntiles-1)) and its parity then reads as "done"
    wait(kBarDone, 0);
    fence_after_sync();
    __syncwarp();
    const float inv = dsum > 0.f ? __fdividef(1.f, dsum) : 0.f;
    __half* orow = o + ((size_t)(row_ok ? row : 0) * Hq + h) * D;
#pragma unroll 1
    for (uint32_t c0 = 0; c0 < D; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(t_o + c0, r);
      tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t w[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const __half2 hp = __floats2half2_rn(__uint_as_float(r[c * 8 + 2 * i]) * inv, __uint_as_float(r[c * 8 + 2 * i + 1]) * inv);
            w[i] = *reinterpret_cast<const uint32_t*>(&hp);
          }
          *reinterpret_cast<uint4*>(orow + c0 + c * 8) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (tid == 0 && *abort_flag) {  // a wait ran out: o is not valid.  Sticky, host-visible (quest_device_error)
    atomicOr(&g_pf_tma_timeouts, 1u);
    device_error_raise(err_word, QUEST_DEVERR_PREFILL);
  }
  if (warp == 4) tmem_free<kTmTmemCols>(tbase);  // the allocating warp frees
}

// ---- host side: tensor maps through the driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

static bool p_in_tmem() { return debug_option(QUEST_DBG_PREFILL_P_SMEM) == 0; }  // test / A-B hook: P through shared memory

template <int D, bool kPTmem>
static int launch_tma(const __half* q, __half* o, uint32_t qo_len, const __half* kv, const int32_t* kv_indices,
                      uint32_t num_pages, uint32_t kv_len, int causal, uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t p_shift,
                      float sc, cudaStream_t s) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return QUEST_ERR_UNSUPPORTED;
  static const cudaError_t attr = cudaFuncSetAttribute(prefill_tma_kernel<D, kPTmem>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                       (int)TmSmem<D>::total);
  if (attr != cudaSuccess) return (int)attr;
  CUtensorMap qmap, kvmap;
  const cuuint32_t ones[5] = {1, 1, 1, 1, 1};
  {
    const cuuint64_t dims[3] = {(cuuint64_t)D, Hq, qo_len};
    const cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)Hq * D * 2};
    const cuuint32_t box[3] = {64, 1, (cuuint32_t)kTmBQ};
    if (enc(&qmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, (void*)q, dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return QUEST_ERR_UNSUPPORTED;
  }
  {
    // the pool's page count is not part of the operator's arguments: the outermost extent only bounds the
    // hardware's range check, so it is set to the largest page id an int32 table can name at this page stride
    const cuuint64_t page_bytes = (cuuint64_t)2 * P * Hkv * D * 2;
    cuuint64_t max_pages = ((cuuint64_t)1 << 46) / page_bytes;
    if (max_pages > 0x7fffffffull) max_pages = 0x7fffffffull;
    const cuuint64_t dims[5] = {(cuuint64_t)D, Hkv, P, 2, max_pages};
    const cuuint64_t strides[4] = {(cuuint64_t)D * 2, (cuuint64_t)Hkv * D * 2, (cuuint64_t)P * Hkv * D * 2, page_bytes};
    const cuuint32_t box[5] = {64, 1, P < 64 ? P : 64u, 1, 1};
    if (enc(&kvmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, (void*)kv, dims, strides, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return QUEST_ERR_UNSUPPORTED;
  }
  dim3 grid((qo_len + kTmBQ - 1) / kTmBQ, Hq);
  // programmatic dependent launch: barrier init and tensor-memory allocation run before griddepcontrol.wait, i.e. under the tail
  // of the launch before this one (the RoPE + append producer in quest_prefill_layer_step); everything that reads memory is after it
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kTmThreads);
  cfg.dynamicSmemBytes = TmSmem<D>::total;
  cfg.stream = s;
  cudaLaunchAttribute lattr[1];
  lattr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  lattr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = lattr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, prefill_tma_kernel<D, kPTmem>, qmap, kvmap, o, qo_len, kv_indices, num_pages, kv_len,
                                           causal, Hq, Hkv, p_shift, sc, device_error_word());
  return e == cudaSuccess ? QUEST_OK : (int)e;
}

// QUEST_OK: launched; QUEST_ERR_UNSUPPORTED: shape not covered here (the caller falls back); else a CUDA error code.
int launch_prefill_tma(const __half* q, __half* o, uint32_t qo_len, const __half* kv, const int32_t* kv_indices,
                       uint32_t num_pages, uint32_t kv_len, int causal, uint32_t Hq, uint32_t Hkv, uint32_t D, uint32_t P,
                       float sc, cudaStream_t s) {
  if (P < 8 || (P & (P - 1)) || P > 256) return QUEST_ERR_UNSUPPORTED;
  if (((uintptr_t)q | (uintptr_t)kv) & 15) return QUEST_ERR_UNSUPPORTED;
  uint32_t p_shift = 0;
  while ((1u << p_shift) < P) ++p_shift;
#define QUEST_TMA_ARGS q, o, qo_len, kv, kv_indices, num_pages, kv_len, causal, Hq, Hkv, P, p_shift, sc, s
  const bool pt = p_in_tmem();
  switch (D) {
    case 64: return pt ? launch_tma<64, true>(QUEST_TMA_ARGS) : launch_tma<64, false>(QUEST_TMA_ARGS);
    case 128: return pt ? launch_tma<128, true>(QUEST_TMA_ARGS) : launch_tma<128, false>(QUEST_TMA_ARGS);
    default: return QUEST_ERR_UNSUPPORTED;
  }
#undef QUEST_TMA_ARGS
}

#ifdef QUEST_PF_CLOCKS
extern "C" int quest_debug_prefill_clocks(unsigned long long* out) {
  return (int)cudaMemcpyFromSymbol(out, g_pf_clk, sizeof(g_pf_clk));
}
#endif

extern "C" int quest_debug_prefill_first_timeout(unsigned int* out4) {
  return (int)cudaMemcpyFromSymbol(out4, g_pf_tma_first, 16);
}

unsigned int prefill_tma_timeouts() {
  unsigned int v = 0;
  if (cudaMemcpyFromSymbol(&v, g_pf_tma_timeouts, sizeof(v)) != cudaSuccess) return ~0u;
  return v;
}

}  // namespace quest
