// Paged causal prefill attention.
// Replaces BatchPrefillWithPagedKVCacheKernel (reference kernels/include/prefill/prefill.cuh:674-869,
// entry :1007-1115; shim quest/ops/csrc/batch_prefill.cu:27-117).  Query i of qo_len sits at position
// kv_len - qo_len + i and attends to kv positions <= its own (causal) or to all (non-causal).
// Two kernels:
//   prefill_umma_kernel  tcgen05 tensor cores: a CTA owns 128 queries of one head (one query per thread = one TMEM
//                        lane), streams 64-token K/V tiles of the paged cache into 128B-swizzled shared memory with
//                        cp.async (double buffered), S = Q K^T and O += P V run as tcgen05.mma with the
//                        accumulators in TMEM, the online softmax reads S with tcgen05.ld, one row per thread.
//   prefill_simt_kernel  CUDA cores, 16 queries per CTA: short query chunks (qo_len <= 16) and the A/B baseline.
#include <cstring>

#include "common.cuh"
#include "umma.cuh"

namespace quest {

constexpr int kPfThreads = 128;
constexpr int kPfBQ = 16;   // queries per CTA
constexpr int kPfBK = 32;   // kv tokens per smem tile
constexpr int kPfTPQ = kPfThreads / kPfBQ;  // 8 threads per query

template <int D>
__global__ void __launch_bounds__(kPfThreads)
prefill_simt_kernel(const __half* __restrict__ q, __half* __restrict__ o, uint32_t qo_len,
               const __half* __restrict__ kv, const int32_t* __restrict__ kv_indices, uint32_t kv_len,
               int causal, uint32_t Hq, uint32_t Hkv, uint32_t P, float sm_scale_log2) {
  constexpr int DPT = D / kPfTPQ;  // dims per thread
  __shared__ __align__(16) __half k_s[kPfBK][D];
  __shared__ __align__(16) __half v_s[kPfBK][D];
  pdl_wait();
  const uint32_t h = blockIdx.y, kvh = h / (Hq / Hkv);
  const uint32_t q0 = blockIdx.x * kPfBQ;
  const uint32_t tid = threadIdx.x, qi = tid / kPfTPQ, sub = tid % kPfTPQ;
  const uint32_t qidx = q0 + qi;
  const bool q_ok = qidx < qo_len;
  const uint32_t qpos = kv_len - qo_len + (q_ok ? qidx : 0);  // absolute position
  float qv[DPT], acc[DPT];
#pragma unroll
  for (int i = 0; i < DPT; ++i) {
    qv[i] = q_ok ? __half2float(q[((size_t)qidx * Hq + h) * D + sub * DPT + i]) * sm_scale_log2 : 0.f;
    acc[i] = 0.f;
  }
  float m = -1e30f, d = 0.f;
  const uint32_t last_q = min(q0 + kPfBQ, qo_len) - 1;
  const uint32_t kv_end = causal ? min(kv_len, kv_len - qo_len + last_q + 1) : kv_len;
  const size_t slab = (size_t)P * Hkv * D;

  for (uint32_t t0 = 0; t0 < kv_end; t0 += kPfBK) {
    __syncthreads();
    // stage a [kPfBK, D] tile of K and V: one 16-byte vector per thread per step
    for (uint32_t v = tid; v < kPfBK * (D / 8); v += kPfThreads) {
      const uint32_t r = v / (D / 8), c = v % (D / 8);
      const uint32_t t = t0 + r;
      H8 kk, vv;
      if (t < kv_end) {
        const size_t page = (size_t)__ldg(kv_indices + t / P);
        const __half* kp = kv + nhd_off(page, 0, t % P, kvh, P, Hkv, D) + c * 8;
        kk = ld_plain(kp);
        vv = ld_plain(kp + slab);
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) kk.h[i] = vv.h[i] = __float2half2_rn(0.f);
      }
      st_plain(&k_s[r][c * 8], kk);
      st_plain(&v_s[r][c * 8], vv);
    }
    __syncthreads();
    const uint32_t nt = min((uint32_t)kPfBK, kv_end - t0);
    for (uint32_t r = 0; r < nt; ++r) {
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < DPT; i += 2) {
        const float2 kf = __half22float2(*reinterpret_cast<const __half2*>(&k_s[r][sub * DPT + i]));
        s = fmaf(qv[i], kf.x, s);
        s = fmaf(qv[i + 1], kf.y, s);
      }
#pragma unroll
      for (int off = kPfTPQ / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
      const bool vis = q_ok && (!causal || t0 + r <= qpos);
      if (vis) {
        const float mn = fmaxf(m, s);
        const float sc = ex2(m - mn), p = ex2(s - mn);
        m = mn;
        d = d * sc + p;
#pragma unroll
        for (int i = 0; i < DPT; i += 2) {
          const float2 vf = __half22float2(*reinterpret_cast<const __half2*>(&v_s[r][sub * DPT + i]));
          acc[i] = fmaf(p, vf.x, acc[i] * sc);
          acc[i + 1] = fmaf(p, vf.y, acc[i + 1] * sc);
        }
      }
    }
  }
  if (q_ok) {
#pragma unroll
    for (int i = 0; i < DPT; ++i)
      o[((size_t)qidx * Hq + h) * D + sub * DPT + i] = __float2half_rn(d > 0.f ? __fdividef(acc[i], d) : 0.f);
  }
}

// ------------------------------------------------------------------------------------------------------------
// tcgen05 version
// ------------------------------------------------------------------------------------------------------------
constexpr int kUmBQ = 128;      // queries per CTA = TMEM lanes = threads
constexpr int kUmBK = 64;       // kv tokens per tile
constexpr int kUmThreads = 128;
constexpr uint32_t kUmTmemCols = 256;  // S: 64 fp32 columns, O: D fp32 columns
constexpr float kUmLazyLog2 = 8.f;     // O is rescaled only when a row's maximum grows by more than 2^8

__device__ unsigned int g_pf_timeouts;  // barrier waits that ran out (0 in a healthy run)

// Shared memory map (bytes from the 1024-aligned base).  Every operand block is [rows][64 halves] = rows x 128 B,
// 128B-swizzled; an operand with D = 128 columns is two such blocks.
template <int D>
struct UmSmem {
  static constexpr uint32_t kQ = (D / 64) * kUmBQ * 128;
  static constexpr uint32_t kTile = (D / 64) * kUmBK * 128;  // one K or V tile
  static constexpr uint32_t kP = kUmBQ * 128;
  static constexpr uint32_t off_q = 0, off_k = kQ, off_v = off_k + 2 * kTile, off_p = off_v + 2 * kTile;
  static constexpr uint32_t off_bar = off_p + kP, total = off_bar + 64;
};

// cp.async one [64 tokens, D] tile of K and of V (kv head kvh) into the swizzled blocks; tokens >= kv_end are zero.
template <int D>
__device__ __forceinline__ void um_load_kv_tile(uint32_t k_saddr, uint32_t v_saddr, const __half* __restrict__ kv,
                                                const int32_t* __restrict__ kv_indices, uint32_t t0, uint32_t kv_end,
                                                uint32_t kvh, uint32_t Hkv, uint32_t P, size_t slab, uint32_t tid) {
  constexpr uint32_t CPR = D / 8;  // 16-byte chunks per token row
#pragma unroll
  for (uint32_t i = 0; i < kUmBK * CPR / kUmThreads; ++i) {
    const uint32_t v = i * kUmThreads + tid, r = v / CPR, c = v % CPR;
    const uint32_t t = t0 + r;
    const bool ok = t < kv_end;
    const uint32_t tt = ok ? t : 0;  // keeps the (unused) source address inside the cache
    const size_t page = (size_t)__ldg(kv_indices + tt / P);
    const __half* kp = kv + nhd_off(page, 0, tt % P, kvh, P, Hkv, D) + c * 8;
    const uint32_t off = (c / 8) * (kUmBK * 128) + umma::sw128_off(r, c % 8);
    umma::cp_async16(k_saddr + off, kp, ok ? 16u : 0u);
    umma::cp_async16(v_saddr + off, kp + slab, ok ? 16u : 0u);
  }
}

template <int D>
__global__ void __launch_bounds__(kUmThreads, 2)
prefill_umma_kernel(const __half* __restrict__ q, __half* __restrict__ o, uint32_t qo_len,
                    const __half* __restrict__ kv, const int32_t* __restrict__ kv_indices, uint32_t kv_len,
                    int causal, uint32_t Hq, uint32_t Hkv, uint32_t P, float sm_scale_log2,
                    unsigned int* __restrict__ err_word) {
  using namespace umma;
  using L = UmSmem<D>;
  extern __shared__ __align__(1024) uint8_t um_smem[];
  const uint32_t sbase = smem_u32(um_smem);
  const uint32_t q_s = sbase + L::off_q, k_s = sbase + L::off_k, v_s = sbase + L::off_v, p_s = sbase + L::off_p;
  const uint32_t bar_s = sbase + L::off_bar, bar_o = bar_s + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(um_smem + L::off_bar + 16);

  const uint32_t tid = threadIdx.x, warp = tid / 32;
  const uint32_t qtile = gridDim.x - 1 - blockIdx.x;  // longest (latest) query tiles first
  const uint32_t q0 = qtile * kUmBQ, h = blockIdx.y, kvh = h / (Hq / Hkv);
  const uint32_t row = q0 + tid;
  const bool row_ok = row < qo_len;
  const uint32_t pos0 = kv_len - qo_len;  // position of query 0
  const uint32_t last_row = min(q0 + kUmBQ, qo_len) - 1;
  const uint32_t kv_end = causal ? pos0 + last_row + 1 : kv_len;
  const uint32_t ntiles = (kv_end + kUmBK - 1) / kUmBK;
  const uint32_t lim = (causal && row_ok) ? pos0 + row : kv_end - 1;  // last kv position this row sees
  const uint32_t mask_from = causal ? pos0 + q0 : kv_end - 1;         // tiles reaching past this need the mask
  const size_t slab = (size_t)P * Hkv * D;

  if (warp == 0) tmem_alloc<kUmTmemCols>(smem_u32(tmem_slot));
  if (tid == 0) {
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    mbar_init_fence();
  }
  pdl_wait();
  {  // Q tile: rows past qo_len are zero
    constexpr uint32_t CPR = D / 8;
#pragma unroll
    for (uint32_t i = 0; i < kUmBQ * CPR / kUmThreads; ++i) {
      const uint32_t v = i * kUmThreads + tid, r = v / CPR, c = v % CPR;
      const bool ok = q0 + r < qo_len;
      const __half* src = q + ((size_t)(ok ? q0 + r : 0) * Hq + h) * D + c * 8;
      cp_async16(q_s + (c / 8) * (kUmBQ * 128) + sw128_off(r, c % 8), src, ok ? 16u : 0u);
    }
  }
  um_load_kv_tile<D>(k_s, v_s, kv, kv_indices, 0, kv_end, kvh, Hkv, P, slab, tid);
  cp_async_commit();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = *tmem_slot;
  const uint32_t lane_addr = (warp * 32u) << 16;
  const uint32_t t_s = tbase + lane_addr, t_o = tbase + 64 + lane_addr;  // per-warp lane window
  constexpr uint32_t kIdescS = idesc_f16(kUmBQ, kUmBK, 0), kIdescO = idesc_f16(kUmBQ, D, 1);

  uint32_t sreg[kUmBK];
  float m_used = -INFINITY, dsum = 0.f;
  bool ok = (sbase & 1023u) == 0;

  for (uint32_t j = 0; j < ntiles; ++j) {
    const uint32_t b = j & 1, t0 = j * kUmBK;
    cp_async_wait<0>();
    fence_smem_to_async();
    __syncthreads();  // K/V tile j (and Q) are in shared memory for the tensor core; S of tile j-1 has been read
    if (tid == 0) {
      fence_after_sync();
#pragma unroll
      for (uint32_t kk = 0; kk < D / 16; ++kk) {
        const uint64_t a = desc_k_major(q_s + (kk / 4) * (kUmBQ * 128) + (kk % 4) * 32);
        const uint64_t bd = desc_k_major(k_s + b * L::kTile + (kk / 4) * (kUmBK * 128) + (kk % 4) * 32);
        mma_f16_ss(tbase, a, bd, kIdescS, kk > 0);
      }
      mma_commit(bar_s);
    }
    if (j > 0) {  // O += P V of tile j-1 is done: its V buffer and the P tile are free, O may be rescaled
      ok &= mbar_wait(bar_o, (j - 1) & 1);
      fence_after_sync();
    }
    if (j + 1 < ntiles)
      um_load_kv_tile<D>(k_s + (b ^ 1) * L::kTile, v_s + (b ^ 1) * L::kTile, kv, kv_indices, t0 + kUmBK, kv_end, kvh,
                         Hkv, P, slab, tid);
    cp_async_commit();
    ok &= mbar_wait(bar_s, j & 1);
    fence_after_sync();
    __syncwarp();
    tmem_ld32(t_s, sreg);
    tmem_ld32(t_s + 32, sreg + 32);
    tmem_ld_wait();

    float tmax = -INFINITY;
    if (t0 + kUmBK - 1 > mask_from) {
#pragma unroll
      for (int i = 0; i < kUmBK; ++i) {
        const float s = (t0 + i <= lim) ? __uint_as_float(sreg[i]) : -INFINITY;
        sreg[i] = __float_as_uint(s);
        tmax = fmaxf(tmax, s);
      }
    } else {
#pragma unroll
      for (int i = 0; i < kUmBK; ++i) tmax = fmaxf(tmax, __uint_as_float(sreg[i]));
    }
    // The running maximum is only raised when it would grow by more than 2^8: p <= 256 stays exact enough in
    // fp16, and the accumulator in TMEM is touched rarely.  d and O use the same (stale) maximum, so O/d is exact.
    float sc = 1.f;
    const bool grow = (tmax - m_used) * sm_scale_log2 > kUmLazyLog2;
    if (grow) {
      sc = ex2((m_used - tmax) * sm_scale_log2);
      m_used = tmax;
      dsum *= sc;
    }
    const float mneg = -m_used * sm_scale_log2;
    float ps[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < kUmBK / 8; ++c) {
      uint32_t w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float p0 = ex2(fmaf(__uint_as_float(sreg[c * 8 + 2 * i]), sm_scale_log2, mneg));
        const float p1 = ex2(fmaf(__uint_as_float(sreg[c * 8 + 2 * i + 1]), sm_scale_log2, mneg));
        ps[i] += p0 + p1;
        const __half2 hp = __floats2half2_rn(p0, p1);
        w[i] = *reinterpret_cast<const uint32_t*>(&hp);
      }
      asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" :: "r"(p_s + sw128_off(tid, c)), "r"(w[0]), "r"(w[1]),
                   "r"(w[2]), "r"(w[3]) : "memory");
    }
    dsum += (ps[0] + ps[1]) + (ps[2] + ps[3]);
    if (j > 0 && __any_sync(0xffffffffu, grow)) {
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
    fence_before_sync();
    fence_smem_to_async();
    __syncthreads();  // P is in shared memory, every thread's TMEM reads/writes are done
    if (tid == 0) {
      fence_after_sync();
#pragma unroll
      for (uint32_t kk = 0; kk < kUmBK / 16; ++kk) {
        const uint64_t a = desc_k_major(p_s + kk * 32);
        const uint64_t bd = desc_mn_major(v_s + b * L::kTile + kk * 2048, kUmBK * 128);
        mma_f16_ss(tbase + 64, a, bd, kIdescO, (j | kk) > 0);
      }
      mma_commit(bar_o);
    }
  }

  ok &= mbar_wait(bar_o, (ntiles - 1) & 1);
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
  if (!ok) {  // a wait ran out: o is not valid.  Sticky, host-visible (quest_device_error)
    atomicOr(&g_pf_timeouts, 1u);
    device_error_raise(err_word, QUEST_DEVERR_PREFILL);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_free<kUmTmemCols>(tbase);
}

// prefill_tma.cu: the warp-specialised TMA + tcgen05 kernel (page sizes 2^n >= 8)
int launch_prefill_tma(const __half* q, __half* o, uint32_t qo_len, const __half* kv, const int32_t* kv_indices,
                       uint32_t num_pages, uint32_t kv_len, int causal, uint32_t Hq, uint32_t Hkv, uint32_t D, uint32_t P,
                       float sc, cudaStream_t s);
unsigned int prefill_tma_timeouts();

// QUEST_DBG_PREFILL_IMPL (quest_debug_set_option; test / A-B hook): 0 = by shape (TMA kernel where it applies, else the
// cp.async tcgen05 kernel; SIMT for qo_len <= 16), 1 = SIMT, 2 = cp.async tcgen05 kernel, 3 = TMA kernel (no fallback)
static int prefill_impl_override() { return debug_option(QUEST_DBG_PREFILL_IMPL); }

template <int D>
static cudaError_t launch_prefill_umma(const __half* q, __half* o, uint32_t qo_len, const __half* kv,
                                       const int32_t* kv_indices, uint32_t kv_len, int causal, uint32_t Hq,
                                       uint32_t Hkv, uint32_t P, float sc, cudaStream_t s) {
  static const cudaError_t attr = cudaFuncSetAttribute(prefill_umma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                       (int)UmSmem<D>::total);
  if (attr != cudaSuccess) return attr;
  dim3 grid((qo_len + kUmBQ - 1) / kUmBQ, Hq);
  prefill_umma_kernel<D><<<grid, kUmThreads, UmSmem<D>::total, s>>>(q, o, qo_len, kv, kv_indices, kv_len, causal, Hq, Hkv,
                                                                    P, sc, device_error_word());
  return cudaSuccess;
}

}  // namespace quest

// Test hook: barrier waits of the tcgen05 prefill kernel that timed out since the library was loaded (0 = healthy).
extern "C" int quest_debug_prefill_timeouts(void) {
  unsigned int v = 0;
  if (cudaMemcpyFromSymbol(&v, quest::g_pf_timeouts, sizeof(v)) != cudaSuccess) return -1;
  return (int)(v | (quest::prefill_tma_timeouts() << 1));
}

extern "C" int quest_prefill(const void* q, void* o, uint32_t qo_len, const void* kv_data, const int32_t* kv_indices,
                             uint32_t num_kv_pages, uint32_t kv_last_page_len, int causal, uint32_t num_qo_heads,
                             uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size, void* stream) {
  if (!q || !o || !kv_data || !kv_indices || !num_qo_heads || !num_kv_heads || !page_size || !num_kv_pages)
    return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (kv_last_page_len < 1 || kv_last_page_len > page_size) return QUEST_ERR_BAD_ARG;
  const uint64_t kv_len64 = (uint64_t)(num_kv_pages - 1) * page_size + kv_last_page_len;
  if (qo_len == 0) return QUEST_OK;
  if (qo_len > kv_len64) return QUEST_ERR_BAD_ARG;
  if (quest::device_error_peek() & QUEST_DEVERR_PREFILL) return QUEST_ERR_DEVICE;  // an earlier prefill gave up on a wait
  const uint32_t kv_len = (uint32_t)kv_len64;
  const float sc = quest::kLog2e / sqrtf((float)head_dim);
  cudaStream_t s = (cudaStream_t)stream;
  const int force = quest::prefill_impl_override();
  if (force == 3 || (force == 0 && qo_len > (uint32_t)quest::kPfBQ)) {
    const int rc = quest::launch_prefill_tma((const __half*)q, (__half*)o, qo_len, (const __half*)kv_data, kv_indices, num_kv_pages,
                                             kv_len, causal, num_qo_heads, num_kv_heads, head_dim, page_size, sc, s);
    if (rc == QUEST_OK) QUEST_RETURN_LAUNCH();
    if (rc != QUEST_ERR_UNSUPPORTED || force == 3) return rc;
  }
  const bool use_umma = force == 2 || (force == 0 && qo_len > (uint32_t)quest::kPfBQ);
  if (use_umma) {
    cudaError_t e;
    switch (head_dim) {
      case 64:
        e = quest::launch_prefill_umma<64>((const __half*)q, (__half*)o, qo_len, (const __half*)kv_data, kv_indices, kv_len,
                                           causal, num_qo_heads, num_kv_heads, page_size, sc, s);
        break;
      case 128:
        e = quest::launch_prefill_umma<128>((const __half*)q, (__half*)o, qo_len, (const __half*)kv_data, kv_indices, kv_len,
                                            causal, num_qo_heads, num_kv_heads, page_size, sc, s);
        break;
      default: return QUEST_ERR_UNSUPPORTED;
    }
    if (e != cudaSuccess) return (int)e;
    QUEST_RETURN_LAUNCH();
  }
  dim3 grid((qo_len + quest::kPfBQ - 1) / quest::kPfBQ, num_qo_heads);
#define QUEST_PF(DD)                                                                                                   \
  quest::prefill_simt_kernel<DD><<<grid, quest::kPfThreads, 0, s>>>((const __half*)q, (__half*)o, qo_len,              \
                                                                    (const __half*)kv_data, kv_indices, kv_len, causal, \
                                                                    num_qo_heads, num_kv_heads, page_size, sc)
  switch (head_dim) {
    case 64: QUEST_PF(64); break;
    case 128: QUEST_PF(128); break;
    default: return QUEST_ERR_UNSUPPORTED;
  }
#undef QUEST_PF
  QUEST_RETURN_LAUNCH();
}
