// Paged causal prefill attention, v1 (SIMT, fp32 accumulate).
// Replaces BatchPrefillWithPagedKVCacheKernel (reference kernels/include/prefill/prefill.cuh:674-869,
// entry :1007-1115; shim quest/ops/csrc/batch_prefill.cu:27-117).  Query i of qo_len sits at position
// kv_len - qo_len + i and attends to kv positions <= its own (causal) or to all (non-causal).
// This version stages K/V tiles of the paged cache in shared memory and runs an online softmax on CUDA
// cores; it is the correctness baseline for the tcgen05 version of this dense contraction.
#include "common.cuh"

namespace quest {

constexpr int kPfThreads = 128;
constexpr int kPfBQ = 16;   // queries per CTA
constexpr int kPfBK = 32;   // kv tokens per smem tile
constexpr int kPfTPQ = kPfThreads / kPfBQ;  // 8 threads per query

template <int D>
__global__ void __launch_bounds__(kPfThreads)
prefill_kernel(const __half* __restrict__ q, __half* __restrict__ o, uint32_t qo_len,
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

}  // namespace quest

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
  const uint32_t kv_len = (uint32_t)kv_len64;
  dim3 grid((qo_len + quest::kPfBQ - 1) / quest::kPfBQ, num_qo_heads);
  const float sc = quest::kLog2e / sqrtf((float)head_dim);
  cudaStream_t s = (cudaStream_t)stream;
#define QUEST_PF(DD)                                                                                              \
  quest::prefill_kernel<DD><<<grid, quest::kPfThreads, 0, s>>>((const __half*)q, (__half*)o, qo_len,              \
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
