// Page-criticality estimate: score[h,p] = sum_d max(q[h,d]*kmax[p,h,d], q[h,d]*kmin[p,h,d]).
// Replaces MaxPossibleSampleWithPagedKVCacheKernel (reference kernels/include/decode/decode_attn.cuh:245-401,
// inner loop :137-168).  HBM-bound stream over the metadata pool: the whole [pages, heads, D] max and
// min slabs are read once, fully coalesced (a row = one (page, kv head) = D halves; consecutive rows are
// contiguous in the NHD pool), with kUnroll independent 128-bit loads in flight per thread.
#include "common.cuh"

namespace quest {

template <int D, int kThreads, int kUnroll>
__global__ void __launch_bounds__(kThreads)
estimate_kernel(const __half* __restrict__ q, __half* __restrict__ o, uint32_t len,
                const __half* __restrict__ meta, const int32_t* __restrict__ meta_indices,
                uint32_t Hq, uint32_t Hkv, uint32_t P) {
  constexpr int LPR = D / 8;                 // lanes per row (reference bdx)
  constexpr int kRowsPerPass = kThreads / LPR;
  extern __shared__ float q_s[];             // [Hq][D] fp32
  const uint32_t G = Hq / Hkv;
  const uint32_t tid = threadIdx.x;
  const uint32_t lane_in_row = tid % LPR;
  const uint32_t row_in_pass = tid / LPR;
  const uint32_t total_rows = len * Hkv;
  const uint32_t base = blockIdx.x * (kRowsPerPass * kUnroll);

  pdl_wait();
  // Issue the metadata loads first; q staging overlaps with their latency.
  H8 mx[kUnroll], mn[kUnroll];
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const uint32_t r = base + u * kRowsPerPass + row_in_pass;
    if (r < total_rows) {
      const uint32_t p = r / Hkv, kvh = r - p * Hkv;
      const uint32_t mp = p / P, e = p - mp * P;
      const size_t page = (size_t)__ldg(meta_indices + mp);
      const __half* a = meta + nhd_off(page, 0, e, kvh, P, Hkv, D) + lane_in_row * 8;
      mx[u] = ld_stream(a);
      mn[u] = ld_stream(a + (size_t)P * Hkv * D);
    }
  }
  for (uint32_t i = tid; i < Hq * D; i += kThreads) q_s[i] = __half2float(q[i]);
  __syncthreads();
  pdl_trigger();

#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const uint32_t r = base + u * kRowsPerPass + row_in_pass;
    const bool valid = r < total_rows;      // whole LPR-lane group is uniformly valid
    const uint32_t p = valid ? r / Hkv : 0, kvh = valid ? r - p * Hkv : 0;
    float fmx[8], fmn[8];
    if (valid) {
      h8_to_float(mx[u], fmx);
      h8_to_float(mn[u], fmn);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) fmx[i] = fmn[i] = 0.f;
    }
    for (uint32_t g = 0; g < G; ++g) {
      const uint32_t h = kvh * G + g;
      const float4* qp = reinterpret_cast<const float4*>(q_s + h * D + lane_in_row * 8);
      const float4 q0 = qp[0], q1 = qp[1];
      const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      // Reference order (decode_attn.cuh:152-160): 8 sequential fp32 adds, then xor butterfly.
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) s = __fadd_rn(s, fmaxf(__fmul_rn(qv[i], fmx[i]), __fmul_rn(qv[i], fmn[i])));
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
      if (valid && lane_in_row == 0) o[(size_t)h * len + p] = __float2half_rn(s);
    }
  }
}

template <int D>
static int launch_estimate(const void* q, void* o, uint32_t len, const void* meta,
                           const int32_t* meta_indices, uint32_t Hq, uint32_t Hkv, uint32_t P,
                           cudaStream_t stream) {
  constexpr int kThreads = 256, kUnroll = 8;
  constexpr int kRowsPerCta = (kThreads / (D / 8)) * kUnroll;
  const uint32_t total_rows = len * Hkv;
  const uint32_t grid = (total_rows + kRowsPerCta - 1) / kRowsPerCta;
  const size_t smem = (size_t)Hq * D * sizeof(float);
  auto kern = estimate_kernel<D, kThreads, kUnroll>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, (const __half*)q, (__half*)o, len, (const __half*)meta,
                                     meta_indices, Hq, Hkv, P);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

}  // namespace quest

extern "C" int quest_estimate(const void* q, void* o, uint32_t len, const void* meta_data,
                              const int32_t* meta_indices, uint32_t num_qo_heads, uint32_t num_kv_heads,
                              uint32_t head_dim, uint32_t page_size, void* stream) {
  if (!q || !o || !meta_data || !meta_indices || !num_qo_heads || !num_kv_heads || !page_size)
    return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (len == 0) return QUEST_OK;
  if ((size_t)num_qo_heads * head_dim * sizeof(float) > 200 * 1024) return QUEST_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  switch (head_dim) {
    case 64: return quest::launch_estimate<64>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, s);
    case 128: return quest::launch_estimate<128>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, s);
    case 256: return quest::launch_estimate<256>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, s);
    default: return QUEST_ERR_UNSUPPORTED;
  }
}
