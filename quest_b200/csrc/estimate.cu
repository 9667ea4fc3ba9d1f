// Page-criticality estimate: score[h,p] = sum_d max(q[h,d]*kmax[p,h,d], q[h,d]*kmin[p,h,d]).
// Replaces MaxPossibleSampleWithPagedKVCacheKernel (reference kernels/include/decode/decode_attn.cuh:245-401,
// inner loop :137-168).
//
// HBM-bound stream over the metadata pool, read exactly once.  Persistent warps, each software-pipelined
// kDepth items deep in registers (an item = one 128-bit load of the max slab + one of the min slab per lane
// = 32/LPR metadata rows), items interleaved across all warps of the grid so neighbouring warps read
// neighbouring 512-byte runs.  ~128 KB per SM stay in flight for the whole kernel; the arithmetic of one
// item hides under the loads of the next kDepth.
//
// Arithmetic (bit-identical to the reference's fp32 order, 8 sequential adds per lane then an xor butterfly):
// products of two fp16 values are exact in fp32, and multiplication by q is monotone, so
//   max(q*kmax, q*kmin) == q * (q >= 0 ? max(kmax,kmin) : min(kmax,kmin))
// for all finite inputs with no assumption that kmax >= kmin (signed zeros cannot change the running sum,
// which is never -0).  The select runs on packed halves (HMNMX2 + one LOP3 per pair against a per-head sign
// mask staged in shared memory), and s + q*x is one FFMA because the product is exact.
#include <stdlib.h>

#include "common.cuh"

namespace quest {

template <int D, int kEstThreads, int kEstDepth>
__global__ void __launch_bounds__(kEstThreads, 2)
estimate_kernel(const __half* __restrict__ q, __half* __restrict__ o, uint32_t len,
                const __half* __restrict__ meta, const int32_t* __restrict__ meta_indices,
                uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t n_meta_smem, uint32_t meta_stable) {
  constexpr int LPR = D / 8;     // lanes per metadata row (reference bdx)
  constexpr int RPL = 32 / LPR;  // rows per warp-wide load
  extern __shared__ __align__(16) uint8_t est_smem[];
  float* q_s = reinterpret_cast<float*>(est_smem);                          // [Hq][D] fp32
  uint32_t* m_s = reinterpret_cast<uint32_t*>(q_s + (size_t)Hq * D);        // [Hq][D/2] 0xffff where q >= 0
  int32_t* idx_s = reinterpret_cast<int32_t*>(m_s + (size_t)Hq * D / 2);    // [n_meta_smem] metadata page table
  const uint32_t G = Hq / Hkv;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lir = lane % LPR, rsub = lane / LPR;
  const uint32_t total_rows = len * Hkv;
  const uint32_t total_items = (total_rows + RPL - 1) / RPL;
  const uint32_t n_warps = gridDim.x * (kEstThreads / 32);
  const uint32_t w_global = warp * gridDim.x + blockIdx.x;  // warps of one CTA are gridDim.x items apart
  const size_t slab = (size_t)P * Hkv * D;

  // In the layer step the kernel that precedes this one (append) only touches the metadata entry of the
  // last page, which is never scored: with meta_stable the pipeline is primed before the
  // programmatic-dependency wait (q is only read after it).  Otherwise wait first.
  if (!meta_stable) pdl_wait();
  const uint32_t n_meta = (len + P - 1) / P;
  for (uint32_t i = tid; i < n_meta && i < n_meta_smem; i += kEstThreads) idx_s[i] = __ldg(meta_indices + i);
  __syncthreads();

  // incremental (page, kv head, metadata page, entry) of this lane's row, advanced by n_warps*RPL rows per item
  const uint32_t step_rows = n_warps * RPL;
  const uint32_t dp = step_rows / Hkv, dh = step_rows % Hkv;
  const uint32_t dmp = dp / P, de = dp % P;
  uint32_t r = w_global * RPL + rsub;
  uint32_t p = r / Hkv, kvh = r % Hkv, mp = p / P, e = p % P;

  H8 mx[kEstDepth], mn[kEstDepth];
  uint32_t ld_p = p, ld_kvh = kvh, ld_mp = mp, ld_e = e, ld_r = r;
  auto issue = [&](int s) {
    if (ld_r < total_rows) {
      const size_t page = (size_t)(ld_mp < n_meta_smem ? idx_s[ld_mp] : __ldg(meta_indices + ld_mp));
      const __half* a = meta + nhd_off(page, 0, ld_e, ld_kvh, P, Hkv, D) + lir * 8;
      mx[s] = ld_stream(a);
      mn[s] = ld_stream(a + slab);
    }
    // advance to this lane's row in the next item
    ld_r += step_rows;
    ld_kvh += dh;
    uint32_t carry = 0;
    if (ld_kvh >= Hkv) {
      ld_kvh -= Hkv;
      carry = 1;
    }
    ld_p += dp + carry;
    ld_e += de + carry;
    ld_mp += dmp;
    while (ld_e >= P) {
      ld_e -= P;
      ++ld_mp;
    }
  };
#pragma unroll
  for (int s = 0; s < kEstDepth; ++s) issue(s);

  if (meta_stable) pdl_wait();
  for (uint32_t i = tid; i < Hq * D / 2; i += kEstThreads) {
    const __half2 h2 = reinterpret_cast<const __half2*>(q)[i];
    const float2 f = __half22float2(h2);
    q_s[2 * i] = f.x;
    q_s[2 * i + 1] = f.y;
    // sign-bit test so that -0 selects like the reference's fmaxf on (+-0) products: either choice sums alike
    const uint32_t bits = *reinterpret_cast<const uint32_t*>(&h2);
    m_s[i] = ((bits & 0x8000u) ? 0u : 0xffffu) | ((bits & 0x80000000u) ? 0u : 0xffff0000u);
  }
  __syncthreads();
  pdl_trigger();

  for (uint32_t it = w_global; it < total_items; it += n_warps * kEstDepth) {
#pragma unroll
    for (int s = 0; s < kEstDepth; ++s) {
      const bool valid = r < total_rows;  // uniform over the LPR lanes of a row
      uint32_t hi[4], lo[4];
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const __half2 a = __hmax2(mx[s].h[i], mn[s].h[i]);
          const __half2 b = __hmin2(mx[s].h[i], mn[s].h[i]);
          hi[i] = *reinterpret_cast<const uint32_t*>(&a);
          lo[i] = *reinterpret_cast<const uint32_t*>(&b);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) hi[i] = lo[i] = 0;
      }
      const uint32_t cur_p = p, cur_kvh = kvh;
      issue(s);  // refill this stage with the item kEstDepth ahead
      r += step_rows;
      kvh += dh;
      uint32_t carry = 0;
      if (kvh >= Hkv) {
        kvh -= Hkv;
        carry = 1;
      }
      p += dp + carry;

      for (uint32_t g = 0; g < G; ++g) {
        const uint32_t h = cur_kvh * G + g;
        const float4* qp = reinterpret_cast<const float4*>(q_s + h * D + lir * 8);
        const float4 q0 = qp[0], q1 = qp[1];
        const uint4 msk = *reinterpret_cast<const uint4*>(m_s + h * (D / 2) + lir * 4);
        const float qv[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        const uint32_t mk[4] = {msk.x, msk.y, msk.z, msk.w};
        float sacc = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t sel = (hi[i] & mk[i]) | (lo[i] & ~mk[i]);
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&sel));
          // reference order (decode_attn.cuh:152-160): 8 sequential fp32 adds of exact products
          sacc = __fmaf_rn(qv[2 * i], f.x, sacc);
          sacc = __fmaf_rn(qv[2 * i + 1], f.y, sacc);
        }
#pragma unroll
        for (int off = LPR / 2; off > 0; off >>= 1) sacc = __fadd_rn(sacc, __shfl_xor_sync(0xffffffffu, sacc, off));
        if (valid && lir == 0) o[(size_t)h * len + cur_p] = __float2half_rn(sacc);
      }
    }
  }
}

template <int D, int kEstThreads, int kEstDepth>
static int launch_estimate(const void* q, void* o, uint32_t len, const void* meta,
                           const int32_t* meta_indices, uint32_t Hq, uint32_t Hkv, uint32_t P,
                           uint32_t flags, cudaStream_t stream) {
  constexpr int RPL = 32 / (D / 8);
  static int num_sms = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (num_sms <= 0) num_sms = 148;
  }
  const uint32_t total_items = (len * Hkv + RPL - 1) / RPL;
  const uint32_t warps_per_cta = kEstThreads / 32;
  uint32_t grid = 2 * num_sms;
  const uint32_t need = (total_items + warps_per_cta - 1) / warps_per_cta;
  if (grid > need) grid = need;
  const uint32_t n_meta = (len + P - 1) / P;
  const uint32_t n_meta_smem = n_meta < 8192 ? n_meta : 8192;
  const size_t smem = (size_t)Hq * D * sizeof(float) + (size_t)Hq * D / 2 * sizeof(uint32_t) + (size_t)n_meta_smem * 4;
  auto kern = estimate_kernel<D, kEstThreads, kEstDepth>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kEstThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, (const __half*)q, (__half*)o, len, (const __half*)meta,
                                     meta_indices, Hq, Hkv, P, n_meta_smem, flags & QUEST_EST_META_STABLE);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

}  // namespace quest

extern "C" int quest_estimate(const void* q, void* o, uint32_t len, const void* meta_data,
                              const int32_t* meta_indices, uint32_t num_qo_heads, uint32_t num_kv_heads,
                              uint32_t head_dim, uint32_t page_size, uint32_t flags, void* stream) {
  if (!q || !o || !meta_data || !meta_indices || !num_qo_heads || !num_kv_heads || !page_size)
    return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (len == 0) return QUEST_OK;
  if ((size_t)num_qo_heads * head_dim * 6 > 160 * 1024) return QUEST_ERR_UNSUPPORTED;
  cudaStream_t s = (cudaStream_t)stream;
  static int cfg = -1;  // tuning hook: QUEST_EST_CFG selects the (threads, depth) variant for head_dim 128
  if (cfg < 0) {
    const char* e = getenv("QUEST_EST_CFG");
    cfg = e ? atoi(e) : 0;
  }
  if (head_dim == 128 && cfg == 1) return quest::launch_estimate<128, 512, 3>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
  if (head_dim == 128 && cfg == 2) return quest::launch_estimate<128, 384, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
  if (head_dim == 128 && cfg == 3) return quest::launch_estimate<128, 512, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
  switch (head_dim) {
    case 64: return quest::launch_estimate<64, 448, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
    case 128: return quest::launch_estimate<128, 448, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
    case 256: return quest::launch_estimate<256, 448, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
    default: return QUEST_ERR_UNSUPPORTED;
  }
}
