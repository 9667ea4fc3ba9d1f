// Sparse paged decode attention over per-head selected pages, split-KV with an in-kernel LSE combine.
// Replaces BatchDecodeWithPagedKVCacheKernel<partition_kv=true> + VariableLengthMergeStates + the
// BatchDecodeHandler planner (reference kernels/include/decode/decode_attn.cuh:440-646, 753-893, 962-1002;
// decode_handler.cuh:73-121; flashinfer cascade.cuh:279-340, state.cuh:53-62).
//
// Unit of work = (q head, chunk of pages).  Each group of D/8 lanes owns whole 16-byte-per-lane token
// rows; a tile of kTile tokens issues 2*kTile independent 128-bit loads per lane before any use, so a
// CTA has its whole chunk in flight at once.  Partials go to a caller-owned workspace; the last CTA to
// arrive for a head (atomic ticket) merges them -- no separate merge launch.
#include "common.cuh"

namespace quest {

constexpr int kDecThreads = 128;
constexpr int kTile = 8;
constexpr float kNegBig = -1e30f;

// workspace layout: [Hq*nchunks*D] fp32 o | [Hq*nchunks*2] fp32 (m,d) | [Hq] u32 tickets
__host__ __device__ inline size_t ws_md_off(uint32_t Hq, uint32_t D, uint32_t nchunks) {
  return (size_t)Hq * nchunks * D * sizeof(float);
}
__host__ __device__ inline size_t ws_ticket_off(uint32_t Hq, uint32_t D, uint32_t nchunks) {
  return ws_md_off(Hq, D, nchunks) + (size_t)Hq * nchunks * 2 * sizeof(float);
}

template <int D>
__global__ void __launch_bounds__(kDecThreads)
decode_kernel(const __half* __restrict__ q, __half* __restrict__ o, float* __restrict__ lse,
              const __half* __restrict__ kv, const int32_t* __restrict__ indices, uint32_t idx_stride,
              uint32_t n_sel, uint32_t last_page_idx, uint32_t last_page_len, uint32_t Hkv, uint32_t G,
              uint32_t P, uint32_t ppc, float sm_scale_log2, uint8_t* __restrict__ ws) {
  constexpr int LPR = D / 8;
  constexpr int NG = kDecThreads / LPR;
  __shared__ float s_o[NG][D];
  __shared__ float s_md[NG][2];
  __shared__ uint32_t s_last;
  const uint32_t chunk = blockIdx.x, nchunks = gridDim.x, h = blockIdx.y, Hq = gridDim.y;
  const uint32_t kvh = h / G;
  const uint32_t tid = threadIdx.x, gid = tid / LPR, lir = tid % LPR;
  const uint32_t total_pages = n_sel + 1;  // + forced last page
  const uint32_t pg0 = chunk * ppc, pg1 = min(pg0 + ppc, total_pages);
  const uint32_t ntok = (pg1 - pg0) * P;
  const size_t slab = (size_t)P * Hkv * D;  // K -> V offset inside a page

  float qv[8];
  {
    const H8 qh = ld_plain(q + (size_t)h * D + lir * 8);
    h8_to_float(qh, qv);
#pragma unroll
    for (int i = 0; i < 8; ++i) qv[i] *= sm_scale_log2;
  }
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  float m = kNegBig, d = 0.f;

  pdl_wait();  // indices come from the top-k kernel
  const int32_t* idx_row = indices + (size_t)h * idx_stride;

  // Warp-uniform trip count (the shuffles below use the full mask); out-of-range tokens are masked.
  for (uint32_t base = 0; base < ntok; base += NG * kTile) {
    const uint32_t t0 = base + gid * kTile;
    H8 kk[kTile], vv[kTile];
    bool ok[kTile];
#pragma unroll
    for (int j = 0; j < kTile; ++j) {
      const uint32_t t = t0 + j;
      const uint32_t pg = pg0 + t / P, e = t % P;
      ok[j] = t < ntok && (pg < n_sel || e < last_page_len);
      if (ok[j]) {
        const size_t page = pg < n_sel ? (size_t)__ldg(idx_row + pg) : (size_t)last_page_idx;
        const __half* kp = kv + nhd_off(page, 0, e, kvh, P, Hkv, D) + lir * 8;
        kk[j] = ld_stream(kp);
        vv[j] = ld_stream(kp + slab);
      }
    }
    float s[kTile];
    float mt = m;
#pragma unroll
    for (int j = 0; j < kTile; ++j) {
      float kf[8];
      float a = 0.f;
      if (ok[j]) {
        h8_to_float(kk[j], kf);
#pragma unroll
        for (int i = 0; i < 8; ++i) a = fmaf(qv[i], kf[i], a);
      }
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
      s[j] = ok[j] ? a : kNegBig;
      mt = fmaxf(mt, s[j]);
    }
    const float scale = ex2(m - mt);
    m = mt;
    d *= scale;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] *= scale;
#pragma unroll
    for (int j = 0; j < kTile; ++j) {
      if (ok[j]) {
        const float p = ex2(s[j] - m);
        d += p;
        float vf[8];
        h8_to_float(vv[j], vf);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fmaf(p, vf[i], acc[i]);
      }
    }
  }

  // CTA merge of the NG group states.
#pragma unroll
  for (int i = 0; i < 8; ++i) s_o[gid][lir * 8 + i] = acc[i];
  if (lir == 0) {
    s_md[gid][0] = m;
    s_md[gid][1] = d;
  }
  __syncthreads();
  constexpr int CPT = (D + kDecThreads - 1) / kDecThreads;  // output columns per thread
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
      const uint32_t col = tid + c * kDecThreads;
      if (col < D) osum[c] += s_o[g][col] * w;
    }
  }
  if (nchunks == 1) {
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const uint32_t col = tid + c * kDecThreads;
      if (col < D) o[(size_t)h * D + col] = __float2half_rn(__fdividef(osum[c], dsum));
    }
    if (lse && tid == 0) lse[h] = M + lg2(dsum);
    return;
  }
  float* ws_o = reinterpret_cast<float*>(ws);
  float* ws_md = reinterpret_cast<float*>(ws + ws_md_off(Hq, D, nchunks));
  uint32_t* ticket = reinterpret_cast<uint32_t*>(ws + ws_ticket_off(Hq, D, nchunks));
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    const uint32_t col = tid + c * kDecThreads;
    if (col < D) ws_o[((size_t)h * nchunks + chunk) * D + col] = osum[c];
  }
  if (tid == 0) {
    ws_md[((size_t)h * nchunks + chunk) * 2] = M;
    ws_md[((size_t)h * nchunks + chunk) * 2 + 1] = dsum;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(ticket + h, 1u) == nchunks - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  float MM = kNegBig;
  for (uint32_t c = 0; c < nchunks; ++c) MM = fmaxf(MM, __ldcg(ws_md + ((size_t)h * nchunks + c) * 2));
  float dd = 0.f, oo[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) oo[c] = 0.f;
  for (uint32_t cc = 0; cc < nchunks; ++cc) {
    const float w = ex2(__ldcg(ws_md + ((size_t)h * nchunks + cc) * 2) - MM);
    dd += __ldcg(ws_md + ((size_t)h * nchunks + cc) * 2 + 1) * w;
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const uint32_t col = tid + c * kDecThreads;
      if (col < D) oo[c] += __ldcg(ws_o + ((size_t)h * nchunks + cc) * D + col) * w;
    }
  }
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    const uint32_t col = tid + c * kDecThreads;
    if (col < D) o[(size_t)h * D + col] = __float2half_rn(__fdividef(oo[c], dd));
  }
  if (tid == 0) {
    if (lse) lse[h] = MM + lg2(dd);
    ticket[h] = 0;  // ready for the next launch
  }
}

template <int D>
static int launch_decode(const void* q, void* o, float* lse, const void* kv, const int32_t* indices,
                         uint32_t idx_stride, uint32_t n_sel, uint32_t last_page_idx, uint32_t last_page_len,
                         uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t ppc, void* ws, cudaStream_t stream) {
  const uint32_t nchunks = (n_sel + 1 + ppc - 1) / ppc;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(nchunks, Hq);
  cfg.blockDim = dim3(kDecThreads);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const float sm_scale_log2 = kLog2e / sqrtf((float)D);
  cudaError_t e = cudaLaunchKernelEx(&cfg, decode_kernel<D>, (const __half*)q, (__half*)o, lse, (const __half*)kv,
                                     indices, idx_stride, n_sel, last_page_idx, last_page_len, Hkv, Hq / Hkv, P,
                                     ppc, sm_scale_log2, (uint8_t*)ws);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

}  // namespace quest

extern "C" int quest_decode_plan(uint32_t num_sel_pages, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                 uint32_t head_dim, uint32_t page_size, uint32_t num_sms,
                                 uint32_t* pages_per_chunk, uint32_t* num_chunks) {
  if (!pages_per_chunk || !num_chunks || !num_qo_heads || !num_kv_heads || !page_size) return QUEST_ERR_BAD_ARG;
  if (head_dim != 64 && head_dim != 128 && head_dim != 256) return QUEST_ERR_UNSUPPORTED;
  if (num_sms == 0) num_sms = 148;
  const uint32_t total = num_sel_pages + 1;
  const uint32_t slots = num_sms * 4;  // resident 128-thread CTAs per SM at ~128 registers/thread
  // Smallest chunk that keeps every CTA resident in one wave, but at least one kTile per lane group.
  const uint32_t groups = quest::kDecThreads / (head_dim / 8);
  uint32_t min_ppc = (groups * quest::kTile + page_size - 1) / page_size;
  if (min_ppc == 0) min_ppc = 1;
  uint32_t ppc = min_ppc;
  while (ppc < total && (uint64_t)num_qo_heads * ((total + ppc - 1) / ppc) > slots) ++ppc;
  if (ppc > total) ppc = total;
  *pages_per_chunk = ppc;
  *num_chunks = (total + ppc - 1) / ppc;
  return QUEST_OK;
}

extern "C" size_t quest_decode_workspace_bytes(uint32_t num_qo_heads, uint32_t head_dim, uint32_t num_chunks) {
  return quest::ws_ticket_off(num_qo_heads, head_dim, num_chunks) + (size_t)num_qo_heads * sizeof(uint32_t);
}

extern "C" int quest_decode_run(const void* q, void* o, float* lse, const void* kv_data, const int32_t* indices,
                                uint32_t idx_row_stride, uint32_t num_sel_pages, uint32_t last_page_idx,
                                uint32_t last_page_len, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                uint32_t head_dim, uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                size_t workspace_bytes, void* stream) {
  if (!q || !o || !kv_data || !num_qo_heads || !num_kv_heads || !page_size || !pages_per_chunk)
    return QUEST_ERR_BAD_ARG;
  if (num_sel_pages && !indices) return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (last_page_len < 1 || last_page_len > page_size) return QUEST_ERR_BAD_ARG;
  const uint32_t nchunks = (num_sel_pages + 1 + pages_per_chunk - 1) / pages_per_chunk;
  if (nchunks > 1) {
    if (!workspace || workspace_bytes < quest_decode_workspace_bytes(num_qo_heads, head_dim, nchunks))
      return QUEST_ERR_WORKSPACE;
  }
  cudaStream_t s = (cudaStream_t)stream;
#define QUEST_DEC(DD)                                                                                         \
  return quest::launch_decode<DD>(q, o, lse, kv_data, indices, idx_row_stride, num_sel_pages, last_page_idx, \
                                  last_page_len, num_qo_heads, num_kv_heads, page_size, pages_per_chunk,     \
                                  workspace, s)
  switch (head_dim) {
    case 64: QUEST_DEC(64);
    case 128: QUEST_DEC(128);
    case 256: QUEST_DEC(256);
    default: return QUEST_ERR_UNSUPPORTED;
  }
#undef QUEST_DEC
}
