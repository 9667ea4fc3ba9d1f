// RoPE (in place), RMSNorm and the LSE merge of partial attention states.
// RoPE replaces QKApplyRotaryInPlaceKernel (reference kernels/include/decode/decode_page.cuh:644-728,
// flashinfer rope.cuh:64-83): Llama rotate-half, freq_i = (1/scale) * (1/theta)^(2(i mod D/2)/D) with
// the same fast intrinsics (__powf, __sincosf) so results track the reference kernel.
// RMSNorm replaces rmsnorm_twoPassAlgo_e8 (reference quest/ops/csrc/rms_norm.cu:82-158).
#include "common.cuh"

namespace quest {

// One thread per 16-byte vector of one (token, head) row; grid-stride over rows of q then k.
template <int D>
__global__ void __launch_bounds__(256)
rope_kernel(__half* __restrict__ q, __half* __restrict__ k, uint32_t seq_len, uint32_t past_len, uint32_t Hq,
            uint32_t Hkv, float rcp_scale, float rcp_theta, const SeqDesc* __restrict__ descs) {
  constexpr int LPR = D / 8;
  pdl_wait();
  const uint32_t lir = threadIdx.x % LPR;
  float freq[8];
  rope_freq8<D>(lir * 8, rcp_scale, rcp_theta, freq);
  const size_t q_rows = (size_t)seq_len * Hq, rows = q_rows + (size_t)seq_len * Hkv;
  const size_t stride = (size_t)gridDim.x * (blockDim.x / LPR);
  for (size_t r = (size_t)blockIdx.x * (blockDim.x / LPR) + threadIdx.x / LPR; r < rows; r += stride) {
    const bool is_q = r < q_rows;
    const size_t rr = is_q ? r : r - q_rows;
    const uint32_t H = is_q ? Hq : Hkv;
    // descs: row i is the new token of sequence i, rotated at that sequence's own position (read from its device record)
    const uint32_t pos = descs ? descs[rr / H].rope_pos : past_len + (uint32_t)(rr / H);
    __half* x = (is_q ? q : k) + rr * D;
    const bool lo_half = lir * 8 < D / 2;
    float a[8], b[8];
    h8_to_float(ld_plain(x + lir * 8), a);
    h8_to_float(ld_plain(x + (lo_half ? lir * 8 + D / 2 : lir * 8 - D / 2)), b);
    __syncwarp();  // partner lanes have read their halves before anyone overwrites them
    const H8 out = rope_rotate8(a, b, lo_half, float(pos), freq);
    st_plain(x + lir * 8, out);
  }
}

__global__ void __launch_bounds__(256)
rms_norm_kernel(const __half* __restrict__ in, const __half* __restrict__ w, __half* __restrict__ out,
                uint32_t cols, float eps) {
  __shared__ float red[8];
  pdl_wait();
  const size_t row = blockIdx.x;
  const __half* x = in + row * cols;
  float ss = 0.f;
  for (uint32_t c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
    float f[8];
    h8_to_float(ld_plain(x + c), f);
#pragma unroll
    for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float inv = rsqrtf(tot / (float)cols + eps);
  for (uint32_t c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
    float f[8], g[8];
    h8_to_float(ld_plain(x + c), f);
    h8_to_float(ld_plain(w + c), g);
    H8 o;
#pragma unroll
    for (int i = 0; i < 4; ++i) o.h[i] = __floats2half2_rn(f[2 * i] * inv * g[2 * i], f[2 * i + 1] * inv * g[2 * i + 1]);
    st_plain(out + row * cols + c, o);
  }
}

// o = sum_i o_i 2^(lse_i - L) / sum_i 2^(lse_i - L): merge of normalised partial states.
// Part i's o block starts at o_parts + i*so, its lse block at lse_parts + i*sl (dense: so = H*D, sl = H).
__global__ void merge_states_kernel(const float* __restrict__ o_parts, const float* __restrict__ lse_parts,
                                    uint32_t n, size_t so, size_t sl, __half* __restrict__ o,
                                    float* __restrict__ lse, uint32_t H, uint32_t D) {
  pdl_wait();
  const uint32_t h = blockIdx.x;
  float L = -1e30f;
  for (uint32_t i = 0; i < n; ++i) L = fmaxf(L, lse_parts[i * sl + h]);
  float den = 0.f;
  for (uint32_t i = 0; i < n; ++i) den += ex2(lse_parts[i * sl + h] - L);
  for (uint32_t c = threadIdx.x; c < D; c += blockDim.x) {
    float acc = 0.f;
    for (uint32_t i = 0; i < n; ++i)
      acc += o_parts[i * so + (size_t)h * D + c] * ex2(lse_parts[i * sl + h] - L);
    o[(size_t)h * D + c] = __float2half_rn(__fdividef(acc, den));
  }
  if (lse && threadIdx.x == 0) lse[h] = L + lg2(den);
}

}  // namespace quest

static int rope_launch(void* q, void* k, uint32_t seq_len, uint32_t past_kv_len, uint32_t num_qo_heads,
                       uint32_t num_kv_heads, uint32_t head_dim, float rope_scale, float rope_theta, const void* descs,
                       void* stream) {
  if (!q || !k || !num_qo_heads || !num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (seq_len == 0) return QUEST_OK;
  const float rs = 1.f / rope_scale, rt = 1.f / rope_theta;
  const size_t rows = (size_t)seq_len * (num_qo_heads + num_kv_heads);
  const uint32_t rows_per_cta = 256 / (head_dim / 8);
  const uint32_t grid = (uint32_t)((rows + rows_per_cta - 1) / rows_per_cta < 148 * 8 ? (rows + rows_per_cta - 1) / rows_per_cta : 148 * 8);
  cudaStream_t s = (cudaStream_t)stream;
  switch (head_dim) {
    case 64: quest::rope_kernel<64><<<grid, 256, 0, s>>>((__half*)q, (__half*)k, seq_len, past_kv_len, num_qo_heads, num_kv_heads, rs, rt, (const quest::SeqDesc*)descs); break;
    case 128: quest::rope_kernel<128><<<grid, 256, 0, s>>>((__half*)q, (__half*)k, seq_len, past_kv_len, num_qo_heads, num_kv_heads, rs, rt, (const quest::SeqDesc*)descs); break;
    case 256: quest::rope_kernel<256><<<grid, 256, 0, s>>>((__half*)q, (__half*)k, seq_len, past_kv_len, num_qo_heads, num_kv_heads, rs, rt, (const quest::SeqDesc*)descs); break;
    default: return QUEST_ERR_UNSUPPORTED;
  }
  QUEST_RETURN_LAUNCH();
}

extern "C" int quest_rms_norm(const void* in, const void* weight, void* out, uint32_t rows, uint32_t cols,
                              float epsilon, void* stream) {
  if (!in || !weight || !out) return QUEST_ERR_BAD_ARG;
  if (cols % 8) return QUEST_ERR_UNSUPPORTED;
  if (rows == 0) return QUEST_OK;
  quest::rms_norm_kernel<<<rows, 256, 0, (cudaStream_t)stream>>>((const __half*)in, (const __half*)weight,
                                                                 (__half*)out, cols, epsilon);
  QUEST_RETURN_LAUNCH();
}

extern "C" int quest_merge_states(const float* o_parts, const float* lse_parts, uint32_t n, uint32_t part_stride,
                                  void* o, float* lse, uint32_t num_heads, uint32_t head_dim, void* stream) {
  if (!o_parts || !lse_parts || !o || !n || !num_heads || !head_dim) return QUEST_ERR_BAD_ARG;
  if (part_stride && part_stride < num_heads * head_dim) return QUEST_ERR_BAD_ARG;
  const size_t so = part_stride ? part_stride : (size_t)num_heads * head_dim, sl = part_stride ? part_stride : num_heads;
  quest::merge_states_kernel<<<num_heads, 128, 0, (cudaStream_t)stream>>>(o_parts, lse_parts, n, so, sl, (__half*)o,
                                                                          lse, num_heads, head_dim);
  QUEST_RETURN_LAUNCH();
}

extern "C" int quest_rope_inplace(void* q, void* k, uint32_t seq_len, uint32_t past_kv_len, uint32_t num_qo_heads,
                                  uint32_t num_kv_heads, uint32_t head_dim, float rope_scale, float rope_theta,
                                  void* stream) {
  return rope_launch(q, k, seq_len, past_kv_len, num_qo_heads, num_kv_heads, head_dim, rope_scale, rope_theta, nullptr, stream);
}

extern "C" int quest_rope_decode_batch(void* q, void* k, const QuestSeqDesc* descs, uint32_t batch, uint32_t num_qo_heads,
                                       uint32_t num_kv_heads, uint32_t head_dim, float rope_scale, float rope_theta,
                                       void* stream) {
  if (!descs) return QUEST_ERR_BAD_ARG;
  return rope_launch(q, k, batch, 0, num_qo_heads, num_kv_heads, head_dim, rope_scale, rope_theta, descs, stream);
}
