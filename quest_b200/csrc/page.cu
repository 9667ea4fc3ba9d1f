// append_kv fused with the per-page min/max Key metadata update, decode (1 token) and prefill (range).
// Replaces AppendPagedKVCacheDecodeKernel / AppendPagedKVCachePrefillKernel
// (reference kernels/include/decode/decode_page.cuh:398-449, 471-562).
// In the NHD pool a page's K slab [P][H][D] has the same layout as the source rows [tokens][H][D], so
// the prefill variant is a coalesced 128-bit copy with the min/max reduction folded in.
#include "common.cuh"

namespace quest {

// One thread per 16-byte column vector of the [H*D] row.
__global__ void __launch_bounds__(256)
append_decode_kernel(const __half* __restrict__ k, const __half* __restrict__ v, __half* __restrict__ kv,
                     const int32_t* __restrict__ kv_indices, uint32_t num_kv_pages, uint32_t kv_last_len,
                     __half* __restrict__ meta, const int32_t* __restrict__ meta_indices,
                     uint32_t num_meta_pages, uint32_t meta_last_len, uint32_t row_vecs /* H*D/8 */,
                     uint32_t P) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  pdl_wait();
  if (c >= row_vecs) return;
  append_decode_column(k, v, kv, kv_indices, num_kv_pages, kv_last_len, meta, meta_indices, num_meta_pages, meta_last_len,
                       row_vecs, P, c);
}

// grid = (pages touched, column slices); each thread owns one 16-byte column of the page.
template <int kBatch>
__global__ void __launch_bounds__(256)
append_prefill_kernel(const __half* __restrict__ k, const __half* __restrict__ v, uint32_t append_len,
                      __half* __restrict__ kv, const int32_t* __restrict__ kv_indices,
                      uint32_t num_kv_pages, uint32_t kv_last_len, __half* __restrict__ meta,
                      const int32_t* __restrict__ meta_indices, uint32_t row_vecs, uint32_t P) {
  const uint32_t c = blockIdx.y * blockDim.x + threadIdx.x;
  pdl_wait();
  if (c >= row_vecs) return;
  const size_t row = (size_t)row_vecs * 8;
  const uint32_t seq_len = (num_kv_pages - 1) * P + kv_last_len;
  const uint32_t append_start = seq_len - append_len;
  const uint32_t j = append_start / P + blockIdx.x;  // logical KV page of this CTA
  if (j >= num_kv_pages) return;
  const uint32_t e0 = append_start > j * P ? append_start - j * P : 0;
  const uint32_t e1 = min(P, seq_len - j * P);
  const size_t page = (size_t)__ldg(kv_indices + j);
  const size_t mpage = (size_t)__ldg(meta_indices + j / P);
  const uint32_t mentry = j % P;
  __half* mxp = meta + ((mpage * 2) * P + mentry) * row + c * 8;
  __half* mnp = mxp + (size_t)P * row;
  H8 mx, mn;
  if (e0 > 0) {  // page already held tokens: resume from stored bounds (decode_page.cuh:509-523)
    mx = ld_plain(mxp);
    mn = ld_plain(mnp);
  } else {
    mx = h8_fill(-kMaxHalf);
    mn = h8_fill(kMaxHalf);
  }
  const ptrdiff_t tok0 = (ptrdiff_t)j * P - (ptrdiff_t)append_start;  // source row of entry 0 (may be < 0; e >= e0 keeps tok0+e >= 0)
  __half* kdst = kv + ((page * 2) * P) * row + c * 8;
  __half* vdst = kdst + (size_t)P * row;
  for (uint32_t e = e0; e < e1; e += kBatch) {
    H8 kk[kBatch], vv[kBatch];
#pragma unroll
    for (int i = 0; i < kBatch; ++i)
      if (e + i < e1) {
        const size_t src = (size_t)(tok0 + (ptrdiff_t)(e + i)) * row + c * 8;
        kk[i] = ld_stream(k + src);
        vv[i] = ld_stream(v + src);
      }
#pragma unroll
    for (int i = 0; i < kBatch; ++i)
      if (e + i < e1) {
        h8_max(mx, kk[i]);
        h8_min(mn, kk[i]);
        st_plain(kdst + (size_t)(e + i) * row, kk[i]);
        st_plain(vdst + (size_t)(e + i) * row, vv[i]);
      }
  }
  st_plain(mxp, mx);
  st_plain(mnp, mn);
}


// ---- prefill producer: RoPE(q, k) + append_kv + page min/max in ONE launch -------------------
// Replaces the reference's three steps ahead of prefill attention -- QKApplyRotaryInPlaceKernel over q and k
// (quest/ops/csrc/page.cu:212-253, decode_page.cuh:644-728), then AppendPagedKVCachePrefillKernel (page.cu:101-210,
// decode_page.cuh:471-562) re-reading the rotated k -- with one pass: k and v are read once, k is rotated in registers, the
// rounded fp16 result goes to the page AND into the page's running max/min, q is rotated in place by the same launch's trailing
// CTAs.  k itself is written back only when the caller asks for the reference's in-place side effect (QUEST_PF_WRITE_BACK_K).
//
// grid.x = pages touched + q CTAs.  A page CTA: each thread owns the PAIR of 16-byte columns (d0..d0+7, d0+D/2..) of one kv
// head, i.e. both partners of the rotate-half rule, so every k element is loaded once and every sin/cos computed once.
template <int D, int kBatch>
__global__ void __launch_bounds__(256, 3)
rope_append_prefill_kernel(__half* __restrict__ q, __half* __restrict__ k, const __half* __restrict__ v, uint32_t append_len,
                           __half* __restrict__ kv, const int32_t* __restrict__ kv_indices, uint32_t num_kv_pages,
                           uint32_t kv_last_len, __half* __restrict__ meta, const int32_t* __restrict__ meta_indices,
                           uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t n_page_ctas, float rcp_scale, float rcp_theta,
                           uint32_t write_back_k) {
  constexpr uint32_t PPR = D / 16;  // column pairs per head row
  pdl_wait();
  pdl_trigger();  // the prefill kernel's CTAs may take the slots this grid's last wave frees; they wait for its completion
  const uint32_t seq_len = (num_kv_pages - 1) * P + kv_last_len;
  const uint32_t append_start = seq_len - append_len;  // = RoPE position of the first new token
  float freq[8];
  if (blockIdx.x >= n_page_ctas) {  // ---- q rows: rotate in place, grid-stride
    const uint32_t pr = threadIdx.x % PPR;
    rope_freq8<D>(pr * 8, rcp_scale, rcp_theta, freq);
    const size_t rows = (size_t)append_len * Hq;
    const size_t stride = (size_t)(gridDim.x - n_page_ctas) * (blockDim.x / PPR);
    for (size_t r = (size_t)(blockIdx.x - n_page_ctas) * (blockDim.x / PPR) + threadIdx.x / PPR; r < rows; r += stride) {
      __half* x = q + r * D + pr * 8;
      float lo[8], hi[8];
      h8_to_float(ld_plain(x), lo);
      h8_to_float(ld_plain(x + D / 2), hi);
      const float pos = float(append_start + (uint32_t)(r / Hq));
      st_plain(x, rope_rotate8(lo, hi, true, pos, freq));
      st_plain(x + D / 2, rope_rotate8(hi, lo, false, pos, freq));
    }
    return;
  }
  // ---- one KV page: rows [e0, e1) of logical page j
  const uint32_t pairs = Hkv * PPR;
  const size_t row = (size_t)Hkv * D;
  const uint32_t j = append_start / P + blockIdx.x;
  if (j >= num_kv_pages) return;
  const uint32_t e0 = append_start > j * P ? append_start - j * P : 0;
  const uint32_t e1 = min(P, seq_len - j * P);
  const size_t page = (size_t)__ldg(kv_indices + j);
  const size_t mpage = (size_t)__ldg(meta_indices + j / P);
  const uint32_t mentry = j % P;
  const ptrdiff_t tok0 = (ptrdiff_t)j * P - (ptrdiff_t)append_start;  // source row of entry 0
  for (uint32_t c = threadIdx.x; c < pairs; c += blockDim.x) {
    const uint32_t head = c / PPR, pr = c % PPR;
    const size_t col = (size_t)head * D + pr * 8;  // low column; its partner is D/2 further on
    rope_freq8<D>(pr * 8, rcp_scale, rcp_theta, freq);
    __half* mxp = meta + ((mpage * 2) * P + mentry) * row + col;
    __half* mnp = mxp + (size_t)P * row;
    H8 mx[2], mn[2];
    if (e0 > 0) {  // page already held tokens: resume from stored bounds (decode_page.cuh:509-523)
      mx[0] = ld_plain(mxp), mx[1] = ld_plain(mxp + D / 2);
      mn[0] = ld_plain(mnp), mn[1] = ld_plain(mnp + D / 2);
    } else {
      mx[0] = mx[1] = h8_fill(-kMaxHalf);
      mn[0] = mn[1] = h8_fill(kMaxHalf);
    }
    __half* kdst = kv + ((page * 2) * P) * row + col;
    __half* vdst = kdst + (size_t)P * row;
    for (uint32_t e = e0; e < e1; e += kBatch) {
      H8 kl[kBatch], kh[kBatch], vl[kBatch], vh[kBatch];
#pragma unroll
      for (int i = 0; i < kBatch; ++i)
        if (e + i < e1) {
          const size_t src = (size_t)(tok0 + (ptrdiff_t)(e + i)) * row + col;
          kl[i] = ld_stream(k + src);
          kh[i] = ld_stream(k + src + D / 2);
          vl[i] = ld_stream(v + src);
          vh[i] = ld_stream(v + src + D / 2);
        }
#pragma unroll
      for (int i = 0; i < kBatch; ++i)
        if (e + i < e1) {
          float lo[8], hi[8];
          h8_to_float(kl[i], lo);
          h8_to_float(kh[i], hi);
          const float pos = float(j * P + e + i);
          const H8 rl = rope_rotate8(lo, hi, true, pos, freq), rh = rope_rotate8(hi, lo, false, pos, freq);
          h8_max(mx[0], rl), h8_min(mn[0], rl);
          h8_max(mx[1], rh), h8_min(mn[1], rh);
          st_plain(kdst + (size_t)(e + i) * row, rl);
          st_plain(kdst + (size_t)(e + i) * row + D / 2, rh);
          st_plain(vdst + (size_t)(e + i) * row, vl[i]);
          st_plain(vdst + (size_t)(e + i) * row + D / 2, vh[i]);
          if (write_back_k) {
            const size_t src = (size_t)(tok0 + (ptrdiff_t)(e + i)) * row + col;
            st_plain(k + src, rl);
            st_plain(k + src + D / 2, rh);
          }
        }
    }
    st_plain(mxp, mx[0]), st_plain(mxp + D / 2, mx[1]);
    st_plain(mnp, mn[0]), st_plain(mnp + D / 2, mn[1]);
  }
}


// fp16 metadata entries of KV pages [first, end) -> their e4m3 copies, rounded outward (common.cuh).  grid = (pages, column
// slices); a thread owns one 8-dim column of the entry.
__global__ void __launch_bounds__(256)
metadata_to_fp8_kernel(const __half* __restrict__ meta, const int32_t* __restrict__ meta_indices, uint32_t first,
                       uint8_t* __restrict__ meta8, uint32_t row_vecs, uint32_t P) {
  const uint32_t c = blockIdx.y * blockDim.x + threadIdx.x;
  pdl_wait();
  if (c >= row_vecs) return;
  const uint32_t j = first + blockIdx.x;
  const size_t row = (size_t)row_vecs * 8;
  const size_t mpage = (size_t)__ldg(meta_indices + j / P);
  const size_t ent = ((mpage * 2) * P + j % P) * row + c * 8;
  const H8 mx = ld_ordered_h8(meta + ent), mn = ld_ordered_h8(meta + ent + (size_t)P * row);
  *reinterpret_cast<uint2*>(meta8 + ent) = h8_to_e4m3_up(mx);
  *reinterpret_cast<uint2*>(meta8 + ent + (size_t)P * row) = h8_to_e4m3_down(mn);
}

}  // namespace quest

static bool bad_dims(uint32_t H, uint32_t D, uint32_t P) { return !H || !P || (D != 64 && D != 128 && D != 256); }

extern "C" int quest_append_decode(const void* k, const void* v, void* kv_data, const int32_t* kv_indices,
                                   uint32_t num_kv_pages, uint32_t kv_last_page_len, void* meta_data,
                                   const int32_t* meta_indices, uint32_t num_meta_pages,
                                   uint32_t meta_last_page_len, uint32_t num_heads, uint32_t head_dim,
                                   uint32_t page_size, void* stream) {
  if (!k || !v || !kv_data || !kv_indices || !meta_data || !meta_indices) return QUEST_ERR_BAD_ARG;
  if (bad_dims(num_heads, head_dim, page_size)) return QUEST_ERR_UNSUPPORTED;
  if (!num_kv_pages || !num_meta_pages || kv_last_page_len < 1 || kv_last_page_len > page_size ||
      meta_last_page_len < 1 || meta_last_page_len > page_size)
    return QUEST_ERR_BAD_ARG;
  const uint32_t row_vecs = num_heads * head_dim / 8;
  const uint32_t threads = 256, grid = (row_vecs + threads - 1) / threads;
  // programmatic launch: on the op-by-op path this kernel follows the previous layer's decode kernel, which releases its
  // dependents before its merge tail; the kernel waits (griddepcontrol.wait) before it touches memory
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(threads);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, quest::append_decode_kernel, (const __half*)k, (const __half*)v, (__half*)kv_data,
                                     kv_indices, num_kv_pages, kv_last_page_len, (__half*)meta_data, meta_indices, num_meta_pages,
                                     meta_last_page_len, row_vecs, page_size);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

extern "C" int quest_append_prefill(const void* k, const void* v, uint32_t append_len, void* kv_data,
                                    const int32_t* kv_indices, uint32_t num_kv_pages,
                                    uint32_t kv_last_page_len, void* meta_data, const int32_t* meta_indices,
                                    uint32_t num_meta_pages, uint32_t num_heads, uint32_t head_dim,
                                    uint32_t page_size, void* stream) {
  if (!k || !v || !kv_data || !kv_indices || !meta_data || !meta_indices) return QUEST_ERR_BAD_ARG;
  if (bad_dims(num_heads, head_dim, page_size)) return QUEST_ERR_UNSUPPORTED;
  if (!num_kv_pages || kv_last_page_len < 1 || kv_last_page_len > page_size) return QUEST_ERR_BAD_ARG;
  const uint64_t seq_len = (uint64_t)(num_kv_pages - 1) * page_size + kv_last_page_len;
  if (append_len == 0) return QUEST_OK;
  if (append_len > seq_len) return QUEST_ERR_BAD_ARG;
  if ((uint64_t)num_meta_pages * page_size < num_kv_pages) return QUEST_ERR_BAD_ARG;
  const uint32_t first_page = (uint32_t)((seq_len - append_len) / page_size);
  const uint32_t row_vecs = num_heads * head_dim / 8;
  const uint32_t threads = 256;
  dim3 grid(num_kv_pages - first_page, (row_vecs + threads - 1) / threads);
  quest::append_prefill_kernel<8><<<grid, threads, 0, (cudaStream_t)stream>>>(
      (const __half*)k, (const __half*)v, append_len, (__half*)kv_data, kv_indices, num_kv_pages,
      kv_last_page_len, (__half*)meta_data, meta_indices, row_vecs, page_size);
  QUEST_RETURN_LAUNCH();
}

extern "C" int quest_rope_append_prefill(void* q, void* k, const void* v, uint32_t append_len, void* kv_data,
                                         const int32_t* kv_indices, uint32_t num_kv_pages, uint32_t kv_last_page_len,
                                         void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                                         uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                                         float rope_scale, float rope_theta, uint32_t flags, void* stream) {
  if (!q || !k || !v || !kv_data || !kv_indices || !meta_data || !meta_indices || !num_qo_heads) return QUEST_ERR_BAD_ARG;
  if (bad_dims(num_kv_heads, head_dim, page_size)) return QUEST_ERR_UNSUPPORTED;
  if (!num_kv_pages || kv_last_page_len < 1 || kv_last_page_len > page_size) return QUEST_ERR_BAD_ARG;
  if (!(rope_scale > 0.f) || !(rope_theta > 0.f)) return QUEST_ERR_BAD_ARG;
  const uint64_t seq_len = (uint64_t)(num_kv_pages - 1) * page_size + kv_last_page_len;
  if (append_len == 0) return QUEST_OK;
  if (append_len > seq_len) return QUEST_ERR_BAD_ARG;
  if ((uint64_t)num_meta_pages * page_size < num_kv_pages) return QUEST_ERR_BAD_ARG;
  const uint32_t first_page = (uint32_t)((seq_len - append_len) / page_size);
  const uint32_t n_page_ctas = num_kv_pages - first_page;
  const uint32_t pairs = num_kv_heads * head_dim / 16;
  uint32_t threads = (pairs + 31) / 32 * 32;
  if (threads > 256) threads = 256;
  if (threads < head_dim / 16 * 4) threads = 64;  // the q CTAs want a few rows per pass
  const uint64_t q_rows = (uint64_t)append_len * num_qo_heads;
  const uint32_t rows_per_cta = threads / (head_dim / 16);
  uint64_t q_ctas = (q_rows + rows_per_cta - 1) / rows_per_cta;
  if (q_ctas > 148 * 8) q_ctas = 148 * 8;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(n_page_ctas + (uint32_t)q_ctas);
  cfg.blockDim = dim3(threads);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const float rs = 1.f / rope_scale, rt = 1.f / rope_theta;
  const uint32_t wb = flags & QUEST_PF_WRITE_BACK_K;
  cudaError_t e;
#define QUEST_RAP(DD)                                                                                                    \
  e = cudaLaunchKernelEx(&cfg, quest::rope_append_prefill_kernel<DD, 2>, (__half*)q, (__half*)k, (const __half*)v,       \
                         append_len, (__half*)kv_data, kv_indices, num_kv_pages, kv_last_page_len, (__half*)meta_data,   \
                         meta_indices, num_qo_heads, num_kv_heads, page_size, n_page_ctas, rs, rt, wb)
  switch (head_dim) {
    case 64: QUEST_RAP(64); break;
    case 128: QUEST_RAP(128); break;
    default: QUEST_RAP(256); break;
  }
#undef QUEST_RAP
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

// fp8 metadata (no counterpart in the reference): e4m3 copies of the metadata entries of KV pages
// [first_kv_page, end_kv_page), max rounded up and min rounded down.  Call it after quest_append_prefill /
// quest_rope_append_prefill for the pages they touched (quest_decode_layer_step_fp8 keeps the copy current from then on).
extern "C" int quest_metadata_to_fp8(const void* meta_data, const int32_t* meta_indices, uint32_t first_kv_page,
                                     uint32_t end_kv_page, void* meta8_data, uint32_t num_heads, uint32_t head_dim,
                                     uint32_t page_size, void* stream) {
  if (!meta_data || !meta_indices || !meta8_data) return QUEST_ERR_BAD_ARG;
  if (bad_dims(num_heads, head_dim, page_size)) return QUEST_ERR_UNSUPPORTED;
  if (first_kv_page > end_kv_page) return QUEST_ERR_BAD_ARG;
  if (first_kv_page == end_kv_page) return QUEST_OK;
  const uint32_t row_vecs = num_heads * head_dim / 8;
  dim3 grid(end_kv_page - first_kv_page, (row_vecs + 255) / 256);
  quest::metadata_to_fp8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const __half*)meta_data, meta_indices, first_kv_page,
                                                                       (uint8_t*)meta8_data, row_vecs, page_size);
  QUEST_RETURN_LAUNCH();
}
