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
  quest::append_decode_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(
      (const __half*)k, (const __half*)v, (__half*)kv_data, kv_indices, num_kv_pages, kv_last_page_len,
      (__half*)meta_data, meta_indices, num_meta_pages, meta_last_page_len, row_vecs, page_size);
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
