// Thin C-ABI harness around the REFERENCE's own CUDA kernels, compiled for sm_100a from the sources where they
// lie under /root/reference (nothing is copied).  TEST / BASELINE INFRASTRUCTURE ONLY: the output
// (oracle/_ref/libquest_ref.so) is loaded by tests/test_vs_ref_kernels.py as the GPU oracle and by bench.py /
// tools as the "kernel to beat"; the product never links or loads it.
// Mirrors what quest/ops/csrc/{page,estimate,approx_attn}.cu do with torch tensors (NHD layout, half, int32 ids,
// batch 1); the raft top-k cannot be built here (rmm/spdlog are not vendored).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "decode/decode_attn.cuh"
#include "decode/decode_handler.cuh"
#include "decode/decode_page.cuh"

using namespace flashinfer;
using kv_t = paged_kv_t<PageStorage::kIndices, QKVLayout::kNHD, half, int32_t>;

static BatchDecodeHandler g_handler;

extern "C" {

int ref_append_decode(void* k, void* v, void* kv_data, int32_t* kv_indices, int32_t* kv_indptr, uint32_t kv_last_len,
                      int32_t kv_last_idx, void* meta_data, int32_t* meta_indices, int32_t* meta_indptr,
                      uint32_t meta_last_len, int32_t meta_last_idx, uint32_t H, uint32_t D, uint32_t P) {
  kv_t kv(H, P, D, 1, 0, kv_last_len, kv_last_idx, (half*)kv_data, kv_indices, kv_indptr);
  kv_t meta(H, P, D, 1, 0, meta_last_len, meta_last_idx, (half*)meta_data, meta_indices, meta_indptr);
  return (int)AppendPagedKVCacheDecode<PageStorage::kIndices, QKVLayout::kNHD, half, int32_t>(kv, meta, (half*)k, (half*)v, nullptr);
}

int ref_append_prefill(void* k, void* v, int32_t* append_indptr, void* kv_data, int32_t* kv_indices, int32_t* kv_indptr,
                       uint32_t kv_last_len, int32_t kv_last_idx, void* meta_data, int32_t* meta_indices,
                       int32_t* meta_indptr, uint32_t meta_last_len, int32_t meta_last_idx, uint32_t H, uint32_t D,
                       uint32_t P) {
  kv_t kv(H, P, D, 1, 0, kv_last_len, kv_last_idx, (half*)kv_data, kv_indices, kv_indptr);
  kv_t meta(H, P, D, 1, 0, meta_last_len, meta_last_idx, (half*)meta_data, meta_indices, meta_indptr);
  return (int)AppendPagedKVCachePrefill<PageStorage::kIndices, QKVLayout::kNHD, half, int32_t>(kv, meta, (half*)k, (half*)v,
                                                                                             append_indptr, nullptr);
}

int ref_estimate(void* q, void* o, void* meta_data, int32_t* meta_indices, int32_t* meta_indptr, uint32_t meta_last_len,
                 int32_t meta_last_idx, uint32_t H, uint32_t D, uint32_t P) {
  kv_t meta(H, P, D, 1, 0, meta_last_len, meta_last_idx, (half*)meta_data, meta_indices, meta_indptr);
  return (int)MaxPossibleSampleWithPagedKVCache<PageStorage::kIndices, QKVLayout::kNHD, half, half, int32_t>(
      (half*)q, meta, (half*)o, H, RotaryMode::kNone);
}

int ref_decode_begin(int32_t* indptr, uint32_t Hq, uint32_t Hkv, uint32_t D, uint32_t P) {
  return (int)g_handler.BeginForward<PageStorage::kIndices, QKVLayout::kNHD, half, half, int32_t>(indptr, 1, Hq, Hkv, D, P,
                                                                                                 RotaryMode::kNone);
}

int ref_decode_forward(void* q, void* o, void* kv_data, int32_t* indices, int32_t* indptr, uint32_t page_budget,
                       uint32_t last_len, int32_t last_idx, uint32_t Hq, uint32_t Hkv, uint32_t D, uint32_t P) {
  kv_t kv(Hkv, P, D, 1, page_budget, last_len, last_idx, (half*)kv_data, indices, indptr);
  return (int)BatchDecodeWithPagedKVCacheWrapper<PageStorage::kIndices, QKVLayout::kNHD, half, half, int32_t>(
      &g_handler, (half*)q, kv, (half*)o, nullptr, Hq, RotaryMode::kNone, 1.f, 1e4f, nullptr);
}

int ref_decode_end(void) {
  g_handler.EndForward();
  return 0;
}

}  // extern "C"
