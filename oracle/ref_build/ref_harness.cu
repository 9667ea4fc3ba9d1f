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
#include "prefill/prefill.cuh"

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

// The reference's prefill kernel in the one configuration the headline model uses (LongChat-7B: 32 MHA heads x 128,
// page 16, causal, long query chunk -> num_frags_x = 2); other shapes return -1.  Timing baseline for
// tools/prefill_bench.py and one more GPU oracle.  qo_indptr_d: device int32 [0, qo_len].
int ref_prefill(void* q, void* o, uint32_t qo_len, int32_t* qo_indptr_d, void* kv_data, int32_t* kv_indices,
                int32_t* kv_indptr, uint32_t kv_last_len, int32_t kv_last_idx, uint32_t Hq, uint32_t Hkv, uint32_t D,
                uint32_t P) {
  if (Hq != Hkv || D != 128 || P != 16 || qo_len <= 64) return -1;
  kv_t kv(Hkv, P, D, 1, 0, kv_last_len, kv_last_idx, (half*)kv_data, kv_indices, kv_indptr);
  int32_t qo_indptr_h[2] = {0, (int32_t)qo_len};
  auto [frags_x, num_qo_tiles, request_h, tile_h] = split_qo_indptr(qo_indptr_h, 1, 1);
  if (frags_x != 2) return -1;
  static int32_t* idx_d = nullptr;
  static size_t idx_cap = 0;
  static uint32_t planned_qo = 0;  // the tile lists are uploaded once per qo_len (the reference uploads them every call)
  if (planned_qo != qo_len) {
    if (idx_cap < request_h.size()) {
      if (idx_d) cudaFree(idx_d);
      idx_cap = request_h.size() * 2;
      if (cudaMalloc(&idx_d, sizeof(int32_t) * 2 * idx_cap) != cudaSuccess) return -2;
    }
    cudaMemcpy(idx_d, request_h.data(), sizeof(int32_t) * request_h.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(idx_d + idx_cap, tile_h.data(), sizeof(int32_t) * tile_h.size(), cudaMemcpyHostToDevice);
    planned_qo = qo_len;
  }
  return (int)BatchPrefillWithPagedKVCacheDispatched<PageStorage::kIndices, QKVLayout::kNHD, 2, 16, 1, 128, RotaryMode::kNone,
                                                     false, true, half, half, int32_t>(
      (half*)q, idx_d, idx_d + idx_cap, qo_indptr_d, kv, (half*)o, nullptr, nullptr, num_qo_tiles, 1.f, 1e4f, nullptr);
}

}  // extern "C"
