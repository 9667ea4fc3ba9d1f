// Library-level entry points: identification, launch counter, device-error words, test switches, and the layer step
// that chains estimate -> top-k -> sparse decode with programmatic dependent launch (one call per layer per token;
// reference call sites quest/models/QuestAttention.py:135-158).
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

namespace quest {
unsigned long long g_launch_count = 0;
int estimate_dispatch(const void* q, void* o, uint32_t len, const void* meta_data, const int32_t* meta_indices,
                      uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                      uint32_t flags, const AppendArgs& ap, cudaStream_t s, const BatchArgs& ba = BatchArgs{},
                      uint32_t batch = 1, uint32_t* hist = nullptr, uint32_t* epoch = nullptr);  // estimate.cu
bool topk_hist_covers(uint32_t max_len);  // topk.cu
int topk_from_hist(const void* scores, uint32_t* hist, const int32_t* in_idx, uint32_t in_stride, int32_t* out_idx,
                   uint32_t num_heads, uint32_t max_len, uint32_t k, const BatchArgs& ba, uint32_t batch, cudaStream_t s,
                   const Handoff& ho = Handoff{}, const HistOut& xo = HistOut{});  // topk.cu
bool decode_handoff_fits(uint32_t n_sel, uint32_t D, uint32_t ppc);  // decode.cu
int decode_run_handoff(const void* q, void* o, const void* kv, const int32_t* indices, uint32_t n_sel, uint32_t last_page_idx,
                       uint32_t last_page_len, uint32_t Hq, uint32_t Hkv, uint32_t D, uint32_t P, uint32_t ppc, void* ws,
                       cudaStream_t s);  // decode.cu
// Page-id hand-over of the layer step (common.cuh, Handoff): on when the step selects from the bin totals, the plan fits and the
// workspace holds it; QUEST_DBG_NO_HANDOFF switches it off (A/B and tests).  Rows of up to 2048 scored pages only: measured
// 18.63 -> 18.48 us per layer step at 2047 pages, but 30.75 -> 31.1 at 4095 and 52.6 -> 53.2 at 8191, where the decode CTAs
// would sit polling on their SMs through a selection of several microseconds (profiles/r02_b_experiments.md).
static Handoff step_handoff(void* workspace, size_t workspace_bytes, uint32_t* hist, uint32_t len, uint32_t k, uint32_t Hq, uint32_t D,
                            uint32_t ppc) {
  Handoff ho = {};
  if (!hist || len > 2048 || debug_option(QUEST_DBG_NO_HANDOFF) || !pdl_enabled() || !decode_handoff_fits(k, D, ppc)) return ho;
  const uint32_t nchunks = (k + 1 + ppc - 1) / ppc;
  if (workspace_bytes < ws_bytes(Hq, D, nchunks)) return ho;  // quest_decode_run would refuse it anyway
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  ho.epoch = reinterpret_cast<uint32_t*>(ws + kWsEpochOff);
  ho.slots = ws + ws_o_off(Hq, nchunks);
  ho.nchunks = nchunks, ho.ppc = ppc, ho.slot_bytes = D * (uint32_t)sizeof(float);
  return ho;
}
// the layer step's bin totals live in its decode workspace (common.cuh); nullptr: the workspace has no room for them
static uint32_t* step_hist(void* workspace, size_t workspace_bytes, uint32_t Hq, uint32_t max_len) {
  if (!workspace || workspace_bytes < ws_md_off(Hq) || !topk_hist_covers(max_len)) return nullptr;
  return reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(workspace) + ws_hist_off(Hq));
}
int topk_batch(const void* scores, int32_t* out_idx, uint32_t num_heads, uint32_t max_len, uint32_t k, const BatchArgs& ba,
               uint32_t batch, cudaStream_t s);  // topk.cu
int decode_batch(const void* q, void* o, const int32_t* indices, uint32_t n_sel, uint32_t Hq, uint32_t Hkv, uint32_t D,
                 uint32_t P, uint32_t ppc, void* ws, const BatchArgs& ba, uint32_t batch, cudaStream_t s);  // decode.cu

static int g_debug_options[QUEST_DBG_COUNT] = {};
int debug_option(int option) { return (option >= 0 && option < QUEST_DBG_COUNT) ? g_debug_options[option] : 0; }

// ---- sticky device-error words: mapped pinned host memory, so the host reads them without touching the device
static unsigned int* g_err_host = nullptr;
static unsigned int* g_err_dev = nullptr;
static std::once_flag g_err_once;
unsigned int* device_error_word() {
  std::call_once(g_err_once, [] {
    void* h = nullptr;
    if (cudaHostAlloc(&h, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
      (void)cudaGetLastError();
      return;
    }
    for (int i = 0; i < 16; ++i) reinterpret_cast<volatile unsigned int*>(h)[i] = 0u;
    void* d = nullptr;
    if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) {
      (void)cudaGetLastError();
      cudaFreeHost(h);
      return;
    }
    g_err_host = reinterpret_cast<unsigned int*>(h);
    g_err_dev = reinterpret_cast<unsigned int*>(d);
  });
  return g_err_dev;
}
unsigned int device_error_peek() {
  if (!g_err_host) return 0u;
  unsigned int bits = 0;
  for (int i = 0; i < QUEST_DEVERR_COUNT; ++i)
    if (reinterpret_cast<volatile unsigned int*>(g_err_host)[i]) bits |= 1u << i;
  return bits;
}
}  // namespace quest

extern "C" uint32_t quest_device_error(void) {
  quest::device_error_word();  // allocates on first use (call once at load time, outside any stream capture)
  return quest::device_error_peek();
}
extern "C" void quest_device_error_clear(void) {
  if (quest::g_err_host)
    for (int i = 0; i < QUEST_DEVERR_COUNT; ++i) reinterpret_cast<volatile unsigned int*>(quest::g_err_host)[i] = 0u;
}
extern "C" int quest_debug_set_option(int option, int value) {
  if (option < 0 || option >= QUEST_DBG_COUNT) return QUEST_ERR_BAD_ARG;
  const int old = quest::g_debug_options[option];
  quest::g_debug_options[option] = value;
  return old;
}

extern "C" const char* quest_version(void) { return "quest_b200 0.1.0 sm_100a"; }
extern "C" uint64_t quest_launch_count(void) { return quest::g_launch_count; }
extern "C" void quest_launch_count_reset(void) { quest::g_launch_count = 0; }

// estimate + top-k + sparse decode without the append (the three calls QuestAttention.forward makes after append_kv,
// quest/models/QuestAttention.py:135-158), chained with programmatic dependent launch.
extern "C" int quest_sparse_decode_step(const void* q, void* o, const void* kv_data, const int32_t* kv_indices,
                                        uint32_t len, uint32_t last_page_idx, uint32_t last_page_len,
                                        const void* meta_data, const int32_t* meta_indices, void* scores,
                                        int32_t* topk_idx, uint32_t k, uint32_t num_qo_heads,
                                        uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                                        uint32_t pages_per_chunk, void* workspace, size_t workspace_bytes,
                                        void* stream) {
  if (!q || !o || !kv_data || !kv_indices || !meta_data || !meta_indices || !scores || !topk_idx) return QUEST_ERR_BAD_ARG;
  if (!num_qo_heads || !num_kv_heads || !page_size || !pages_per_chunk || num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (k > len) return QUEST_ERR_BAD_ARG;
  if (k == 0 || len == 0)
    return quest_decode_run(q, o, nullptr, kv_data, topk_idx, k, k, last_page_idx, last_page_len, num_qo_heads, num_kv_heads,
                            head_dim, page_size, pages_per_chunk, workspace, workspace_bytes, stream);
  // the estimate kernel counts the scores into per-head bin totals, from which the top-k kernel selects in one pass
  if (quest::device_error_peek() & QUEST_DEVERR_STEP) return QUEST_ERR_DEVICE;  // an earlier step lost its page-id hand-over
  uint32_t* hist = quest::step_hist(workspace, workspace_bytes, num_qo_heads, len);
  quest::Handoff ho = quest::step_handoff(workspace, workspace_bytes, hist, len, k, num_qo_heads, head_dim, pages_per_chunk);
  int rc = quest::estimate_dispatch(q, scores, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size,
                                    QUEST_EST_META_STABLE | QUEST_EST_META_ORDERED, quest::AppendArgs{}, (cudaStream_t)stream,
                                    quest::BatchArgs{}, 1, hist, ho.epoch);
  if (rc == QUEST_ERR_UNSUPPORTED && hist) {  // shapes only the generic estimate kernel takes
    hist = nullptr;
    ho = quest::Handoff{};
    rc = quest_estimate(q, scores, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size,
                        QUEST_EST_META_STABLE | QUEST_EST_META_ORDERED, stream);
  }
  if (rc) return rc;
  // payload ids = the sequence's page table, shared by all heads (row stride 0)
  rc = hist ? quest::topk_from_hist(scores, hist, kv_indices, 0, topk_idx, num_qo_heads, len, k, quest::BatchArgs{}, 1, (cudaStream_t)stream, ho)
            : quest_topk(scores, kv_indices, 0, nullptr, topk_idx, num_qo_heads, len, k, stream);
  if (rc) return rc;
  if (ho.epoch)
    return quest::decode_run_handoff(q, o, kv_data, topk_idx, k, last_page_idx, last_page_len, num_qo_heads, num_kv_heads, head_dim,
                                     page_size, pages_per_chunk, workspace, (cudaStream_t)stream);
  return quest_decode_run(q, o, nullptr, kv_data, topk_idx, k, k, last_page_idx, last_page_len, num_qo_heads,
                          num_kv_heads, head_dim, page_size, pages_per_chunk, workspace, workspace_bytes, stream);
}

// One decode step of one layer: append_kv + estimate + top-k + sparse decode (the four calls of
// quest/models/QuestAttention.py:100-150 in the reference).  The append rides on one extra CTA of the estimate launch, so the
// step is three launches.
static int layer_step_impl(const void* q, const void* k_new, const void* v_new, void* o, void* kv_data,
                           const int32_t* kv_indices, uint32_t num_kv_pages, uint32_t kv_last_page_idx,
                           uint32_t kv_last_page_len, void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                           uint32_t meta_last_page_len, void* scores, int32_t* topk_idx, uint32_t k,
                           uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                           uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                           size_t workspace_bytes, void* stream, void* meta8 /* e4m3 metadata pool to score from, or nullptr */) {
  if (!q || !k_new || !v_new || !o || !kv_data || !kv_indices || !meta_data || !meta_indices) return QUEST_ERR_BAD_ARG;
  if (!num_kv_pages || !num_meta_pages || !num_qo_heads || !num_kv_heads || !page_size || !pages_per_chunk) return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (kv_last_page_len < 1 || kv_last_page_len > page_size || meta_last_page_len < 1 || meta_last_page_len > page_size)
    return QUEST_ERR_BAD_ARG;
  if (num_meta_pages != (num_kv_pages + page_size - 1) / page_size ||
      meta_last_page_len != (num_kv_pages - 1) % page_size + 1)
    return QUEST_ERR_BAD_ARG;  // one metadata entry per KV page
  const uint32_t len = num_kv_pages - 1;  // scored pages: all but the last
  if (k > len || (k && (!scores || !topk_idx))) return QUEST_ERR_BAD_ARG;
  quest::AppendArgs ap = {(const __half*)k_new, (const __half*)v_new, (__half*)kv_data, kv_indices, (__half*)meta_data,
                          num_kv_pages, kv_last_page_len, num_meta_pages, meta_last_page_len, (uint8_t*)meta8};
  int rc;
  quest::Handoff ho = {};
  if (k == 0) {  // nothing to select: attend to the last page only
    rc = quest_append_decode(k_new, v_new, kv_data, kv_indices, num_kv_pages, kv_last_page_len, meta_data, meta_indices,
                             num_meta_pages, meta_last_page_len, num_kv_heads, head_dim, page_size, stream);
    if (rc == QUEST_OK && meta8)  // keep the e4m3 copy of the entry just updated current
      rc = quest_metadata_to_fp8(meta_data, meta_indices, num_kv_pages - 1, num_kv_pages, meta8, num_kv_heads, head_dim, page_size, stream);
  } else {
    // the estimate kernel counts the scores into per-head bin totals, from which the top-k kernel selects in one pass
    if (quest::device_error_peek() & QUEST_DEVERR_STEP) return QUEST_ERR_DEVICE;  // an earlier step lost its page-id hand-over
    uint32_t* hist = quest::step_hist(workspace, workspace_bytes, num_qo_heads, len);
    ho = quest::step_handoff(workspace, workspace_bytes, hist, len, k, num_qo_heads, head_dim, pages_per_chunk);
    const uint32_t fl = QUEST_EST_META_STABLE | QUEST_EST_META_ORDERED | (meta8 ? quest::kEstMetaFp8 : 0u);
    const void* scored = meta8 ? meta8 : meta_data;  // the append CTA still updates the fp16 pool (ap.meta) and its e4m3 copy
    rc = quest::estimate_dispatch(q, scores, len, scored, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size, fl, ap,
                                  (cudaStream_t)stream, quest::BatchArgs{}, 1, hist, ho.epoch);
    if (rc == QUEST_ERR_UNSUPPORTED && hist && !meta8) {  // shapes only the generic estimate kernel takes (it launches the append itself)
      hist = nullptr;
      ho = quest::Handoff{};
      rc = quest::estimate_dispatch(q, scores, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size,
                                    QUEST_EST_META_STABLE | QUEST_EST_META_ORDERED, ap, (cudaStream_t)stream);
    }
    if (rc) return rc;
    // payload ids = the sequence's page table, shared by all heads (row stride 0)
    rc = hist ? quest::topk_from_hist(scores, hist, kv_indices, 0, topk_idx, num_qo_heads, len, k, quest::BatchArgs{}, 1, (cudaStream_t)stream, ho)
              : quest_topk(scores, kv_indices, 0, nullptr, topk_idx, num_qo_heads, len, k, stream);
  }
  if (rc) return rc;
  if (ho.epoch)
    return quest::decode_run_handoff(q, o, kv_data, topk_idx, k, kv_last_page_idx, kv_last_page_len, num_qo_heads, num_kv_heads,
                                     head_dim, page_size, pages_per_chunk, workspace, (cudaStream_t)stream);
  return quest_decode_run(q, o, nullptr, kv_data, topk_idx, k, k, kv_last_page_idx, kv_last_page_len, num_qo_heads, num_kv_heads,
                          head_dim, page_size, pages_per_chunk, workspace, workspace_bytes, stream);
}

extern "C" int quest_decode_layer_step(const void* q, const void* k_new, const void* v_new, void* o, void* kv_data,
                                       const int32_t* kv_indices, uint32_t num_kv_pages, uint32_t kv_last_page_idx,
                                       uint32_t kv_last_page_len, void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                                       uint32_t meta_last_page_len, void* scores, int32_t* topk_idx, uint32_t k,
                                       uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                                       uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  return layer_step_impl(q, k_new, v_new, o, kv_data, kv_indices, num_kv_pages, kv_last_page_idx, kv_last_page_len, meta_data,
                         meta_indices, num_meta_pages, meta_last_page_len, scores, topk_idx, k, num_qo_heads, num_kv_heads,
                         head_dim, page_size, pages_per_chunk, workspace, workspace_bytes, stream, nullptr);
}

// The same step scoring from the e4m3 copy of the metadata (flag-gated variant, no counterpart in
// the reference): the estimate streams half the bytes; the append CTA updates the fp16 entry of the last page and its e4m3 copy.
extern "C" int quest_decode_layer_step_fp8(const void* q, const void* k_new, const void* v_new, void* o, void* kv_data,
                                           const int32_t* kv_indices, uint32_t num_kv_pages, uint32_t kv_last_page_idx,
                                           uint32_t kv_last_page_len, void* meta_data, void* meta8_data, const int32_t* meta_indices,
                                           uint32_t num_meta_pages, uint32_t meta_last_page_len, void* scores, int32_t* topk_idx,
                                           uint32_t k, uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                                           uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                           size_t workspace_bytes, void* stream) {
  if (!meta8_data) return QUEST_ERR_BAD_ARG;
  return layer_step_impl(q, k_new, v_new, o, kv_data, kv_indices, num_kv_pages, kv_last_page_idx, kv_last_page_len, meta_data,
                         meta_indices, num_meta_pages, meta_last_page_len, scores, topk_idx, k, num_qo_heads, num_kv_heads,
                         head_dim, page_size, pages_per_chunk, workspace, workspace_bytes, stream, meta8_data);
}

// The layer step for a batch of sequences, driven by per-sequence records in device memory (see include/quest_b200.h).
extern "C" int quest_decode_layer_step_batch(const void* q, const void* k_new, const void* v_new, void* o,
                                             const QuestSeqDesc* descs, uint32_t batch, size_t kv_layer_offset,
                                             size_t meta_layer_offset, uint32_t max_kv_pages, void* scores,
                                             size_t scores_seq_stride, int32_t* topk_idx, uint32_t k,
                                             uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                                             uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                             size_t workspace_seq_bytes, void* stream) {
  if (!q || !k_new || !v_new || !o || !descs || !batch || !scores || !topk_idx || !workspace) return QUEST_ERR_BAD_ARG;
  if (!num_qo_heads || !num_kv_heads || !page_size || !pages_per_chunk || num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (max_kv_pages < 2 || k == 0 || batch > 65535) return QUEST_ERR_BAD_ARG;
  const uint32_t max_len = max_kv_pages - 1;
  if (scores_seq_stride < (size_t)num_qo_heads * max_len) return QUEST_ERR_BAD_ARG;
  const uint32_t nchunks = (k + 1 + pages_per_chunk - 1) / pages_per_chunk;
  if (workspace_seq_bytes % 16 || workspace_seq_bytes < quest_decode_workspace_bytes(num_qo_heads, head_dim, nchunks))
    return QUEST_ERR_WORKSPACE;
  quest::BatchArgs ba = {reinterpret_cast<const quest::SeqDesc*>(descs), kv_layer_offset, meta_layer_offset,
                         scores_seq_stride, workspace_seq_bytes};
  // the append rides on the estimate launch; its pointers and lengths are taken from the records inside the kernel
  quest::AppendArgs ap = {(const __half*)k_new, (const __half*)v_new, nullptr, nullptr, nullptr, 0, 0, 0, 0, nullptr};
  cudaStream_t s = (cudaStream_t)stream;
  const uint32_t flags = QUEST_EST_META_STABLE | QUEST_EST_META_ORDERED;
  uint32_t* hist = quest::step_hist(workspace, workspace_seq_bytes, num_qo_heads, max_len);
  int rc = quest::estimate_dispatch(q, scores, max_len, nullptr, nullptr, num_qo_heads, num_kv_heads, head_dim, page_size,
                                    flags, ap, s, ba, batch, hist);
  if (rc) return rc;
  rc = hist ? quest::topk_from_hist(scores, hist, nullptr, 0, topk_idx, num_qo_heads, max_len, k, ba, batch, s)
            : quest::topk_batch(scores, topk_idx, num_qo_heads, max_len, k, ba, batch, s);
  if (rc) return rc;
  return quest::decode_batch(q, o, topk_idx, k, num_qo_heads, num_kv_heads, head_dim, page_size, pages_per_chunk, workspace,
                             ba, batch, s);
}

// The prompt's attention layer as one call: fused RoPE + append + metadata, then the causal paged prefill, chained with
// programmatic dependent launch (the prefill kernel's barrier / tensor-memory set-up runs under the producer's tail).
extern "C" int quest_prefill_layer_step(void* q, void* k, const void* v, void* o, uint32_t append_len, void* kv_data,
                                        const int32_t* kv_indices, uint32_t num_kv_pages, uint32_t kv_last_page_len,
                                        void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                                        uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                                        float rope_scale, float rope_theta, uint32_t flags, void* stream) {
  if (!o) return QUEST_ERR_BAD_ARG;
  int rc = quest_rope_append_prefill(q, k, v, append_len, kv_data, kv_indices, num_kv_pages, kv_last_page_len, meta_data,
                                     meta_indices, num_meta_pages, num_qo_heads, num_kv_heads, head_dim, page_size, rope_scale,
                                     rope_theta, flags, stream);
  if (rc != QUEST_OK) return rc;
  return quest_prefill(q, o, append_len, kv_data, kv_indices, num_kv_pages, kv_last_page_len, 1, num_qo_heads, num_kv_heads,
                       head_dim, page_size, stream);
}

// estimate (optionally carrying the step's append_kv) + selection from the bin totals, with the selected SCORES as well as the
// payload ids and with output row strides: phase 1 of the sequence-sharded step (quest_b200.parallel.SeqShardedDecode.candidates),
// where a rank's best pages travel as scores and its ids go into its columns of the global payload row.  Two launches chained
// with programmatic dependent launch instead of append + estimate + top-k + two copies.  No counterpart in the reference (it
// has no multi-GPU path); the kernels are those of quest_decode_layer_step.
extern "C" int quest_estimate_select(const void* q, const void* k_new, const void* v_new, void* kv_data, const int32_t* kv_indices,
                                     uint32_t num_kv_pages, uint32_t kv_last_page_len, void* meta_data,
                                     const int32_t* meta_indices, uint32_t num_meta_pages, uint32_t meta_last_page_len,
                                     uint32_t len, void* scores, void* out_val, uint32_t out_val_stride, int32_t* out_idx,
                                     uint32_t out_idx_stride, uint32_t k, uint32_t num_qo_heads, uint32_t num_kv_heads,
                                     uint32_t head_dim, uint32_t page_size, uint32_t flags, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  if (!q || !meta_data || !meta_indices || !scores || !out_idx || !kv_indices) return QUEST_ERR_BAD_ARG;
  if (!num_qo_heads || !num_kv_heads || !page_size || num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (k == 0 || k > len || len > num_kv_pages) return QUEST_ERR_BAD_ARG;
  if ((out_val_stride && out_val_stride < k) || (out_idx_stride && out_idx_stride < k)) return QUEST_ERR_BAD_ARG;
  quest::AppendArgs ap = {};
  if (k_new != nullptr) {
    if (!v_new || !kv_data || !num_kv_pages || !num_meta_pages) return QUEST_ERR_BAD_ARG;
    if (kv_last_page_len < 1 || kv_last_page_len > page_size || meta_last_page_len < 1 || meta_last_page_len > page_size)
      return QUEST_ERR_BAD_ARG;
    ap = {(const __half*)k_new, (const __half*)v_new, (__half*)kv_data, kv_indices, (__half*)meta_data,
          num_kv_pages, kv_last_page_len, num_meta_pages, meta_last_page_len, nullptr};
  }
  uint32_t* hist = quest::step_hist(workspace, workspace_bytes, num_qo_heads, len);
  if (!hist) return QUEST_ERR_UNSUPPORTED;  // no room for the bin totals / rows beyond their reach: the caller uses the separate ops
  int rc = quest::estimate_dispatch(q, scores, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size,
                                    flags & (QUEST_EST_META_ORDERED | QUEST_EST_META_STABLE), ap, (cudaStream_t)stream,
                                    quest::BatchArgs{}, 1, hist);
  if (rc) return rc;  // QUEST_ERR_UNSUPPORTED before anything was launched: shapes only the generic estimate kernel takes
  const quest::HistOut xo = {(uint16_t*)out_val, out_val_stride, out_idx_stride};
  return quest::topk_from_hist(scores, hist, kv_indices, 0, out_idx, num_qo_heads, len, k, quest::BatchArgs{}, 1, (cudaStream_t)stream,
                               quest::Handoff{}, xo);
}

// Device-side record updater for the descriptor-driven step: advances every sequence's record by one token -- a page fills, the
// next one (whose id the host already put into the device page table when it reserved the tokens) becomes the last page, the
// metadata entry / page counters follow, the RoPE position moves on.  With it as the first node of the captured step, a run of
// pre-reserved tokens replays without any host work between tokens (quest_b200.utils.batch.BatchDecodeStep.run_ahead / advance).
// The reference rebuilds these quantities on the host for every token (quest/utils/controller.py:72-129).
namespace quest {
__global__ void seq_desc_advance_kernel(SeqDesc* __restrict__ descs, uint32_t batch, uint32_t P) {
  pdl_wait();
  const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  SeqDesc d = descs[b];
  if (d.kv_last_page_len >= P) {       // the new token opens a page ...
    d.num_kv_pages += 1;
    d.kv_last_page_len = 1;
    if (d.meta_last_page_len >= P) {   // ... whose metadata entry opens a metadata page
      d.num_meta_pages += 1;
      d.meta_last_page_len = 1;
    } else {
      d.meta_last_page_len += 1;
    }
  } else {
    d.kv_last_page_len += 1;
  }
  d.rope_pos += 1;
  descs[b] = d;
}
}  // namespace quest

extern "C" int quest_seq_desc_advance(QuestSeqDesc* descs, uint32_t batch, uint32_t page_size, void* stream) {
  if (!descs || !batch || !page_size) return QUEST_ERR_BAD_ARG;
  quest::seq_desc_advance_kernel<<<(batch + 63) / 64, 64, 0, (cudaStream_t)stream>>>(reinterpret_cast<quest::SeqDesc*>(descs), batch,
                                                                                  page_size);
  QUEST_RETURN_LAUNCH();
}
