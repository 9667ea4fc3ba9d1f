/*
 * quest_b200.h -- C ABI of libquest_b200.so, the sm_100a implementation of Quest's
 * query-aware sparse decode-attention path.
 *
 * Every entry point replaces one function of the reference's pybind module `quest._kernels`
 * (reference: quest/ops/csrc/bsk_ops.h:23-117, bsk_ops.cu:4-20).  The boundary is plain C:
 * raw device pointers, explicit sizes, a CUDA stream as `void*`; no torch types, no exceptions.
 *
 * Conventions
 *   - return value: 0 = success, >0 = a cudaError_t from the launch, <0 = QUEST_ERR_* below.
 *   - all tensors are fp16 (the reference dispatches Half only,
 *     quest/ops/csrc/pytorch_extension_utils.h:25-35) unless stated; page ids are int32.
 *   - paged pools use the reference's NHD layout (quest/utils/kv_cache.py:20-23):
 *       pool[page][2][page_size][num_kv_heads][head_dim]; slab 0 = K (metadata: per-dim max),
 *       slab 1 = V (metadata: per-dim min)   (kernels/include/decode/decode_page.cuh:196-239,443-446)
 *   - `stream` is a cudaStream_t; kernels are enqueued, never synchronised.
 *   - head_dim in {64, 128, 256}; page_size >= 1; num_qo_heads % num_kv_heads == 0.
 */
#ifndef QUEST_B200_H_
#define QUEST_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QUEST_OK 0
#define QUEST_ERR_BAD_ARG (-1)      /* null pointer / zero size / inconsistent sizes            */
#define QUEST_ERR_UNSUPPORTED (-2)  /* head_dim / group size / length outside the built variants */
#define QUEST_ERR_WORKSPACE (-3)    /* caller-provided workspace too small                       */
#define QUEST_ERR_DEVICE (-4)       /* an earlier kernel of this library gave up on a bounded wait: its
                                       output was not valid (see quest_device_error)             */
#define QUEST_ERR_SCRATCH_LIMIT (-5) /* quest_decode_run_static: the plan needs more than the built-in scratch array */

/* Library identification: "quest_b200 <version> sm_100a". */
const char* quest_version(void);

/* ---- RoPE ------------------------------------------------------------------------------------
 * Replaces apply_rope_in_place (bsk_ops.h:23-27, page.cu:212-253, decode_page.cuh:644-728).
 * q: [seq_len, num_qo_heads, head_dim], k: [seq_len, num_kv_heads, head_dim], rotated in place with
 * Llama rotate-half at positions past_kv_len + i. */
int quest_rope_inplace(void* q, void* k, uint32_t seq_len, uint32_t past_kv_len,
                       uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                       float rope_scale, float rope_theta, void* stream);

/* ---- RMSNorm ---------------------------------------------------------------------------------
 * Replaces rms_norm_forward (bsk_ops.h:29-32, rms_norm.cu:160-212). in/out: [rows, cols]. */
int quest_rms_norm(const void* in, const void* weight, void* out, uint32_t rows, uint32_t cols,
                   float epsilon, void* stream);

/* ---- append_kv + per-page min/max metadata -----------------------------------------------------
 * Replaces append_kv_cache_decode (bsk_ops.h:64-76, page.cu:6-99, decode_page.cuh:398-449).
 * Writes the new token's K,V row (k, v: [1, num_heads, head_dim]) at
 * (kv_indices[num_kv_pages-1], kv_last_page_len-1) and folds K into the metadata entry
 * (meta_indices[num_meta_pages-1], meta_last_page_len-1): max=hmax(max,k), min=hmin(min,k);
 * the entry is initialised to -/+65504 first when kv_last_page_len == 1. */
int quest_append_decode(const void* k, const void* v,
                        void* kv_data, const int32_t* kv_indices, uint32_t num_kv_pages,
                        uint32_t kv_last_page_len,
                        void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                        uint32_t meta_last_page_len,
                        uint32_t num_heads, uint32_t head_dim, uint32_t page_size, void* stream);

/* Replaces append_kv_cache_prefill (bsk_ops.h:50-62, page.cu:101-210, decode_page.cuh:471-562).
 * k, v: [append_len, num_heads, head_dim] are the LAST append_len tokens of a sequence that now
 * occupies num_kv_pages pages with kv_last_page_len tokens in the last one.  Scatters the rows into
 * their pages and writes per-page per-dim min/max into metadata entry
 * (meta_indices[j / page_size], j % page_size) for every touched KV page j; a page that already
 * held tokens before this call resumes from its stored min/max. */
int quest_append_prefill(const void* k, const void* v, uint32_t append_len,
                         void* kv_data, const int32_t* kv_indices, uint32_t num_kv_pages,
                         uint32_t kv_last_page_len,
                         void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                         uint32_t num_heads, uint32_t head_dim, uint32_t page_size, void* stream);

/* ---- estimate ----------------------------------------------------------------------------------
 * Replaces estimate_attn_score (bsk_ops.h:41-48, estimate.cu:6-84, decode_attn.cuh:137-168,245-401).
 * o[h, p] = fp16( sum_d max(q[h,d]*max[p,h/g,d], q[h,d]*min[p,h/g,d]) ) for the first `len`
 * KV pages (the caller excludes the last page: len = num_kv_pages - 1).  fp32 products/sums in the
 * reference's order (8 dims per lane sequentially, then xor-butterfly over head_dim/8 lanes), so
 * scores are bit-identical to the reference kernel.  q: [num_qo_heads, head_dim]; o: [num_qo_heads, len].
 * flags: QUEST_EST_META_STABLE promises that the kernel enqueued immediately before this one on `stream`
 * does not write metadata entries [0, len) (true in the layer step: append only touches the last page's
 * entry), which lets the metadata stream start before the programmatic dependency on that kernel resolves. */
#define QUEST_EST_META_STABLE 1u
/* QUEST_EST_META_ORDERED promises max >= min elementwise in every scored entry (true for metadata written by
 * quest_append_*): the per-element max/min pre-pass is skipped; results are identical on such data. */
#define QUEST_EST_META_ORDERED 2u
int quest_estimate(const void* q, void* o, uint32_t len,
                   const void* meta_data, const int32_t* meta_indices,
                   uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                   uint32_t page_size, uint32_t flags, void* stream);

/* ---- top-k -------------------------------------------------------------------------------------
 * Replaces topk_filtering (bsk_ops.h:34-39, topk.cu:7-46, kernels/include/topk/decode_select_k.cuh:25-62,
 * raft select_radix.cuh radix_topk_one_block_kernel).  Per head: the k largest of `len` fp16 scores
 * and their payload ids.  Deterministic: every score > tau (the k-th largest) is selected, ties at
 * tau are taken lowest position first, and the output is in ascending input position.
 * in_idx: [num_heads, len] with row stride in_idx_row_stride elements (0 = one row shared by all
 * heads); NULL = payload is the position itself.  out_val: [num_heads, k] fp16 (may be NULL),
 * out_idx: [num_heads, k].  Needs no scratch for len <= quest_topk_max_len(). */
uint32_t quest_topk_max_len(void);
int quest_topk(const void* scores, const int32_t* in_idx, uint32_t in_idx_row_stride,
               void* out_val, int32_t* out_idx, uint32_t num_heads, uint32_t len, uint32_t k,
               void* stream);

/* ---- sparse paged decode attention -------------------------------------------------------------
 * Replaces BatchDecodeWithPagedKVCachePyTorchWrapper::{BeginForward,Forward,EndForward}
 * (bsk_ops.h:88-117, approx_attn.cu:27-151, decode_handler.cuh:39-244, decode_attn.cuh:440-646,
 * 675-1005, flashinfer cascade.cuh:279-340).  The stateful handler becomes a pure host plan plus a
 * caller-owned workspace.
 *
 * quest_decode_plan: chooses pages_per_chunk for num_sel_pages selected pages (+1 forced last page)
 * so that num_qo_heads * num_chunks fills the GPU; no device access, no sync. */
int quest_decode_plan(uint32_t num_sel_pages, uint32_t num_qo_heads, uint32_t num_kv_heads,
                      uint32_t head_dim, uint32_t page_size, uint32_t num_sms,
                      uint32_t* pages_per_chunk, uint32_t* num_chunks);
/* Bytes of workspace for a plan; the workspace must be zero-filled once before first use
 * (the kernel restores its arrival counters to zero). */
size_t quest_decode_workspace_bytes(uint32_t num_qo_heads, uint32_t head_dim, uint32_t num_chunks);
/* q, o: [num_qo_heads, head_dim]; indices: [num_qo_heads, num_sel_pages] page ids with row stride
 * idx_row_stride; the page last_page_idx holding last_page_len tokens is always attended in
 * addition.  lse (fp32 [num_qo_heads], base-2, may be NULL).
 * Sequence-sharded use: an id < 0 in `indices` is a selected page that another shard holds and
 * contributes no rows; last_page_len = 0 means this shard does not hold the last page.  A head with no rows at
 * all gets o = 0 and lse = -inf, the identity of quest_merge_states. */
int quest_decode_run(const void* q, void* o, float* lse,
                     const void* kv_data, const int32_t* indices, uint32_t idx_row_stride,
                     uint32_t num_sel_pages, uint32_t last_page_idx, uint32_t last_page_len,
                     uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                     uint32_t page_size, uint32_t pages_per_chunk,
                     void* workspace, size_t workspace_bytes, void* stream);

/* Same, for one shard of a sequence-sharded step: the normalised partial output stays fp32 ([num_qo_heads, head_dim]) and
 * lse is required, so that the cross-GPU merge (quest_merge_states, flashinfer state.cuh:53-62) sees what the
 * single-GPU merge sees -- one rounding to fp16, at the very end. */
int quest_decode_run_partial(const void* q, float* o_partial, float* lse,
                             const void* kv_data, const int32_t* indices, uint32_t idx_row_stride,
                             uint32_t num_sel_pages, uint32_t last_page_idx, uint32_t last_page_len,
                             uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                             uint32_t page_size, uint32_t pages_per_chunk,
                             void* workspace, size_t workspace_bytes, void* stream);

/* ---- sparse paged decode without any allocation ----------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same computation and
 * arguments as quest_decode_run, minus the workspace: the split-KV scratch is a file-scope device array inside the
 * library (csrc/decode.cu) of quest_decode_static_scratch_bytes() = 4 MiB, which every GPU that loads the library
 * holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); it enqueues
 * one small cudaMemsetAsync (the words that must start at zero) and the kernel.
 * Limits: a plan whose quest_decode_workspace_bytes(num_qo_heads, head_dim, num_chunks) exceeds the array is refused
 * with QUEST_ERR_SCRATCH_LIMIT (raise pages_per_chunk, or use quest_decode_run).  ALL calls on a GPU share the one
 * array, which keeps the last call's contents: calls that overlap on the GPU (two streams, or two processes' worth of
 * streams in one context) RACE -- serialise them on one stream.  The first call on each device resolves the array's
 * address (cudaGetSymbolAddress, with that device current); make that first call outside stream capture. */
size_t quest_decode_static_scratch_bytes(void);
int quest_decode_run_static(const void* q, void* o, float* lse,
                            const void* kv_data, const int32_t* indices, uint32_t idx_row_stride,
                            uint32_t num_sel_pages, uint32_t last_page_idx, uint32_t last_page_len,
                            uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                            uint32_t page_size, uint32_t pages_per_chunk, void* stream);

/* ---- fused layer step: estimate -> top-k -> sparse decode ---------------------------------------
 * The three calls QuestAttention.forward makes per layer per token (quest/models/QuestAttention.py:
 * 135-158) behind one entry; kernels are chained with programmatic dependent launch.  scores
 * ([num_qo_heads, len] fp16) and topk_idx ([num_qo_heads, k] int32) are caller-provided
 * intermediates and hold the same values the separate calls would produce.
 * kv_indices: the sequence's page table (device, first `len` entries are the scored pages);
 * last_page_idx: id of the always-attended last page (host value, reference controller.py:86). */
int quest_sparse_decode_step(const void* q, void* o,
                             const void* kv_data, const int32_t* kv_indices, uint32_t len,
                             uint32_t last_page_idx, uint32_t last_page_len,
                             const void* meta_data, const int32_t* meta_indices,
                             void* scores, int32_t* topk_idx, uint32_t k,
                             uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                             uint32_t page_size, uint32_t pages_per_chunk,
                             void* workspace, size_t workspace_bytes, void* stream);

/* ---- one decode step of one layer ---------------------------------------------------------------
 * append_kv + estimate + top-k + sparse decode in one call: the four operator calls of
 * quest/models/QuestAttention.py:100-150 (append_kv, decode_estimate, decode_topk, decode_sparse_attn;
 * quest/utils/__init__.py:137-276).  k_new, v_new: [num_kv_heads, head_dim] of the new token; the page tables
 * and lengths already count it (as for quest_append_decode); kv_last_page_idx = kv_indices[num_kv_pages-1]
 * as a host value (the reference passes it the same way, controller.py:86-87).  k pages are selected among
 * the num_kv_pages-1 full pages (k = 0: attend to the last page only); scores [num_qo_heads, num_kv_pages-1]
 * fp16 and topk_idx [num_qo_heads, k] are outputs.  The append rides on the estimate kernel (3 launches). */
int quest_decode_layer_step(const void* q, const void* k_new, const void* v_new, void* o,
                            void* kv_data, const int32_t* kv_indices, uint32_t num_kv_pages,
                            uint32_t kv_last_page_idx, uint32_t kv_last_page_len,
                            void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                            uint32_t meta_last_page_len,
                            void* scores, int32_t* topk_idx, uint32_t k,
                            uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                            uint32_t page_size, uint32_t pages_per_chunk,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- layer step for a batch of sequences, replayable from a CUDA graph ------------------------------------
 * Same work as quest_decode_layer_step for `batch` independent sequences in three launches (grid.y / grid.z = sequence),
 * with everything that changes from token to token read from DEVICE memory: one 64-byte record per sequence.  The host
 * refreshes the records (one small H2D copy per token, e.g. from a pinned buffer inside the captured graph) and replays
 * the graph; no argument of the call changes while max_kv_pages holds.  This is the reference's per-token
 * InferenceController.begin_forward + four operator calls per layer (controller.py:72-129, QuestAttention.py:100-150)
 * with the per-token index tensors replaced by the records (the reference asserts batch 1,
 * QuestAttention.py:68).
 * q, o: [batch, num_qo_heads, head_dim]; k_new, v_new: [batch, num_kv_heads, head_dim]; scores: per sequence
 * [num_qo_heads, num_kv_pages-1] at scores + b*scores_seq_stride (elements); topk_idx: [batch, num_qo_heads, k];
 * workspace: batch blocks of workspace_seq_bytes (>= quest_decode_workspace_bytes, multiple of 16, zero at start).
 * kv_layer_offset / meta_layer_offset: byte offset of this layer inside every sequence's pools.
 * max_kv_pages: upper bound on num_kv_pages over the batch (sizes grids, the top-k variant, shared memory; <= 16385).
 * A sequence with fewer than k scored pages attends to all of them (the rest of its topk_idx row is -1). */
typedef struct QuestSeqDesc {
  const void* kv_data;          /* the sequence's KV pool, layer 0 */
  const void* meta_data;        /* its metadata pool, layer 0 */
  const int32_t* kv_indices;    /* device page tables */
  const int32_t* meta_indices;
  uint32_t num_kv_pages, kv_last_page_len, num_meta_pages, meta_last_page_len; /* the new token already counted */
  uint32_t rope_pos;            /* position of the new token (quest_rope_decode_batch) */
  uint32_t reserved[3];
} QuestSeqDesc;                 /* 64 bytes */
int quest_decode_layer_step_batch(const void* q, const void* k_new, const void* v_new, void* o,
                                  const QuestSeqDesc* descs, uint32_t batch, size_t kv_layer_offset,
                                  size_t meta_layer_offset, uint32_t max_kv_pages, void* scores,
                                  size_t scores_seq_stride, int32_t* topk_idx, uint32_t k,
                                  uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                                  uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                  size_t workspace_seq_bytes, void* stream);
/* RoPE of one new token per sequence at the position in its record: q [batch, num_qo_heads, head_dim],
 * k [batch, num_kv_heads, head_dim] (apply_rope_in_place with a device-resident past_kv_len, page.cu:212-253). */
int quest_rope_decode_batch(void* q, void* k, const QuestSeqDesc* descs, uint32_t batch, uint32_t num_qo_heads,
                            uint32_t num_kv_heads, uint32_t head_dim, float rope_scale, float rope_theta, void* stream);

/* Device-side record updater: advances every record by one token (page / metadata-page roll-over, last-page lengths, RoPE
 * position), so that a captured step whose first node this is replays over a run of tokens the host has reserved in advance
 * (page ids already in the device tables) with no host work in between.  Replaces the per-token host rebuild of
 * quest/utils/controller.py:72-129 for such runs. */
int quest_seq_desc_advance(QuestSeqDesc* descs, uint32_t batch, uint32_t page_size, void* stream);

/* ---- paged causal prefill ----------------------------------------------------------------------
 * Replaces prefill_with_paged_kv_cache (bsk_ops.h:78-86, batch_prefill.cu:27-117, prefill.cuh:674-1115).
 * q, o: [qo_len, num_qo_heads, head_dim]; the KV sequence is the num_kv_pages pages of kv_indices
 * with kv_last_page_len tokens in the last; query i sits at position kv_len - qo_len + i. */
int quest_prefill(const void* q, void* o, uint32_t qo_len,
                  const void* kv_data, const int32_t* kv_indices, uint32_t num_kv_pages,
                  uint32_t kv_last_page_len, int causal,
                  uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                  uint32_t page_size, void* stream);

/* ---- prefill producer and whole prefill layer step -------------------------
 * quest_rope_append_prefill replaces, in ONE launch, the reference's two launches ahead of prefill attention:
 * apply_rope_in_place over q and k (bsk_ops.h:41-48, quest/ops/csrc/page.cu:212-253, decode_page.cuh:644-728) followed by
 * append_kv_cache_prefill (bsk_ops.h:50-62, page.cu:101-210, decode_page.cuh:471-562).  q [append_len, num_qo_heads, head_dim]
 * is rotated in place at positions kv_len - append_len ..; k [append_len, num_kv_heads, head_dim] is rotated in registers and
 * the rounded result is scattered into the pages and folded into the pages' min/max metadata, v is copied; the pool, the
 * metadata and q come out bit-identical to the two-call sequence.  k itself is left untouched unless flags has
 * QUEST_PF_WRITE_BACK_K (the reference's in-place side effect; nothing on the path reads k again).  Cache arguments as for
 * quest_append_prefill: the page counts and kv_last_page_len already include the new tokens. */
#define QUEST_PF_WRITE_BACK_K 1u
int quest_rope_append_prefill(void* q, void* k, const void* v, uint32_t append_len,
                              void* kv_data, const int32_t* kv_indices, uint32_t num_kv_pages,
                              uint32_t kv_last_page_len,
                              void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                              uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                              float rope_scale, float rope_theta, uint32_t flags, void* stream);
/* The prompt's whole attention layer -- quest/models/QuestAttention.py:85-99 (apply_rope_in_place, append_kv,
 * prefill_forward: three operator calls, three launches and a host sync at batch_prefill.cu:84) -- as one call and two launches
 * chained with programmatic dependent launch: quest_rope_append_prefill, then causal quest_prefill of the rotated q over the
 * pages into o [append_len, num_qo_heads, head_dim]. */
int quest_prefill_layer_step(void* q, void* k, const void* v, void* o, uint32_t append_len,
                             void* kv_data, const int32_t* kv_indices, uint32_t num_kv_pages,
                             uint32_t kv_last_page_len,
                             void* meta_data, const int32_t* meta_indices, uint32_t num_meta_pages,
                             uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                             float rope_scale, float rope_theta, uint32_t flags, void* stream);

/* ---- fp8 (e4m3) page metadata: SURVEY.md section 8 row f-4, a flag-gated variant with NO counterpart in the reference ---------
 * (its metadata is fp16 only, decode_page.cuh:79-352).  A second pool of the same shape [pages, 2, page_size, num_heads,
 * head_dim] with ONE BYTE per bound, indexed by the same metadata page table, holds every per-page max rounded UP and every min
 * rounded DOWN to e4m3, so the estimate over it is still an upper bound of every q.k in the page (and >= the fp16 score) while
 * it streams half the bytes.  |x| > 448 saturates (the bound is lost there; fp16 keys of that size do not occur).  Scores are NOT
 * bit-identical to the reference's -- this is why it is a separate entry point and off by default; the selected set is the
 * reference's top-k of the widened bounds.
 * quest_metadata_to_fp8: (re)builds the e4m3 entries of KV pages [first_kv_page, end_kv_page) from the fp16 pool -- call it
 *   after quest_append_prefill / quest_rope_append_prefill / quest_append_decode for the pages they touched.
 * quest_estimate_fp8: quest_estimate over the e4m3 pool (flags: QUEST_EST_META_STABLE only; the pool is ordered by construction).
 * quest_decode_layer_step_fp8: quest_decode_layer_step scoring from meta8_data; the append keeps both pools current. */
int quest_metadata_to_fp8(const void* meta_data, const int32_t* meta_indices, uint32_t first_kv_page, uint32_t end_kv_page,
                          void* meta8_data, uint32_t num_heads, uint32_t head_dim, uint32_t page_size, void* stream);
int quest_estimate_fp8(const void* q, void* o, uint32_t len, const void* meta8_data, const int32_t* meta_indices,
                       uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                       uint32_t flags, void* stream);
int quest_decode_layer_step_fp8(const void* q, const void* k_new, const void* v_new, void* o, void* kv_data,
                                const int32_t* kv_indices, uint32_t num_kv_pages, uint32_t kv_last_page_idx,
                                uint32_t kv_last_page_len, void* meta_data, void* meta8_data, const int32_t* meta_indices,
                                uint32_t num_meta_pages, uint32_t meta_last_page_len, void* scores, int32_t* topk_idx,
                                uint32_t k, uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim,
                                uint32_t page_size, uint32_t pages_per_chunk, void* workspace,
                                size_t workspace_bytes, void* stream);

/* ---- estimate + selection with scores out (phase 1 of the sequence-sharded step) ------------------------------------
 * No counterpart in the reference (it has no multi-GPU path; SURVEY.md section 8 row e).  quest_estimate over the first `len`
 * pages of the table (optionally carrying append_kv of one new token like quest_decode_layer_step: k_new != NULL), then the k
 * best pages per q head selected from the bin totals the estimate publishes: their scores (fp16) into out_val and their
 * payload ids (kv_indices[page]) into out_idx, rows out_val_stride / out_idx_stride elements apart (0 = k), in ascending page
 * order as quest_topk.  flags: QUEST_EST_META_* as for quest_estimate.  workspace: a quest_decode_workspace_bytes() block,
 * zero when first used.  QUEST_ERR_UNSUPPORTED (nothing launched): shapes / lengths the bin-total path does not take --
 * use quest_append_decode + quest_estimate + quest_topk. */
int quest_estimate_select(const void* q, const void* k_new, const void* v_new, void* kv_data, const int32_t* kv_indices,
                          uint32_t num_kv_pages, uint32_t kv_last_page_len, void* meta_data,
                          const int32_t* meta_indices, uint32_t num_meta_pages, uint32_t meta_last_page_len,
                          uint32_t len, void* scores, void* out_val, uint32_t out_val_stride, int32_t* out_idx,
                          uint32_t out_idx_stride, uint32_t k, uint32_t num_qo_heads, uint32_t num_kv_heads,
                          uint32_t head_dim, uint32_t page_size, uint32_t flags, void* workspace, size_t workspace_bytes,
                          void* stream);

/* ---- combine of partial attention states (sequence-sharded multi-GPU path) ----------------------
 * Merges n partial states (o_i fp32 [n, num_heads, head_dim] normalised, lse_i fp32 [n, num_heads]
 * base-2) into o fp16 [num_heads, head_dim]: flashinfer state.cuh:53-62 / cascade.cuh:279-340.
 * part_stride: floats between consecutive parts in BOTH arrays when each rank's (o, lse) pair travels as one
 * packed record of an all-gather (o_parts = base, lse_parts = base + num_heads*head_dim); 0 = two dense arrays.
 * A part with lse = -inf (a shard that held none of a head's pages) has weight 0. */
int quest_merge_states(const float* o_parts, const float* lse_parts, uint32_t n, uint32_t part_stride,
                       void* o, float* lse, uint32_t num_heads, uint32_t head_dim, void* stream);

/* ---- all-gather of a small message over NVLink peer memory (sequence-sharded path) ---------
 * The reference has no multi-GPU path; this replaces the NCCL all-gather that BASELINE.json's north_star names for
 * the two per-layer exchanges (top-k candidates, (o, lse) records) with one kernel per exchange.  Every rank calls
 * it with the same shapes: src (rows x row_bytes, dense, row_bytes % 16 == 0) is stored into slot `rank` of every
 * peer's staging area, published with a system-scope release of this call's epoch into the peer's flag word, and
 * once all `world` epochs have arrived in this rank's own flag words the staging area is unpacked into
 * dst = [rows][world][row_bytes].  peer_staging[p] / peer_flags[p]: rank p's staging area
 * (quest_peer_staging_bytes) and flag words (world x uint32, zero at start) as mapped into THIS process (host
 * arrays of `world` device pointers; peer-mapped symmetric memory, or plain device memory when world == 1).
 * state: (1 + world) x uint32 of local device memory, zero at start: [0] number of waits that ran into timeout_ms
 * (0 = 2000) -- nonzero means dst is not valid; [1 + p] epoch of the exchange with rank p (advanced by the kernel, so
 * the call can be captured in a CUDA graph).  One CTA per peer: the pairwise exchanges overlap.  world <= 8. */
size_t quest_peer_staging_bytes(uint32_t rows, uint32_t row_bytes, uint32_t world);
int quest_peer_allgather(const void* src, uint32_t rows, uint32_t row_bytes, void* const* peer_staging,
                         void* const* peer_flags, void* state, void* dst, uint32_t world, uint32_t rank,
                         uint32_t timeout_ms, void* stream);

/* ---- device-side failures ------------------------------------------------------------------------------
 * The tcgen05 prefill kernels and the peer exchange wait on hardware barriers / remote flags with a wall-time
 * bound (seconds) instead of hanging the GPU.  A kernel whose wait ran out raises a sticky bit in host-mapped
 * memory; quest_device_error() returns the bits raised so far (a plain host read: no synchronisation, so a
 * failure becomes visible once the failing kernel has run).  While a class's bit is set the entry points of that
 * class (quest_prefill; quest_peer_allgather) return QUEST_ERR_DEVICE.  The reference has no counterpart: its
 * kernels cannot time out (they use no asynchronous barriers), and CUDA launch failures surface as TORCH_CHECK
 * errors (quest/ops/csrc/page.cu:91-93). */
#define QUEST_DEVERR_PREFILL 1u
#define QUEST_DEVERR_PEER 2u
#define QUEST_DEVERR_STEP 4u   /* layer step: a decode CTA never received its page ids from the top-k launch */
#define QUEST_DEVERR_COUNT 3
uint32_t quest_device_error(void);
void quest_device_error_clear(void);

/* ---- explicit test / A-B switches (process-wide; nothing in the library reads the environment) -----------
 * quest_debug_set_option(option, value) returns the previous value; all options default to 0 = product path. */
#define QUEST_DBG_PDL_OFF 0          /* 1: launch without programmatic dependent launch                       */
#define QUEST_DBG_EST_ROWS 1         /* 1: generic estimate kernel on shapes the tile kernel takes             */
#define QUEST_DBG_PREFILL_IMPL 2     /* 0: by shape; 1: SIMT; 2: cp.async + tcgen05; 3: TMA + tcgen05 only     */
#define QUEST_DBG_PREFILL_P_SMEM 3   /* 1: TMA prefill hands P to the PV MMA through shared memory             */
#define QUEST_DBG_TOPK_NO_CLUSTER 4  /* 1: one CTA per top-k row for every length                              */
#define QUEST_DBG_TOPK_GENERIC 5     /* 1: shared-memory-key top-k kernel for every length                     */
#define QUEST_DBG_TOPK_NO_HIST 6     /* 1: no selection from bin totals: the layer step and quest_topk use the register radix kernels */
#define QUEST_DBG_EST_PER_Q 7        /* GQA estimate layout for groups of 2 / 4.  1: one lane row per q head (the round-1 layout); groups
                                        of 4 only: 2 = two lane rows x two q rows in registers, 3 = the whole group in one lane row
                                        (0 picks 2 inside a batch-1 layer step, 3 elsewhere)                        */
#define QUEST_DBG_NO_HANDOFF 8       /* 1: layer step without the page-id hand-over (decode waits for the top-k launch) */
#define QUEST_DBG_NO_ROW_COMPACT 9   /* 1: quest_decode_run_partial walks the sparse list chunk by chunk (no row compaction) */
#define QUEST_DBG_COUNT 10
int quest_debug_set_option(int option, int value);

/* Counters for bench/tests: number of kernels this library has launched since load / last reset. */
uint64_t quest_launch_count(void);
void quest_launch_count_reset(void);

#ifdef __cplusplus
}
#endif
#endif /* QUEST_B200_H_ */
