// Shared device helpers for the quest_b200 kernels (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/quest_b200.h"

// Where each kernel releases its programmatic dependents (0 = only at exit, 1 = after its prologue).  Measured on
// B200 (profiles/r01_pdl_trigger_ab.txt, r01_pdl_trigger_ab2.txt; us per layer step): releasing from the estimate
// kernel parks the top-k CTAs on SM slots the scan needs (21.7 -> 26.9); from the top-k kernel it changes nothing;
// from the decode kernel it lets the next launch on the stream -- the next layer's estimate, whose first ring of
// metadata loads does not depend on anything earlier -- start under the decode's merge tail (21.7 -> 20.0).
// Round 2, with the page-id hand-over (the top-k kernel then releases the decode launch only from behind its wait, topk.cu):
// QUEST_TRIG_EST=1 parks the 32 top-k CTAs as the first estimate CTAs retire, 18.37 -> 18.56 us (profiles/r02_b_experiments.md, 16).
#ifndef QUEST_TRIG_EST
#define QUEST_TRIG_EST 0
#endif
#ifndef QUEST_TRIG_TOPK
#define QUEST_TRIG_TOPK 0
#endif
#ifndef QUEST_TRIG_DEC
#define QUEST_TRIG_DEC 1
#endif

namespace quest {

extern unsigned long long g_launch_count;  // defined in capi.cu
int debug_option(int option);  // quest_debug_set_option (capi.cu): explicit test / A-B switches, never the environment
inline int pdl_enabled() { return debug_option(QUEST_DBG_PDL_OFF) == 0; }  // programmatic dependent launch attribute

// Sticky, host-visible error words in mapped pinned memory (capi.cu).  A kernel that gives up on a bounded wait stores
// into its class's word; every later quest_* entry point that depends on that class returns QUEST_ERR_DEVICE until
// quest_device_error_clear(), so corrupt data is never handed back silently.
unsigned int* device_error_word();  // device-visible pointer to QUEST_DEVERR_COUNT words (UVA: valid on every GPU)
unsigned int device_error_peek();   // OR of the QUEST_DEVERR_* bits raised so far; a host read, no synchronisation
#ifdef __CUDACC__
__device__ __forceinline__ void device_error_raise(unsigned int* words, unsigned int bit) {
  if (words == nullptr) return;
  *reinterpret_cast<volatile unsigned int*>(words + (31 - __clz(bit))) = 1u;
  __threadfence_system();
}
#endif

// 16-byte vector of 8 halves: the reference's vec_t<half, 8> (vec_size = max(16/sizeof(half), D/32)).
struct alignas(16) H8 {
  __half2 h[4];
};

// Streaming 128-bit load that bypasses L1 allocation (data is touched once).
__device__ __forceinline__ H8 ld_stream(const void* p) {
  H8 r;
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3])
               : "l"(p));
  return r;
}
// The same load without .nc: ptxas moves non-coherent (read-only) loads freely, also DOWN across griddepcontrol.wait and CTA
// barriers; a kernel that wants loads in flight BEFORE its programmatic-dependency wait issues them with this form.
__device__ __forceinline__ H8 ld_stream_ordered(const void* p) {
  H8 r;
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3])
               : "l"(p)
               : "memory");
  return r;
}
// 8-byte forms (fp8 metadata rows: 8 dims per lane): the bytes land in the first two words of the H8
__device__ __forceinline__ H8 ld_stream8(const void* p) {
  H8 r;
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(u[0]), "=r"(u[1]) : "l"(p));
  return r;
}
__device__ __forceinline__ H8 ld_stream8_ordered(const void* p) {
  H8 r;
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(u[0]), "=r"(u[1]) : "l"(p) : "memory");
  return r;
}
// Predicated form: r keeps its previous contents when !pred (no memory access is made).
__device__ __forceinline__ void ld_stream_if(H8& r, const void* p, bool pred) {
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.u32 p, %5, 0;\n@p ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n}"
      : "+r"(u[0]), "+r"(u[1]), "+r"(u[2]), "+r"(u[3])
      : "l"(p), "r"((uint32_t)pred));
}
__device__ __forceinline__ H8 ld_plain(const void* p) { return *reinterpret_cast<const H8*>(p); }
__device__ __forceinline__ void st_plain(void* p, const H8& v) { *reinterpret_cast<H8*>(p) = v; }

__device__ __forceinline__ void h8_to_float(const H8& v, float* f) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __half22float2(v.h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2(float x) {
  float y;
  asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// c + a.lo*b.lo, then + a.hi*b.hi, each a single-rounding fp32 FMA whose multiplicands are fp16 (PTX fma.rn.f32.f16, sm_100+:
// one FHFMA each, operands taken from the .H0 / .H1 halves of the packed registers, no conversion instructions).  The product of
// two halves is exact in fp32, so this is bit-identical to __fmaf_rn(float(a), float(b), c) in the same order.
__device__ __forceinline__ float fma_h2_seq(uint32_t a, uint32_t b, float c) {
#ifdef QUEST_NO_FHFMA  // A/B build: convert, then two FFMAs (same bits)
  const float2 fa = __half22float2(*reinterpret_cast<const __half2*>(&a));
  const float2 fb = __half22float2(*reinterpret_cast<const __half2*>(&b));
  return __fmaf_rn(fa.y, fb.y, __fmaf_rn(fa.x, fb.x, c));
#else
  float d;
  asm("{\n.reg .b16 al, ah, bl, bh;\n.reg .f32 t;\nmov.b32 {al, ah}, %1;\nmov.b32 {bl, bh}, %2;\n"
      "fma.rn.f32.f16 t, al, bl, %3;\nfma.rn.f32.f16 %0, ah, bh, t;\n}"
      : "=f"(d)
      : "r"(a), "r"(b), "f"(c));
  return d;
#endif
}

// Loads of data the launch immediately before this one produced (q, page ids): plain ld.global, never .nc -- under programmatic
// dependent launch the kernel's lifetime starts before griddepcontrol.wait, the data is not read-only over that lifetime, and
// ptxas does hoist .nc loads above the wait (it did for the estimate kernel's q in the round-1 build).
__device__ __forceinline__ uint4 ld_ordered16(const void* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ H8 ld_ordered_h8(const void* p) {
  const uint4 v = ld_ordered16(p);
  return *reinterpret_cast<const H8*>(&v);
}
__device__ __forceinline__ uint32_t ld_ordered32(const void* p) {
  uint32_t v;
  asm volatile("ld.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Programmatic dependent launch: wait for the producer grid / release the consumer grid.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// NHD pool addressing (reference decode_page.cuh:196-239): element offset of (page, slab, entry, head, 0).
__device__ __forceinline__ size_t nhd_off(size_t page, uint32_t slab, uint32_t entry, uint32_t head,
                                          uint32_t P, uint32_t H, uint32_t D) {
  return (((page * 2 + slab) * P + entry) * H + head) * (size_t)D;
}

__device__ __forceinline__ void h8_max(H8& a, const H8& b) {
#pragma unroll
  for (int i = 0; i < 4; ++i) a.h[i] = __hmax2(a.h[i], b.h[i]);
}
__device__ __forceinline__ void h8_min(H8& a, const H8& b) {
#pragma unroll
  for (int i = 0; i < 4; ++i) a.h[i] = __hmin2(a.h[i], b.h[i]);
}
__device__ __forceinline__ H8 h8_fill(float x) {
  H8 r;
  const __half2 v = __float2half2_rn(x);
#pragma unroll
  for (int i = 0; i < 4; ++i) r.h[i] = v;
  return r;
}

// Llama rotate-half RoPE of 8 consecutive dims (reference decode_page.cuh:644-728, flashinfer rope.cuh:64-83): a = the dims
// themselves, b = their partners D/2 away; out = a*cos + (lo_half ? -b : b)*sin at angle pos*freq[i], with the reference's fast
// intrinsics.  Written with explicit fp32 intrinsics so that every kernel that rotates (rope_norm.cu, page.cu) rounds alike.
__device__ __forceinline__ H8 rope_rotate8(const float a[8], const float b[8], bool lo_half, float pos, const float freq[8]) {
  H8 out;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float r2[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int e = 2 * i + j;
      float sn, cs;
      __sincosf(__fmul_rn(pos, freq[e]), &sn, &cs);
      r2[j] = __fmaf_rn(a[e], cs, __fmul_rn(lo_half ? -b[e] : b[e], sn));
    }
    out.h[i] = __floats2half2_rn(r2[0], r2[1]);
  }
  return out;
}
// freq_i = (1/scale) * (1/theta)^(2 (i mod D/2) / D) for the 8 dims starting at d0
template <int D>
__device__ __forceinline__ void rope_freq8(uint32_t d0, float rcp_scale, float rcp_theta, float freq[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) freq[i] = rcp_scale * __powf(rcp_theta, float(2 * ((d0 + i) % (D / 2))) / float(D));
}

// ---- fp8 (e4m3) copy of the page metadata, SURVEY.md section 8 row f-4 (flag-gated; the reference has fp16 only) -----------
// A second pool of the same shape holds every bound as one e4m3 byte, rounded OUTWARD: max entries up, min entries down, so
// sum_d max(q*max8, q*min8) is still an upper bound of every q.k in the page and the estimate streams half the bytes.
// |x| > 448 (the largest e4m3) saturates to +-448 -- the one case where the bound is lost; fp16 keys of that size do not occur.
__device__ __forceinline__ uint32_t f16x2_to_e4m3x2_rn(uint32_t h2) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f16x2 %0, %1;" : "=h"(r) : "r"(h2));
  return r;
}
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t b2) {
  uint32_t r;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"((uint16_t)b2));
  return r;
}
// smallest e4m3 >= x for both halves of h2 (two bytes, low half in the low byte)
__device__ __forceinline__ uint32_t e4m3x2_round_up(uint32_t h2) {
  const uint32_t b = f16x2_to_e4m3x2_rn(h2);
  const uint32_t y = e4m3x2_to_f16x2(b);
  const __half2 lt = __hlt2(*reinterpret_cast<const __half2*>(&y), *reinterpret_cast<const __half2*>(&h2));  // 1.0 where y < x
  const uint32_t m = *reinterpret_cast<const uint32_t*>(&lt);
  uint32_t lo = b & 0xffu, hi = (b >> 8) & 0xffu;
  // one step towards +inf: magnitude up for positives (not past 448 = 0x7e), magnitude down for negatives
  if (m & 0xffffu) lo = (lo & 0x80u) ? lo - 1u : (lo < 0x7eu ? lo + 1u : lo);
  if (m >> 16) hi = (hi & 0x80u) ? hi - 1u : (hi < 0x7eu ? hi + 1u : hi);
  return lo | (hi << 8);
}
__device__ __forceinline__ uint32_t e4m3x2_round_down(uint32_t h2) { return e4m3x2_round_up(h2 ^ 0x80008000u) ^ 0x8080u; }
__device__ __forceinline__ uint2 h8_to_e4m3_up(const H8& v) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
  return make_uint2(e4m3x2_round_up(u[0]) | (e4m3x2_round_up(u[1]) << 16), e4m3x2_round_up(u[2]) | (e4m3x2_round_up(u[3]) << 16));
}
__device__ __forceinline__ uint2 h8_to_e4m3_down(const H8& v) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
  return make_uint2(e4m3x2_round_down(u[0]) | (e4m3x2_round_down(u[1]) << 16),
                    e4m3x2_round_down(u[2]) | (e4m3x2_round_down(u[3]) << 16));
}
__device__ __forceinline__ H8 e4m3x8_to_h8(uint2 b) {
  H8 r;
  uint32_t* u = reinterpret_cast<uint32_t*>(&r);
  u[0] = e4m3x2_to_f16x2(b.x & 0xffffu), u[1] = e4m3x2_to_f16x2(b.x >> 16);
  u[2] = e4m3x2_to_f16x2(b.y & 0xffffu), u[3] = e4m3x2_to_f16x2(b.y >> 16);
  return r;
}

constexpr float kMaxHalf = 65504.f;  // CUDART_MAX_NORMAL_FP16 (decode_page.cuh:430-431)

// append_kv of one token for one 16-byte column c of the [H*D] row: K/V rows into the last KV page, running
// max/min into the last metadata entry (reference decode_page.cuh:398-449; -+65504 start for a page's first token).
__device__ __forceinline__ void append_decode_column(const __half* __restrict__ k, const __half* __restrict__ v,
                                                     __half* __restrict__ kv, const int32_t* __restrict__ kv_indices,
                                                     uint32_t num_kv_pages, uint32_t kv_last_len, __half* __restrict__ meta,
                                                     const int32_t* __restrict__ meta_indices, uint32_t num_meta_pages,
                                                     uint32_t meta_last_len, uint32_t row_vecs /* H*D/8 */, uint32_t P,
                                                     uint32_t c, uint8_t* __restrict__ meta8 = nullptr) {
  const size_t row = (size_t)row_vecs * 8;  // H*D elements
  const size_t page = (size_t)__ldg(kv_indices + num_kv_pages - 1);
  const uint32_t entry = kv_last_len - 1;
  const size_t mpage = (size_t)__ldg(meta_indices + num_meta_pages - 1);
  const uint32_t mentry = meta_last_len - 1;
  __half* kdst = kv + ((page * 2) * P + entry) * row + c * 8;
  __half* vdst = kdst + (size_t)P * row;
  __half* mxp = meta + ((mpage * 2) * P + mentry) * row + c * 8;
  __half* mnp = mxp + (size_t)P * row;
  const H8 kk = ld_plain(k + c * 8);
  const H8 vv = ld_plain(v + c * 8);
  H8 mx, mn;
  if (entry > 0) {
    mx = ld_plain(mxp);
    mn = ld_plain(mnp);
  } else {
    mx = h8_fill(-kMaxHalf);
    mn = h8_fill(kMaxHalf);
  }
  h8_max(mx, kk);
  h8_min(mn, kk);
  st_plain(kdst, kk);
  st_plain(vdst, vv);
  st_plain(mxp, mx);
  st_plain(mnp, mn);
  if (meta8 != nullptr) {  // fp8 copy of the entry just updated: current by the time the page stops being the last (unscored) one
    uint8_t* m8 = meta8 + ((mpage * 2) * P + mentry) * row + c * 8;
    *reinterpret_cast<uint2*>(m8) = h8_to_e4m3_up(mx);
    *reinterpret_cast<uint2*>(m8 + (size_t)P * row) = h8_to_e4m3_down(mn);
  }
}

// Optional append_kv work carried by another kernel's first CTA (estimate.cu); k == nullptr: none.
struct AppendArgs {
  const __half* k;
  const __half* v;
  __half* kv;
  const int32_t* kv_indices;
  __half* meta;
  uint32_t num_kv_pages, kv_last_len, num_meta_pages, meta_last_len;
  uint8_t* meta8;  // fp8 metadata pool to keep in step with the entry the append updates (nullptr: none)
};

// Descriptor-driven launches (quest_decode_layer_step_batch): everything that changes from token to token lives in one
// 64-byte record per sequence in DEVICE memory (== QuestSeqDesc of include/quest_b200.h), so that a captured CUDA graph of
// the step can be replayed for the next token after the records were refreshed.  descs == nullptr: scalar arguments.
struct SeqDesc {
  const uint8_t* kv_data;        // the sequence's KV pool (all layers), bytes
  const uint8_t* meta_data;      // its metadata pool
  const int32_t* kv_indices;     // device page tables
  const int32_t* meta_indices;
  uint32_t num_kv_pages, kv_last_page_len, num_meta_pages, meta_last_page_len;  // the new token already counted
  uint32_t rope_pos, reserved[3];
};
static_assert(sizeof(SeqDesc) == 64, "SeqDesc is part of the C ABI");
struct BatchArgs {
  const SeqDesc* descs;
  size_t kv_layer_off, meta_layer_off;  // byte offset of the layer inside every sequence's pools
  size_t score_stride;                  // elements between the sequences' score blocks
  size_t ws_stride;                     // bytes between the sequences' decode workspaces
};

constexpr float kLog2e = 1.44269504088896340736f;
constexpr uint32_t kEstMetaFp8 = 0x100u;  // internal estimate flag (beside QUEST_EST_META_*): meta_data is the e4m3 pool

// fp16 bits -> order-preserving 16-bit key (raft select_radix.cuh:82-90 twiddle_in)
__device__ __forceinline__ uint16_t score_key(uint16_t h) { return (h & 0x8000u) ? (uint16_t)~h : (uint16_t)(h | 0x8000u); }
// estimate epilogue of the layer step: count one score of head h in the head's bin totals (fire-and-forget reduction at L2)
__device__ __forceinline__ void hist_count(uint32_t* __restrict__ hist, uint32_t h, __half score) {
  const uint32_t bin = (uint32_t)score_key(__half_as_ushort(score)) >> 5;
  asm volatile("red.global.add.u32 [%0], %1;" ::"l"(hist + (size_t)h * 1024 + (bin >> 1)), "r"(1u << ((bin & 1u) * 16)) : "memory");
}

// Decode workspace layout (caller-owned bytes, zero-filled once when allocated):
//   [0, 4096)                   reserved (zero)
//   [4096, 4096 + 4*Hq)         one merge ticket per q head (self-resetting)
//   ws_hist_off(Hq)             layer step: per q head, how many of its scores fall in each of 2048 key bins (16-bit counts,
//                               4 KB per head); added to by the estimate kernel, read and zeroed by the top-k kernel
//   ws_md_off(Hq)               fp32 (m, d) pairs [Hq][nchunks]
//   ws_o_off(Hq, nchunks)       fp32 partial outputs [Hq][nchunks][D]
// The zero-state words sit at fixed offsets so that launches with different chunk counts can share a workspace.
constexpr size_t kWsSyncBytes = 4096;
constexpr size_t kWsEpochOff = 64;  // layer step: the step counter of the page-id hand-over (inside the reserved, zero-initialised block)

// Page-id hand-over of the layer step (top-k kernel -> decode kernel without a grid dependency).  The decode launch is released
// as soon as the top-k CTAs are running; each (head, chunk) decode CTA polls ITS OWN slot of the split-KV partial-output area,
// where the top-k CTA of that head deposits the chunk's page ids as 8-byte entries (tag : id).  The tag is a NaN bit pattern
// that carries the step counter, so an entry can be told from a stale partial output (never NaN) or an older step without any
// fence: an aligned 8-byte store is single-copy atomic.  Saves the grid completion + flush + launch latency between the two
// kernels and one dependent L2 round trip.  epoch == nullptr: off (the decode kernel waits for the launch before it).
struct Handoff {
  uint32_t* epoch;      // workspace + kWsEpochOff: bumped once per step by the estimate launch
  uint8_t* slots;       // workspace + ws_o_off(Hq, nchunks)
  uint32_t nchunks, ppc, slot_bytes;  // slot of (head, chunk) at slots + (head * nchunks + chunk) * slot_bytes; ppc * 8 <= slot_bytes
};
// Optional outputs of the bin-total top-k beyond the dense [heads, k] id list: the selected scores themselves (fp16 bits) and row
// strides in elements (0 = k), so that a caller can have both written straight into a slice of a larger array (the sequence-
// sharded step's candidate message and payload row).
struct HistOut {
  uint16_t* val;
  uint32_t val_stride, idx_stride;
};
__device__ __forceinline__ uint32_t handoff_tag(uint32_t epoch) { return 0xfff00000u | (epoch & 0xfffffu); }
__device__ __forceinline__ void st_relaxed_u64(void* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const void* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__host__ __device__ inline size_t ws_ticket_off() { return kWsSyncBytes; }
constexpr uint32_t kHistBinShift = 5;                               // bin = 16-bit key >> 5
constexpr uint32_t kHistBins = 1u << (16 - kHistBinShift);          // 2048
constexpr size_t kHistBytesPerHead = kHistBins * 2;                 // 16-bit counts: rows of up to 65535 scores
__host__ __device__ inline size_t ws_hist_off(uint32_t Hq) { return (kWsSyncBytes + (size_t)Hq * 4 + 255) & ~(size_t)255; }
__host__ __device__ inline size_t ws_md_off(uint32_t Hq) { return ws_hist_off(Hq) + (size_t)Hq * kHistBytesPerHead; }
__host__ __device__ inline size_t ws_o_off(uint32_t Hq, uint32_t nchunks) {
  return (ws_md_off(Hq) + (size_t)Hq * nchunks * 2 * sizeof(float) + 255) & ~(size_t)255;
}
__host__ __device__ inline size_t ws_bytes(uint32_t Hq, uint32_t D, uint32_t nchunks) {
  return ws_o_off(Hq, nchunks) + (size_t)Hq * nchunks * D * sizeof(float);
}

}  // namespace quest

#define QUEST_RETURN_LAUNCH()                         \
  do {                                                \
    ++quest::g_launch_count;                          \
    cudaError_t e__ = cudaGetLastError();             \
    return e__ == cudaSuccess ? QUEST_OK : (int)e__;  \
  } while (0)
