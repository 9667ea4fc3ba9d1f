// Page-criticality estimate: score[h,p] = sum_d max(q[h,d]*kmax[p,h,d], q[h,d]*kmin[p,h,d]).
// Replaces MaxPossibleSampleWithPagedKVCacheKernel (reference kernels/include/decode/decode_attn.cuh:245-401,
// inner loop :137-168).
//
// HBM-bound stream over the metadata pool, read exactly once.  Two kernels:
//  * estimate_tile_kernel (the product path): a CTA of 16 warps covers 16*32/(D/8) consecutive metadata rows
//    per pass (for 32 kv heads x 128: one page, all heads = 8 KB contiguous), every lane keeps ONE kv head for
//    the whole kernel, walks its CTA's contiguous page range through a 4-deep register ring (8 x 128-bit loads
//    per lane in flight, ~128 KB per SM) and reduces 4 consecutive pages with a transposed butterfly.  About 40
//    instructions per 1 KB of metadata: the arithmetic hides under the stream (tools/kbench_est.cu measures
//    the variants that led here; profiles/r01_kbench_est_{a,b,c}.txt).
//  * estimate_rows_kernel: any shape (kv heads not dividing the tile), same arithmetic, slower addressing.
//
// Arithmetic (bit-identical to the reference's fp32 order, 8 sequential adds per lane then an xor butterfly):
// products of two fp16 values are exact in fp32, and multiplication by q is monotone, so
//   max(q*kmax, q*kmin) == q * (q >= 0 ? max(kmax,kmin) : min(kmax,kmin))
// for all finite inputs with no assumption that kmax >= kmin (signed zeros cannot change the running sum,
// which is never -0).  The select runs on packed halves (one LOP3 per pair against a sign mask), and
// s + q*x is one mixed-precision FMA (fma.rn.f32.f16 -> FHFMA: fp16 multiplicands straight from the packed registers,
// fp32 addend, one rounding) because the product is exact: no fp16 -> fp32 conversion instructions at all.  With QUEST_EST_META_ORDERED the caller promises
// kmax >= kmin elementwise (true for metadata written by append_kv) and the HMNMX2 pair is skipped.
// The transposed butterfly adds the same operand pairs in the same order as the reference's
// shfl_xor(8,4,2,1) tree (own + partner at every level), so the fp32 sum is unchanged.
#include <stdlib.h>

#include "common.cuh"

namespace quest {

// n / d for n < 2^24 and any d >= 1: m = ceil(2^40 / d); exact while n * d < 2^40.
__host__ __device__ inline uint64_t div40_magic(uint32_t d) { return ((1ull << 40) + d - 1) / d; }
__device__ __forceinline__ uint32_t div40(uint32_t n, uint64_t m) { return (uint32_t)(((uint64_t)n * m) >> 40); }

template <int D, int kEstThreads, int kEstDepth, int kG>
__global__ void __launch_bounds__(kEstThreads, 2)
estimate_rows_kernel(const __half* __restrict__ q, __half* __restrict__ o, uint32_t len,
                const __half* __restrict__ meta, const int32_t* __restrict__ meta_indices,
                uint32_t Hq, uint32_t Hkv, uint32_t P, uint64_t magic_mp /* 1/(P*Hkv) */,
                uint64_t magic_h /* 1/Hkv */, uint32_t meta_stable) {
  constexpr int LPR = D / 8;     // lanes per metadata row (reference bdx)
  constexpr int RPL = 32 / LPR;  // rows per warp-wide load
  extern __shared__ __align__(16) uint8_t est_smem[];
  uint32_t* q_s = reinterpret_cast<uint32_t*>(est_smem);                    // [Hq][D/2] packed halves (region sized for fp32)
  uint32_t* m_s = q_s + (size_t)Hq * D;                                     // [Hq][D/2] 0xffff where q >= 0
  int32_t* idx_s = reinterpret_cast<int32_t*>(m_s + (size_t)Hq * D / 2);    // [ceil(len/P)] metadata page table
  const uint32_t G = kG ? kG : Hq / Hkv;  // kG = 0: any group size
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lir = lane % LPR, rsub = lane / LPR;
  const uint32_t total_rows = len * Hkv;
  const uint32_t n_warps = gridDim.x * (kEstThreads / 32);
  const uint32_t step_rows = n_warps * RPL;
  const uint32_t rows_per_mp = P * Hkv;                       // rows of one metadata page's max slab
  const uint32_t slab_bytes = rows_per_mp * D * 2;            // max slab -> min slab
  const uint64_t mp_bytes = 2ull * slab_bytes;                // one metadata page
  const uint8_t* base = reinterpret_cast<const uint8_t*>(meta) + lir * 16;

  // In the layer step the kernel that precedes this one (append) only touches the metadata entry of the
  // last page, which is never scored: with meta_stable the pipeline is primed before the
  // programmatic-dependency wait (q is only read after it).  Otherwise wait first.
  if (!meta_stable) pdl_wait();
  const uint32_t n_meta = (len + P - 1) / P;
  for (uint32_t i = tid; i < n_meta; i += kEstThreads) idx_s[i] = __ldg(meta_indices + i);
  __syncthreads();

  // warps of one CTA are gridDim.x items apart, so neighbouring CTAs stream neighbouring 512-byte runs
  uint32_t r_ld = (warp * gridDim.x + blockIdx.x) * RPL + rsub;  // this lane's row in the next item to load
  uint32_t r = r_ld;                                             // ... and in the next item to consume
  H8 mx[kEstDepth], mn[kEstDepth];
  auto issue = [&](int s) {
    const uint32_t rc = min(r_ld, total_rows - 1);  // past the end: re-read the last row (never stored)
    const uint32_t mp = div40(rc, magic_mp);
    const uint32_t within = rc - mp * rows_per_mp;
    const uint8_t* a = base + (uint64_t)(uint32_t)idx_s[mp] * mp_bytes + within * (uint32_t)(D * 2);
    mx[s] = ld_stream(a);
    mn[s] = ld_stream(a + slab_bytes);
    r_ld += step_rows;
  };
#pragma unroll
  for (int s = 0; s < kEstDepth; ++s) issue(s);

  if (meta_stable) pdl_wait();
  for (uint32_t i = tid; i < Hq * D / 2; i += kEstThreads) {
    // sign-bit test: for q = -0 either choice gives a +-0 product, which cannot change the sum
    const uint32_t bits = reinterpret_cast<const uint32_t*>(q)[i];
    q_s[i] = bits;
    m_s[i] = ((bits & 0x8000u) ? 0u : 0xffffu) | ((bits & 0x80000000u) ? 0u : 0xffff0000u);
  }
  __syncthreads();
#if QUEST_TRIG_EST
  pdl_trigger();
#endif


  while (r - rsub < total_rows) {
#pragma unroll
    for (int s = 0; s < kEstDepth; ++s) {
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const __half2 a = __hmax2(mx[s].h[i], mn[s].h[i]);
        const __half2 b = __hmin2(mx[s].h[i], mn[s].h[i]);
        hi[i] = *reinterpret_cast<const uint32_t*>(&a);
        lo[i] = *reinterpret_cast<const uint32_t*>(&b);
      }
      issue(s);  // refill this stage with the item kEstDepth ahead
      const bool valid = r < total_rows;  // uniform over the LPR lanes of a row
      const uint32_t rc = min(r, total_rows - 1);
      const uint32_t p = div40(rc, magic_h);
      const uint32_t kvh = rc - p * Hkv;
      r += step_rows;
#pragma unroll
      for (uint32_t g = 0; g < G; ++g) {
        const uint32_t h = kvh * G + g;
        const uint4 qq = *reinterpret_cast<const uint4*>(q_s + h * (D / 2) + lir * 4);
        const uint4 msk = *reinterpret_cast<const uint4*>(m_s + h * (D / 2) + lir * 4);
        const uint32_t qh[4] = {qq.x, qq.y, qq.z, qq.w};
        const uint32_t mk[4] = {msk.x, msk.y, msk.z, msk.w};
        float sacc = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t sel = (hi[i] & mk[i]) | (lo[i] & ~mk[i]);
          // reference order (decode_attn.cuh:152-160): 8 sequential fp32 adds of exact products
          sacc = fma_h2_seq(qh[i], sel, sacc);
        }
#pragma unroll
        for (int off = LPR / 2; off > 0; off >>= 1) sacc = __fadd_rn(sacc, __shfl_xor_sync(0xffffffffu, sacc, off));
        if (valid && lir == 0) o[(size_t)h * len + p] = __float2half_rn(sacc);
      }
    }
  }

}

// ---- tile kernel -------------------------------------------------------------------------------------------
constexpr int kTileWarps = 16;



// Transposed butterfly over the LPR lanes of a row group: 4 per-lane partial sums (rows a[0..3]) -> the lane
// holds the complete sum of row (lir / (LPR/4)).  Same operand pairs and order as xor(LPR/2 .. 1) on each row.
template <int LPR>
__device__ __forceinline__ float reduce4_transposed(const float a[4], uint32_t lir) {
  const bool b_hi = lir & (LPR / 2), b_lo = lir & (LPR / 4);
  const float s0 = b_hi ? a[0] : a[2], s1 = b_hi ? a[1] : a[3];
  const float k0 = b_hi ? a[2] : a[0], k1 = b_hi ? a[3] : a[1];
  const float r0 = __fadd_rn(k0, __shfl_xor_sync(0xffffffffu, s0, LPR / 2));
  const float r1 = __fadd_rn(k1, __shfl_xor_sync(0xffffffffu, s1, LPR / 2));
  const float sd = b_lo ? r0 : r1, kp = b_lo ? r1 : r0;
  float v = __fadd_rn(kp, __shfl_xor_sync(0xffffffffu, sd, LPR / 4));
#pragma unroll
  for (int off = LPR / 8; off > 0; off >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}

// kG = 1: every lane row keeps ONE q head for the whole kernel, q and its sign mask in registers.  MHA, and GQA with a
// power-of-two group (g_shift = log2 G) whose q heads divide the row tile: the G lane rows of a group read the same
// kv-head row (the two rows of one warp-wide load coalesce; rows in different warps are served by L2), so the
// arithmetic per lane and the refill-after-use ring are exactly the MHA ones and DRAM still sees every byte once.
// kG = 0: any group size, one lane row per kv head looping over its q heads, q/mask in smem.
// kG = 100 + G (G = 2, 4): one lane row per KV head with the q rows and sign masks of its G q heads in registers (8 registers per
// q head as packed halves): every metadata byte is loaded once per group and scored G times, 12 instructions per (page, q head)
// plus the shared butterfly, against 4x the whole MHA loop when every q head has its own lane row.
// -DQUEST_EST_CLOCKS: every CTA stamps %globaltimer at its phase boundaries (tools/step_timeline.py)
#ifdef QUEST_EST_CLOCKS
__device__ unsigned long long g_est_clk[1024 * 4];
#define E_STAMP(i)                                                                 \
  do {                                                                             \
    if (threadIdx.x == 0) {                                                        \
      unsigned long long t_;                                                       \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                       \
      g_est_clk[(blockIdx.x & 1023) * 4 + (i)] = t_;                               \
    }                                                                              \
  } while (0)
#else
#define E_STAMP(i) do { } while (0)
#endif

// Groups of 4 in registers: 16 registers of q + 16 of sign masks do not fit the 64-register budget of two 16-warp CTAs per SM,
// so that instantiation runs one CTA per SM at up to 128 registers.  Measured alternatives (32q/8kv, 128K: 11.9 us with one lane
// row per q head): this layout 10.2 us; the same with the ring in shared memory (per-lane cp.async slots, 192 KB in flight per
// SM) 12.45 us; a register ring of 8 stages with the masks rebuilt by PRMT (128 KB in flight at one CTA per SM) 11.3 us;
// two CTAs per SM with the masks rebuilt by PRMT spills (profiles/r02_b_experiments.md).
template <int D, int kOrdered, int kG>
__global__ void __launch_bounds__(kTileWarps * 32, kG == 104 ? 1 : 2)
estimate_tile_kernel(const __half* __restrict__ q_, __half* __restrict__ o_, uint32_t len_,
                     const __half* __restrict__ meta_, const int32_t* __restrict__ meta_indices_,
                     uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t p_magic /* ceil(2^32/P), 0 for P=1 */,
                     uint32_t meta_stable, uint32_t g_shift, const AppendArgs ap_, const BatchArgs ba, uint32_t* __restrict__ hist_,
                     uint32_t* __restrict__ epoch) {
  // descriptor mode: blockIdx.y = sequence; lengths, page tables and pools come from its device record
  const __half* __restrict__ q = q_;
  __half* __restrict__ o = o_;
  uint32_t len = len_;
  const __half* __restrict__ meta = meta_;
  const int32_t* __restrict__ meta_indices = meta_indices_;
  AppendArgs ap = ap_;
  // layer step: every score is also counted in its head's bin totals, from which the top-k kernel finds the threshold
  // without a histogram pass of its own (topk.cu, topk_hist_kernel); one workspace block per sequence
  uint32_t* __restrict__ hist = hist_;
  if (hist != nullptr && ba.descs != nullptr) hist += (size_t)blockIdx.y * (ba.ws_stride / 4);
  if (ba.descs != nullptr) {
    const SeqDesc* d = ba.descs + blockIdx.y;
    const uint32_t npg = d->num_kv_pages;
    len = npg - 1;
    meta = reinterpret_cast<const __half*>(d->meta_data + ba.meta_layer_off);
    meta_indices = d->meta_indices;
    q += (size_t)blockIdx.y * Hq * (D);
    o += (size_t)blockIdx.y * ba.score_stride;
    if (ap.k != nullptr) {
      ap.k += (size_t)blockIdx.y * Hkv * (D);
      ap.v += (size_t)blockIdx.y * Hkv * (D);
      ap.kv = reinterpret_cast<__half*>(const_cast<uint8_t*>(d->kv_data) + ba.kv_layer_off);
      ap.kv_indices = d->kv_indices;
      ap.meta = const_cast<__half*>(meta);
      ap.num_kv_pages = npg;
      ap.kv_last_len = d->kv_last_page_len;
      ap.num_meta_pages = d->num_meta_pages;
      ap.meta_last_len = d->meta_last_page_len;
    }
  }
  constexpr int LPR = D / 8, RPL = 32 / LPR, kThreads = kTileWarps * 32, kDepth = 4;
  constexpr uint32_t T_r = kTileWarps * RPL;  // metadata rows per CTA pass; Hkv divides T_r
  extern __shared__ __align__(16) uint8_t est_smem[];
  int32_t* idx_s = reinterpret_cast<int32_t*>(est_smem);                       // [ceil(len/P)]
  const uint32_t n_meta = (len + P - 1) / P;
  uint32_t* q_s = reinterpret_cast<uint32_t*>(est_smem + (((size_t)n_meta * 4 + 15) & ~(size_t)15));  // kG == 0 only: packed halves
  uint32_t* m_s = q_s + (size_t)Hq * D;
  constexpr int kGr = kG >= 100 ? kG - 100 : 0;  // group size held in registers
  const uint32_t G = kGr ? (uint32_t)kGr : kG ? (uint32_t)kG : Hq / Hkv;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lir = lane % LPR, rsub = lane / LPR;
  constexpr int kQR = kGr ? kGr : 1;                    // q rows a lane keeps in registers
  // lane rows a pass walks per page: one per q head (kG = 1), one per kQR q heads (registers; g_shift > 0: 2^g_shift lane
  // rows of the same warp-wide load share a kv-head row), one per kv head (kG = 0)
  const uint32_t HV = (kG == 1 || kGr) ? Hq / kQR : Hkv;
  const uint32_t PPP = T_r / HV;                        // pages per CTA pass
  const uint32_t rt = warp * RPL + rsub;                // this lane's row in the pass
  const uint32_t pg_in = rt / HV, kvh = rt - pg_in * HV;  // kG = 1: kvh is the lane's q head
  // contiguous page range of this CTA, balanced to one page
  // with an append to carry, the last CTA does only that and the scan is shared by the others
  const uint32_t n_scan = ap.k != nullptr ? gridDim.x - 1 : gridDim.x;
  const bool scans = blockIdx.x < n_scan;
  // (a partition skewed towards the CTAs that start first -- they are placed in blockIdx order as the launch before drains --
  // was measured: no gain, every CTA finishes when the DRAM stream does; profiles/r02_b_experiments.md)
  const uint32_t p_begin = scans ? (uint32_t)(((uint64_t)blockIdx.x * len) / n_scan) : 0u;
  const uint32_t p_end = scans ? (uint32_t)(((uint64_t)(blockIdx.x + 1) * len) / n_scan) : 0u;
  // kOrdered == 2: the metadata is the e4m3 pool (one byte per bound, max >= min by construction): 8 bytes per lane and slab,
  // widened to halves when a stage is consumed; everything after that is the fp16 arithmetic
  constexpr bool kF8 = kOrdered == 2;
  constexpr uint32_t kEB = kF8 ? 1 : 2;  // bytes per bound
  const uint32_t row_bytes = Hkv * D * kEB, slab_bytes = P * row_bytes;
  const uint8_t* base = reinterpret_cast<const uint8_t*>(meta) + (size_t)((kG == 1 || kGr) ? kvh >> g_shift : kvh) * (D * kEB) + lir * (8 * kEB);

  // append_kv of the step's new token rides on one extra CTA of this launch.  The entry it updates (the last
  // page's) is not one of the scored ones; the decode kernel two launches later reads the new K/V row.
  const bool carries_append = !scans;
  E_STAMP(0);
  if (len == 0 && !carries_append) return;  // descriptor mode: a sequence with nothing to score yet
  if (!meta_stable || carries_append) pdl_wait();  // see QUEST_EST_META_STABLE
  if (carries_append) {
    const uint32_t row_vecs = Hkv * D / 8;
    for (uint32_t c = tid; c < row_vecs; c += kThreads)
      append_decode_column(ap.k, ap.v, ap.kv, ap.kv_indices, ap.num_kv_pages, ap.kv_last_len, ap.meta, meta_indices,
                           ap.num_meta_pages, ap.meta_last_len, row_vecs, P, c, ap.meta8);
    return;
  }
  for (uint32_t i = tid; i < n_meta; i += kThreads) idx_s[i] = __ldg(meta_indices + i);
  __syncthreads();

  const uint32_t p_clamp = max(p_end, 1u) - 1u;
  uint32_t p_ld = p_begin + pg_in, p = p_ld;
  H8 mx[kDepth], mn[kDepth];
  auto issue = [&](int s, bool ordered = false) {
    // past the end of THIS CTA's range: re-read its own last page (never stored; an L2 hit) -- running on into the
    // neighbour's pages cost 15 % extra DRAM traffic at 128K, where the neighbour's copy has long left L2
    const uint32_t pc = min(p_ld, p_clamp);
    const uint32_t mp = p_magic ? __umulhi(pc, p_magic) : pc;
    const uint32_t e = pc - mp * P;
    const uint8_t* a = base + (uint64_t)(uint32_t)idx_s[mp] * (2u * slab_bytes) + e * row_bytes;
    if (ordered) {  // the first ring: must be in flight before the launch wait below (see ld_stream_ordered)
      mx[s] = kF8 ? ld_stream8_ordered(a) : ld_stream_ordered(a);
      mn[s] = kF8 ? ld_stream8_ordered(a + slab_bytes) : ld_stream_ordered(a + slab_bytes);
    } else {
      mx[s] = kF8 ? ld_stream8(a) : ld_stream(a);
      mn[s] = kF8 ? ld_stream8(a + slab_bytes) : ld_stream(a + slab_bytes);
    }
    p_ld += PPP;
  };
#pragma unroll
  for (int s = 0; s < kDepth; ++s) issue(s, true);
  E_STAMP(1);  // first ring of metadata loads issued

  if (meta_stable && !carries_append) pdl_wait();
  E_STAMP(2);  // the launch before this one has completed
  // layer step with the page-id hand-over (common.cuh, Handoff): one thread opens the step; nothing reads the counter before the
  // top-k kernel, which waits for this whole launch
  if (epoch != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) *epoch = *epoch + 1u;
  uint32_t qh[kQR][4], mk[kQR][4];
  if (kG == 1 || kGr) {
#pragma unroll
    for (int g = 0; g < kQR; ++g) {
      const uint4 raw = ld_ordered16(q + (size_t)(kvh * kQR + g) * D + lir * 8);  // written by the launch before this one
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        qh[g][i] = w[i];
        // sign-bit test: for q = -0 either choice gives a +-0 product, which cannot change the sum
        mk[g][i] = ((w[i] & 0x8000u) ? 0u : 0xffffu) | ((w[i] & 0x80000000u) ? 0u : 0xffff0000u);
      }
    }
  } else {
    for (uint32_t i = tid; i < Hq * D / 2; i += kThreads) {
      const uint32_t bits = reinterpret_cast<const uint32_t*>(q)[i];
      q_s[i] = bits;
      m_s[i] = ((bits & 0x8000u) ? 0u : 0xffffu) | ((bits & 0x80000000u) ? 0u : 0xffff0000u);
    }
    __syncthreads();
  }
#if QUEST_TRIG_EST
  pdl_trigger();
#endif

  while (kGr && p - pg_in < p_end) {  // group in registers: a stage is scored for its G q heads, then refilled at once
    float sa[kQR][kDepth];
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      uint32_t hi[4], lo[4];
      const H8 vmx = kF8 ? e4m3x8_to_h8(*reinterpret_cast<const uint2*>(&mx[s])) : mx[s];
      const H8 vmn = kF8 ? e4m3x8_to_h8(*reinterpret_cast<const uint2*>(&mn[s])) : mn[s];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (kOrdered) {
          hi[i] = *reinterpret_cast<const uint32_t*>(&vmx.h[i]);
          lo[i] = *reinterpret_cast<const uint32_t*>(&vmn.h[i]);
        } else {
          const __half2 a = __hmax2(vmx.h[i], vmn.h[i]);
          const __half2 b = __hmin2(vmx.h[i], vmn.h[i]);
          hi[i] = *reinterpret_cast<const uint32_t*>(&a);
          lo[i] = *reinterpret_cast<const uint32_t*>(&b);
        }
      }
      issue(s);  // refill: the stage's contents are in hi / lo
#pragma unroll
      for (int g = 0; g < kQR; ++g) {
        float sacc = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) sacc = fma_h2_seq(qh[g][i], (hi[i] & mk[g][i]) | (lo[i] & ~mk[g][i]), sacc);
        sa[g][s] = sacc;
      }
    }
    const uint32_t pp = p + (lir / (LPR / 4)) * PPP;
    const bool st = (lir % (LPR / 4)) == 0 && pp < p_end;
#pragma unroll
    for (int g = 0; g < kQR; ++g) {
      const float v = reduce4_transposed<LPR>(sa[g], lir);
      if (st) {
        const uint32_t h = kvh * kQR + g;
        const __half sv = __float2half_rn(v);
        o[(size_t)h * len + pp] = sv;
        if (hist != nullptr) hist_count(hist, h, sv);
      }
    }
    p += kDepth * PPP;
  }
  while (!kGr && p - pg_in < p_end) {  // uniform over the CTA: the pass's first page is still in range
    uint32_t sel[kDepth][4];
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      const H8 vmx = kF8 ? e4m3x8_to_h8(*reinterpret_cast<const uint2*>(&mx[s])) : mx[s];
      const H8 vmn = kF8 ? e4m3x8_to_h8(*reinterpret_cast<const uint2*>(&mn[s])) : mn[s];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint32_t hi, lo;
        if (kOrdered) {
          hi = *reinterpret_cast<const uint32_t*>(&vmx.h[i]);
          lo = *reinterpret_cast<const uint32_t*>(&vmn.h[i]);
        } else {
          const __half2 a = __hmax2(vmx.h[i], vmn.h[i]);
          const __half2 b = __hmin2(vmx.h[i], vmn.h[i]);
          hi = *reinterpret_cast<const uint32_t*>(&a);
          lo = *reinterpret_cast<const uint32_t*>(&b);
        }
        if (kG == 1) {
          sel[s][i] = (hi & mk[0][i]) | (lo & ~mk[0][i]);
        } else {  // keep both; the select depends on the q head
          sel[s][i] = hi;
          mx[s].h[i] = *reinterpret_cast<const __half2*>(&lo);  // parked until the refill below
        }
      }
      if (kG == 1) issue(s);  // refill this stage with the page kDepth passes ahead
    }
    const uint32_t pp = p + (lir / (LPR / 4)) * PPP;  // page whose score this lane holds after the reduce
    const bool st = (lir % (LPR / 4)) == 0 && pp < p_end;
    if (kG == 1) {
      float sa[kDepth];
#pragma unroll
      for (int s = 0; s < kDepth; ++s) {
        float sacc = 0.f;
#pragma unroll
        // reference order (decode_attn.cuh:152-160): 8 sequential fp32 adds of exact products
        for (int i = 0; i < 4; ++i) sacc = fma_h2_seq(qh[0][i], sel[s][i], sacc);
        sa[s] = sacc;
      }
      const float v = reduce4_transposed<LPR>(sa, lir);
      if (st) {
        const __half sv = __float2half_rn(v);
        o[(size_t)kvh * len + pp] = sv;
        if (hist != nullptr) hist_count(hist, kvh, sv);
      }
    } else {
      for (uint32_t g = 0; g < G; ++g) {
        const uint32_t h = kvh * G + g;
        const uint4 qq = *reinterpret_cast<const uint4*>(q_s + h * (D / 2) + lir * 4);
        const uint4 msk = *reinterpret_cast<const uint4*>(m_s + h * (D / 2) + lir * 4);
        const uint32_t qg[4] = {qq.x, qq.y, qq.z, qq.w};
        const uint32_t mg[4] = {msk.x, msk.y, msk.z, msk.w};
        float sa[kDepth];
#pragma unroll
        for (int s = 0; s < kDepth; ++s) {
          float sacc = 0.f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t lo = *reinterpret_cast<const uint32_t*>(&mx[s].h[i]);
            sacc = fma_h2_seq(qg[i], (sel[s][i] & mg[i]) | (lo & ~mg[i]), sacc);
          }
          sa[s] = sacc;
        }
        const float v = reduce4_transposed<LPR>(sa, lir);
        if (st) {
          const __half sv = __float2half_rn(v);
          o[(size_t)h * len + pp] = sv;
          if (hist != nullptr) hist_count(hist, h, sv);
        }
      }
#pragma unroll
      for (int s = 0; s < kDepth; ++s) issue(s);
    }
    p += kDepth * PPP;
  }
  E_STAMP(3);  // this CTA's pages are scored
}

template <int D>
static int launch_estimate_tile(const void* q, void* o, uint32_t len, const void* meta, const int32_t* meta_indices,
                                uint32_t Hq, uint32_t Hkv, uint32_t P, uint32_t flags, const AppendArgs& ap,
                                cudaStream_t stream, const BatchArgs& ba = BatchArgs{}, uint32_t batch = 1,
                                uint32_t* hist = nullptr, uint32_t* epoch = nullptr) {
  // descriptor mode (ba.descs): `len` is an upper bound over the batch (sizes the grid and the page-table staging)
  constexpr uint32_t T_r = kTileWarps * (32 / (D / 8));
  static int num_sms = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (num_sms <= 0) num_sms = 148;
  }
  const uint32_t G = Hq / Hkv;
  const bool in_regs = (G == 2 || G == 4) && debug_option(QUEST_DBG_EST_PER_Q) != 1;  // the group's q rows in registers
  const bool per_q = !in_regs && (G & (G - 1)) == 0 && T_r % Hq == 0;  // one q head per lane row (MHA, other power-of-two groups)
  uint32_t g_shift = 0;
  while ((1u << g_shift) < G) ++g_shift;
  // A group of 4 either sits in one lane row (4 q rows + sign masks = 32 registers: one CTA per SM) or is split over the two lane
  // rows of one warp-wide load with two q rows each (64 registers, two CTAs per SM, 1.4x the instructions).  Alone the first is
  // faster (10.3 vs 10.9 us, 32q/8kv at 128K); inside a layer step the second is (31.25 vs 31.7 us at 128K, 16.4 vs 16.9 us at 32K:
  // twice the CTAs take the slots the launch before frees).  profiles/r02_b_experiments.md section 12.
  const int layout = debug_option(QUEST_DBG_EST_PER_Q);
  const bool in_step = batch == 1 && (ap.k != nullptr || hist != nullptr);
  const bool split4 = in_regs && G == 4 && T_r % (Hq / 2) == 0 && (layout == 2 || (layout == 0 && in_step && !(flags & kEstMetaFp8)));  // e4m3 pool: the split is slower in the step too
  const uint32_t in_reg_rows = split4 ? 2u : G;   // q rows per lane row
  const uint32_t PPP = T_r / (per_q ? Hq : in_regs ? Hq / in_reg_rows : Hkv);
  const uint32_t occ = (in_regs && G == 4 && !split4) ? 1u : 2u;  // resident CTAs per SM
  uint32_t grid = occ * num_sms / batch;  // CTAs per sequence
  if (grid < 8) grid = 8;
  grid -= (ap.k != nullptr ? 1u : 0u);  // scanning CTAs
  const uint32_t need = (len + PPP - 1) / PPP;  // at least one pass of pages per CTA
  if (grid > need) grid = need;
  if (ap.k != nullptr) ++grid;  // + the CTA that carries the append
  const uint32_t n_meta = (len + P - 1) / P;
  size_t smem = ((size_t)n_meta * 4 + 15) & ~(size_t)15;
  if (!per_q && !in_regs) smem += (size_t)Hq * D * 6;
  if (smem > 100 * 1024) return QUEST_ERR_UNSUPPORTED;
  const bool f8 = flags & kEstMetaFp8;
  if (f8 && ba.descs != nullptr) return QUEST_ERR_UNSUPPORTED;  // the descriptor records name the fp16 pool only
  const bool ord = flags & QUEST_EST_META_ORDERED;
#define QUEST_EST_KERN(KG) (f8 ? estimate_tile_kernel<D, 2, KG> : ord ? estimate_tile_kernel<D, 1, KG> : estimate_tile_kernel<D, 0, KG>)
  auto kern = per_q ? QUEST_EST_KERN(1) : !in_regs ? QUEST_EST_KERN(0) : (G == 2 || split4) ? QUEST_EST_KERN(102) : QUEST_EST_KERN(104);
#undef QUEST_EST_KERN
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid, batch);
  cfg.blockDim = dim3(kTileWarps * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = quest::pdl_enabled();
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const uint32_t p_magic = P <= 1 ? 0u : (uint32_t)((0x100000000ull + P - 1) / P);
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, (const __half*)q, (__half*)o, len, (const __half*)meta, meta_indices,
                                     Hq, Hkv, P, p_magic, flags & QUEST_EST_META_STABLE, per_q ? g_shift : split4 ? 1u : 0u, ap, ba, hist, epoch);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
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
  const size_t smem = (size_t)Hq * D * sizeof(float) + (size_t)Hq * D / 2 * sizeof(uint32_t) + (size_t)n_meta * 4;
  if (smem > 100 * 1024 || (uint64_t)len * Hkv >= (1u << 24)) return QUEST_ERR_UNSUPPORTED;
  const uint32_t G = Hq / Hkv;
  auto kern = G == 1 ? estimate_rows_kernel<D, kEstThreads, kEstDepth, 1> : G == 4 ? estimate_rows_kernel<D, kEstThreads, kEstDepth, 4>
                                                                       : estimate_rows_kernel<D, kEstThreads, kEstDepth, 0>;
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
                                     meta_indices, Hq, Hkv, P, div40_magic(P * Hkv), div40_magic(Hkv),
                                     flags & QUEST_EST_META_STABLE);
  if (e != cudaSuccess) return (int)e;
  QUEST_RETURN_LAUNCH();
}

}  // namespace quest

namespace quest {
// estimate, optionally carrying the step's append_kv (ap.k != nullptr) on the tile kernel's first CTA
int estimate_dispatch(const void* q, void* o, uint32_t len, const void* meta_data, const int32_t* meta_indices,
                      uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                      uint32_t flags, const AppendArgs& ap, cudaStream_t s, const BatchArgs& ba, uint32_t batch,
                      uint32_t* hist, uint32_t* epoch) {
  if ((size_t)num_qo_heads * head_dim * 6 > 160 * 1024) return QUEST_ERR_UNSUPPORTED;
  {  // product path: the tile kernel, whenever the kv heads divide its row tile
    const uint32_t T_r = kTileWarps * (32 / (head_dim / 8));
    const bool force_rows = debug_option(QUEST_DBG_EST_ROWS) != 0;  // test hook: the generic kernel on tile-kernel shapes
    if (!force_rows && len && T_r % num_kv_heads == 0 && (head_dim == 64 || head_dim == 128 || head_dim == 256)) {
      switch (head_dim) {
        case 64: return launch_estimate_tile<64>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, ap, s, ba, batch, hist, epoch);
        case 128: return launch_estimate_tile<128>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, ap, s, ba, batch, hist, epoch);
        default: return launch_estimate_tile<256>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, ap, s, ba, batch, hist, epoch);
      }
    }
  }
  if (ba.descs != nullptr || hist != nullptr || epoch != nullptr || (flags & kEstMetaFp8)) return QUEST_ERR_UNSUPPORTED;  // tile kernel only
  if (ap.k != nullptr) {  // generic path: append as its own launch
    const int rc = quest_append_decode(ap.k, ap.v, ap.kv, ap.kv_indices, ap.num_kv_pages, ap.kv_last_len, ap.meta,
                                       meta_indices, ap.num_meta_pages, ap.meta_last_len, num_kv_heads, head_dim,
                                       page_size, s);
    if (rc) return rc;
    flags &= ~QUEST_EST_META_STABLE;  // the kernel before the estimate now writes the metadata pool
  }
  if (len == 0) return QUEST_OK;
  switch (head_dim) {
    case 64: return launch_estimate<64, 448, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
    case 128: return launch_estimate<128, 448, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
    case 256: return launch_estimate<256, 448, 4>(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, page_size, flags, s);
    default: return QUEST_ERR_UNSUPPORTED;
  }
}
}  // namespace quest

#ifdef QUEST_EST_CLOCKS
extern "C" int quest_debug_estimate_clocks(unsigned long long* host_out, uint32_t n_ctas) {
  return (int)cudaMemcpyFromSymbol(host_out, quest::g_est_clk, (size_t)n_ctas * 4 * 8);
}
#endif

extern "C" int quest_estimate(const void* q, void* o, uint32_t len, const void* meta_data,
                              const int32_t* meta_indices, uint32_t num_qo_heads, uint32_t num_kv_heads,
                              uint32_t head_dim, uint32_t page_size, uint32_t flags, void* stream) {
  if (!q || !o || !meta_data || !meta_indices || !num_qo_heads || !num_kv_heads || !page_size)
    return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (len == 0) return QUEST_OK;
  quest::AppendArgs none = {};
  return quest::estimate_dispatch(q, o, len, meta_data, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size,
                                  flags, none, (cudaStream_t)stream, quest::BatchArgs{}, 1, nullptr, nullptr);
}

// fp8 metadata (no counterpart in the reference, whose metadata is fp16 only): the same estimate over
// the e4m3 pool.  The scores are those of the fp16 arithmetic applied to the widened bounds (the tests restate this on the
// dequantized pool, bit for bit); they are >= the fp16-metadata scores because the pool is rounded outward.
extern "C" int quest_estimate_fp8(const void* q, void* o, uint32_t len, const void* meta8_data, const int32_t* meta_indices,
                                  uint32_t num_qo_heads, uint32_t num_kv_heads, uint32_t head_dim, uint32_t page_size,
                                  uint32_t flags, void* stream) {
  if (!q || !o || !meta8_data || !meta_indices || !num_qo_heads || !num_kv_heads || !page_size) return QUEST_ERR_BAD_ARG;
  if (num_qo_heads % num_kv_heads) return QUEST_ERR_BAD_ARG;
  if (len == 0) return QUEST_OK;
  quest::AppendArgs none = {};
  return quest::estimate_dispatch(q, o, len, meta8_data, meta_indices, num_qo_heads, num_kv_heads, head_dim, page_size,
                                  (flags & QUEST_EST_META_STABLE) | quest::kEstMetaFp8, none, (cudaStream_t)stream,
                                  quest::BatchArgs{}, 1, nullptr, nullptr);
}
