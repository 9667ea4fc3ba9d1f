// tcgen05 / TMEM / mbarrier primitives for the sm_100a tensor-core kernels (prefill.cu): shared-memory matrix
// descriptors for 128-byte-swizzled fp16 tiles, the instruction descriptor of kind::f16, TMEM allocation and
// 32-lane x 32-bit loads/stores, commit -> mbarrier.  Encodings follow the PTX ISA "tcgen05" chapter (matrix
// descriptor: start address [0,14), leading byte offset [16,30), stride byte offset [32,46), version [46,48) = 1,
// swizzle mode [61,64); instruction descriptor: D format [4,6), A/B format [7,10)/[10,13), A/B major [15]/[16],
// N>>3 [17,23), M>>4 [24,29)).
#pragma once
#include <stdint.h>

namespace quest {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- 128-byte swizzle: a tile is stored as blocks of [rows][64 halves]; the 16-byte chunk c of row r lands at
// chunk (c ^ (r & 7)) of that row (the hardware applies the same XOR to address bits [4,7) with bits [7,10), so
// block bases must be 1024-byte aligned).
__device__ __forceinline__ uint32_t sw128_off(uint32_t r, uint32_t c) { return r * 128u + ((c ^ (r & 7u)) << 4); }

constexpr uint64_t kDescFixed = (1ull << 46) | (2ull << 61);  // descriptor version 1 (sm_100), SWIZZLE_128B

// Operand whose reduction (K) dimension is contiguous in shared memory: rows are M (or N) indices, one row = 64
// K-elements = 128 bytes, groups of 8 rows 1024 bytes apart (SBO).  A K=16 slice inside the row is addressed by
// adding 32 bytes per step to `saddr`.  LBO is not used by swizzled K-major layouts (canonical value 1).
__device__ __forceinline__ uint64_t desc_k_major(uint32_t saddr) {
  return kDescFixed | ((uint64_t)(1024u >> 4) << 32) | (1ull << 16) | (uint64_t)((saddr >> 4) & 0x3FFFu);
}
// Operand whose M/N dimension is contiguous: rows are K indices, one row = 64 MN-elements = 128 bytes, groups of 8
// K-rows 1024 bytes apart (SBO), 64-element MN blocks `mn_block_bytes` apart (LBO).
__device__ __forceinline__ uint64_t desc_mn_major(uint32_t saddr, uint32_t mn_block_bytes) {
  return kDescFixed | ((uint64_t)(1024u >> 4) << 32) | ((uint64_t)((mn_block_bytes >> 4) & 0x3FFFu) << 16) |
         (uint64_t)((saddr >> 4) & 0x3FFFu);
}

// kind::f16 instruction descriptor: fp16 A and B, fp32 accumulator, A K-major, B K-major (b_mn = 0) or MN-major (1).
__host__ __device__ constexpr uint32_t idesc_f16(uint32_t M, uint32_t N, uint32_t b_mn) {
  return (1u << 4) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.  accumulate == 0 overwrites D.
__device__ __forceinline__ void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
      :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Same with the A operand in TMEM (lane = row, 16-bit elements packed two per 32-bit column, 8 columns per K = 16).
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}"
      :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The mbarrier receives one arrival when every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void mma_commit(uint32_t bar_saddr) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar_saddr) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes (st.shared, cp.async) -> visible to the tensor core's async-proxy reads
__device__ __forceinline__ void fence_smem_to_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM.  One warp allocates (power of two >= 32 columns) and the same warp frees.
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t result_saddr) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(result_saddr), "n"(kCols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_free(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "n"(kCols) : "memory");
}

// 32 consecutive 32-bit columns of this thread's lane (warp w of the CTA owns lanes 32*(w%4) .. +31; taddr carries
// the lane base in bits [16,32) and the column in [0,16)).  Warp-collective.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31};"
      :: "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
         "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
         "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
         "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]), "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- mbarrier (shared::cta), used as the completion signal of tcgen05.commit
__device__ __forceinline__ void mbar_init(uint32_t bar_saddr, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar_saddr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Waits for the phase with the given parity to complete.  The spin is bounded (a lost arrival shows up as a wrong
// result and a set bit in *timeouts, not as a hung GPU); returns false on timeout.
__device__ __forceinline__ bool mbar_wait(uint32_t bar_saddr, uint32_t parity) {
  unsigned long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    uint32_t done;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(bar_saddr), "r"(parity)
        : "memory");
    if (done) return true;
    if ((spin & 255u) == 255u) {  // bounded by wall time (4 s of %globaltimer), not by a poll count
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000ull) return false;
    }
  }
}

// ---- cp.async (16 bytes, L2 only); src_bytes == 0 writes zeros without touching global memory
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* g, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(saddr), "l"(g), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }


// ---- mbarrier arrivals and an abort-aware wait for warp-specialised pipelines: the first wait that runs out sets
// *abort_flag (shared memory) and every later wait in the CTA returns at once, so a lost signal ends the kernel
// quickly with a wrong result (reported through a global counter) instead of hanging the GPU.
__device__ __forceinline__ void mbar_arrive(uint32_t bar_saddr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar_saddr) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar_saddr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_saddr), "r"(bytes) : "memory");
}
// Bounded by wall time (%globaltimer, kWaitTimeoutNs), not by a poll count: under time slicing, MPS or a profiler's
// replay a legitimate wait can take many polls.
constexpr unsigned long long kWaitTimeoutNs = 4000000000ull;
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool mbar_wait_abortable(uint32_t bar_saddr, uint32_t parity, volatile uint32_t* abort_flag) {
  unsigned long long t0 = 0;
  for (uint32_t spin = 0;; ++spin) {
    uint32_t done;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(done)
        : "r"(bar_saddr), "r"(parity)
        : "memory");
    if (done) return true;
    if ((spin & 63u) == 63u) {
      if (*abort_flag) return true;
      const unsigned long long now = global_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > kWaitTimeoutNs) break;
    }
  }
  *abort_flag = 1u;
  return false;  // this wait is the one that ran out
}

// ---- TMA: one box of a tiled tensor map -> shared memory, completion counted in bytes on an mbarrier.
// Coordinates are innermost first.  Issued by ONE thread.
__device__ __forceinline__ void tma_load_3d(uint32_t dst_saddr, const void* tmap, uint32_t bar_saddr, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2,%3,%4}], [%5];"
      :: "r"(dst_saddr), "l"((uint64_t)tmap), "r"(c0), "r"(c1), "r"(c2), "r"(bar_saddr)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst_saddr, const void* tmap, uint32_t bar_saddr, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2,%3,%4,%5,%6}], [%7];"
      :: "r"(dst_saddr), "l"((uint64_t)tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "r"(bar_saddr)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" :: "l"((uint64_t)tmap) : "memory");
}

// ---- packed fp32x2 arithmetic (FFMA2 / FADD2) and the 3-input maximum (FMNMX3) of sm_100
__device__ __forceinline__ uint64_t pack_f32x2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t add_f32x2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

}  // namespace umma
}  // namespace quest
