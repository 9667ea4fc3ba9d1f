// HBM access-pattern probe for the Quest path on B200 (sm_100a).  Not part of the product: it measures
// what each load mechanism can reach for the three access patterns the path needs, so the kernels'
// design is chosen from data:
//   contiguous : the whole metadata pool as one stream (estimate, page-major)
//   headmajor  : 32-row x 256-B boxes with an 8 KiB row stride (estimate, head-major / KV page of one head)
//   gather     : the same boxes at random pages (sparse decode)
// Mechanisms: LDG.128 to registers, cp.async.bulk 1-D into an mbarrier ring, TMA 2-D boxes into the ring,
// per-row 256-B cp.async.bulk into the ring.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o bw_probe bw_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <algorithm>
#include <random>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma2d_g2s(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}

// ---------------------------------------------------------------------------------------------- LDG
// one-shot: each thread issues U independent 16-B loads then reduces.  grid covers the buffer.
template <int U>
__global__ void __launch_bounds__(256) k_ldg_oneshot(const uint4* __restrict__ src, size_t nvec, uint32_t* sink) {
  size_t base = (size_t)blockIdx.x * 256 * U + threadIdx.x;
  uint4 v[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    size_t i = base + (size_t)u * 256;
    v[u] = make_uint4(0, 0, 0, 0);
    if (i < nvec) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i));
  }
  uint32_t a = 0;
#pragma unroll
  for (int u = 0; u < U; ++u) a ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  if (a == 0x12345678u) *sink = a;
}

// persistent grid-stride with U loads in flight per thread
template <int U>
__global__ void __launch_bounds__(256) k_ldg_persist(const uint4* __restrict__ src, size_t nvec, uint32_t* sink) {
  uint32_t a = 0;
  for (size_t base = (size_t)blockIdx.x * 256 * U + threadIdx.x; base < nvec; base += (size_t)gridDim.x * 256 * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      size_t i = base + (size_t)u * 256;
      v[u] = make_uint4(0, 0, 0, 0);
      if (i < nvec) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) a ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (a == 0x12345678u) *sink = a;
}

template <int U, int THREADS, int PDL>
__global__ void __launch_bounds__(THREADS) k_ldg_persist2(const uint4* __restrict__ src, size_t nvec, uint32_t* sink) {
  uint32_t a = 0;
  if (PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
  if (PDL == 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  for (size_t base = (size_t)blockIdx.x * THREADS * U + threadIdx.x; base < nvec; base += (size_t)gridDim.x * THREADS * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      size_t i = base + (size_t)u * THREADS;
      v[u] = make_uint4(0, 0, 0, 0);
      if (i < nvec) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) a ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (a == 0x12345678u) *sink = a;
}

// ---------------------------------------------------------------------------------------------- ring kernels
// Unit list: each unit is `unit_bytes` landing in one ring stage.  mode 0: 1-D bulk from src + off[u];
// mode 1: one TMA 2-D box at (c0,c1) = (colrow[2u], colrow[2u+1]); mode 2: 32 row copies of 256 B at
// src + off[u] + r*row_stride.  Units are dealt round-robin (u = blockIdx.x + i*gridDim.x) or in
// contiguous blocks per CTA (blocked=1).
struct RingArgs {
  const uint8_t* src;
  const uint64_t* off;
  const int32_t* colrow;
  uint32_t n_units, unit_bytes, stages, mode, blocked, row_stride, poll1;
};

__global__ void __launch_bounds__(288) k_ring(const __grid_constant__ CUtensorMap map, RingArgs a, uint32_t* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + 32;
  uint8_t* data = smem + 1024;
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t ncons_warps = (blockDim.x >> 5) - 1;
  if (tid == 0) {
    for (uint32_t s = 0; s < a.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, ncons_warps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t u0, u1, ustep;
  if (a.blocked) {
    uint32_t per = (a.n_units + gridDim.x - 1) / gridDim.x;
    u0 = blockIdx.x * per; u1 = min(u0 + per, a.n_units); ustep = 1;
  } else { u0 = blockIdx.x; u1 = a.n_units; ustep = gridDim.x; }
  if (warp == ncons_warps) {  // producer warp
    uint32_t s = 0, ph = 0;
    for (uint32_t u = u0; u < u1; u += ustep) {
      if (lane == 0) {
        mbar_wait(empty + s, ph ^ 1);
        mbar_expect(full + s, a.unit_bytes);
      }
      __syncwarp();
      uint8_t* dst = data + (size_t)s * a.unit_bytes;
      if (a.mode == 0) {
        if (lane == 0) bulk_g2s(dst, a.src + a.off[u], a.unit_bytes, full + s);
      } else if (a.mode == 1) {
        if (lane == 0) tma2d_g2s(dst, &map, a.colrow[2 * u], a.colrow[2 * u + 1], full + s);
      } else {
        const uint32_t rows = a.unit_bytes / 256;
        for (uint32_t r = lane; r < rows; r += 32) bulk_g2s(dst + r * 256, a.src + a.off[u] + (size_t)r * a.row_stride, 256, full + s);
      }
      if (++s == a.stages) { s = 0; ph ^= 1; }
    }
  } else {
    uint32_t s = 0, ph = 0, acc = 0;
    const uint32_t nthr = ncons_warps * 32;
    for (uint32_t u = u0; u < u1; u += ustep) {
      if (a.poll1) { if (lane == 0) mbar_wait(full + s, ph); __syncwarp(); } else mbar_wait(full + s, ph);
      const uint4* p = reinterpret_cast<const uint4*>(data + (size_t)s * a.unit_bytes);
      for (uint32_t i = tid; i < a.unit_bytes / 16; i += nthr) { uint4 v = p[i]; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty + s);
      if (++s == a.stages) { s = 0; ph ^= 1; }
    }
    if (acc == 0x12345678u) *sink = acc;
  }
}

// gather with LDG: unit = 32 rows x 256 B; a 16-lane group reads one row (what the v1 decode kernel does)
template <int U>
__global__ void __launch_bounds__(128) k_gather_ldg(const uint8_t* __restrict__ src, const uint64_t* __restrict__ off, uint32_t n_units, uint32_t units_per_cta, uint32_t row_stride, uint32_t* sink) {
  const uint32_t tid = threadIdx.x, g = tid >> 4, l = tid & 15;  // 8 row groups
  uint32_t a = 0;
  const uint32_t u0 = blockIdx.x * units_per_cta, u1 = min(u0 + units_per_cta, n_units);
  for (uint32_t u = u0; u < u1; u += U / 4 ? U / 4 : 1) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {  // U rows per group per iteration => U*8 rows = U/4 units
      uint32_t uu = u + (j * 8 + g) / 32, r = (j * 8 + g) % 32;
      v[j] = make_uint4(0, 0, 0, 0);
      if (uu < u1) {
        const uint8_t* p = src + __ldg(off + uu) + (size_t)r * row_stride + l * 16;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(p));
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) a ^= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
  }
  if (a == 0x12345678u) *sink = a;
}

static uint32_t g_poll1 = 0;
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int NBUF = 8;                       // rotate over 8 buffers (> L2)
  const size_t H = 32, D = 128, ROWB = H * D * 2;          // 8 KiB per token row
  const size_t PAGES = 2048, PAGE_ROWS = 32;               // a "page" = 32 rows (K+V or max+min) = 256 KiB
  const size_t BUF = PAGES * PAGE_ROWS * ROWB;             // 512 MiB per buffer (a KV pool); metadata = first 128 pages
  uint8_t* d;
  CK(cudaMalloc(&d, BUF * NBUF));
  CK(cudaMemset(d, 1, BUF * NBUF));
  uint32_t* sink; CK(cudaMalloc(&sink, 4));
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaStream_t st; CK(cudaStreamCreate(&st));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  EncodeFn enc; cudaDriverEntryPointQueryResult qr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &qr));
  CUtensorMap maps[NBUF];
  for (int b = 0; b < NBUF; ++b) {
    cuuint64_t dims[2] = {H * D, PAGES * PAGE_ROWS};
    cuuint64_t strides[1] = {ROWB};
    cuuint32_t box[2] = {128, 32};
    cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&maps[b], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d + b * BUF, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  }
  const int ITERS = 40;
  auto report = [&](const char* name, size_t bytes, float ms) {
    printf("%-64s %8.2f us  %8.1f GB/s\n", name, ms * 1e3 / ITERS, bytes / (ms / ITERS * 1e-3) / 1e9);
  };
  const size_t META = 128 * PAGE_ROWS * ROWB;  // 32 MiB: the metadata pool at 32K context

  // ---- contiguous, LDG
#define RUN_ONESHOT(U) { size_t nvec = META / 16; int grid = (int)((nvec + 256 * U - 1) / (256 * U)); \
    for (int w = 0; w < 3; ++w) k_ldg_oneshot<U><<<grid, 256, 0, st>>>((const uint4*)(d + (w % NBUF) * BUF), nvec, sink); \
    CK(cudaEventRecord(e0, st)); for (int i = 0; i < ITERS; ++i) k_ldg_oneshot<U><<<grid, 256, 0, st>>>((const uint4*)(d + (i % NBUF) * BUF), nvec, sink); \
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); char nm[96]; snprintf(nm, 96, "contig LDG one-shot U=%d grid=%d", U, grid); report(nm, META, ms); }
  RUN_ONESHOT(4) RUN_ONESHOT(8) RUN_ONESHOT(16)
#define RUN_PERSIST(U, CPS) { size_t nvec = META / 16; int grid = sms * CPS; \
    for (int w = 0; w < 3; ++w) k_ldg_persist<U><<<grid, 256, 0, st>>>((const uint4*)(d + (w % NBUF) * BUF), nvec, sink); \
    CK(cudaEventRecord(e0, st)); for (int i = 0; i < ITERS; ++i) k_ldg_persist<U><<<grid, 256, 0, st>>>((const uint4*)(d + (i % NBUF) * BUF), nvec, sink); \
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); char nm[96]; snprintf(nm, 96, "contig LDG persistent U=%d ctas/sm=%d", U, CPS); report(nm, META, ms); }
  RUN_PERSIST(4, 2) RUN_PERSIST(8, 1) RUN_PERSIST(8, 2) RUN_PERSIST(8, 4) RUN_PERSIST(16, 2)

  // ---- ring kernels
  std::mt19937 rng(1);
  uint64_t* d_off; int32_t* d_cr;
  CK(cudaMalloc(&d_off, sizeof(uint64_t) * 65536)); CK(cudaMalloc(&d_cr, sizeof(int32_t) * 2 * 65536));
  CK(cudaFuncSetAttribute(k_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  auto run_ring = [&](const char* name, std::vector<uint64_t>& off, std::vector<int32_t>& cr, uint32_t unit_bytes, uint32_t stages, uint32_t mode, uint32_t blocked, int grid, int threads) {
    CK(cudaMemcpy(d_off, off.data(), off.size() * 8, cudaMemcpyHostToDevice));
    if (!cr.empty()) CK(cudaMemcpy(d_cr, cr.data(), cr.size() * 4, cudaMemcpyHostToDevice));
    size_t smem = 1024 + (size_t)unit_bytes * stages;
    RingArgs a; a.off = d_off; a.colrow = d_cr; a.n_units = (uint32_t)off.size(); a.unit_bytes = unit_bytes; a.stages = stages; a.mode = mode; a.blocked = blocked; a.row_stride = (uint32_t)ROWB; a.poll1 = g_poll1;
    for (int w = 0; w < 3; ++w) { a.src = d + (w % NBUF) * BUF; k_ring<<<grid, threads, smem, st>>>(maps[w % NBUF], a, sink); }
    CK(cudaGetLastError());
    CK(cudaEventRecord(e0, st));
    for (int i = 0; i < ITERS; ++i) { a.src = d + (i % NBUF) * BUF; k_ring<<<grid, threads, smem, st>>>(maps[i % NBUF], a, sink); }
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    char nm[128]; snprintf(nm, 128, "%s unit=%uK st=%u grid=%d thr=%d %s poll1=%d", name, unit_bytes / 1024, stages, grid, threads, blocked ? "blocked" : "rr", g_poll1);
    report(nm, (size_t)unit_bytes * off.size(), ms);
  };
  std::vector<int32_t> nocr;
  for (uint32_t ub : {8192u, 16384u, 32768u}) {
    std::vector<uint64_t> off; for (size_t o = 0; o + ub <= META; o += ub) off.push_back(o);
    for (int cps : {1, 2}) for (uint32_t blocked : {0u, 1u}) {
      uint32_t stages = std::min<uint32_t>(16, (200 * 1024 / cps) / ub);
      run_ring("contig bulk1d", off, nocr, ub, stages, 0, blocked, sms * cps, 288);
    }
  }
  { std::vector<uint64_t> off; for (size_t o = 0; o + 16384 <= META; o += 16384) off.push_back(o);
    run_ring("contig bulk1d", off, nocr, 16384, 4, 0, 1, sms * 2, 288);
    run_ring("contig bulk1d", off, nocr, 16384, 3, 0, 1, sms * 4, 160);
    run_ring("contig bulk1d", off, nocr, 16384, 6, 0, 1, sms * 2, 160); }
  // head-major metadata: unit = (head h, metadata page m): 32 rows x 256 B
  for (int order = 0; order < 2; ++order) {
    std::vector<uint64_t> off; std::vector<int32_t> cr;
    if (order == 0) { for (size_t h = 0; h < H; ++h) for (size_t m = 0; m < 128; ++m) { off.push_back(m * PAGE_ROWS * ROWB + h * 256); cr.push_back((int)(h * 128)); cr.push_back((int)(m * 32)); } }
    else { for (size_t m = 0; m < 128; ++m) for (size_t h = 0; h < H; ++h) { off.push_back(m * PAGE_ROWS * ROWB + h * 256); cr.push_back((int)(h * 128)); cr.push_back((int)(m * 32)); } }
    const char* on = order == 0 ? "headmajor(h outer)" : "headmajor(m outer)";
    char nm[64];
    for (int cps : {1, 2}) for (uint32_t blocked : {0u, 1u}) {
      snprintf(nm, 64, "%s tma2d", on); run_ring(nm, off, cr, 8192, std::min(16, 24 / cps), 1, blocked, sms * cps, 288);
      snprintf(nm, 64, "%s rows256", on); run_ring(nm, off, cr, 8192, std::min(16, 24 / cps), 2, blocked, sms * cps, 288);
    }
  }
  // gather: 32 heads x 128 random pages of the 2048-page pool
  {
    std::vector<uint64_t> off; std::vector<int32_t> cr;
    for (size_t h = 0; h < H; ++h) {
      std::vector<int> perm(PAGES); for (size_t i = 0; i < PAGES; ++i) perm[i] = (int)i; std::shuffle(perm.begin(), perm.end(), rng);
      std::sort(perm.begin(), perm.begin() + 128);
      for (int i = 0; i < 128; ++i) { off.push_back((size_t)perm[i] * PAGE_ROWS * ROWB + h * 256); cr.push_back((int)(h * 128)); cr.push_back(perm[i] * 32); }
    }
    for (int cps : {1, 2, 4}) for (uint32_t blocked : {0u, 1u}) {
      run_ring("gather tma2d", off, cr, 8192, std::min(16, 24 / cps), 1, blocked, sms * cps, cps == 4 ? 160 : 288);
      run_ring("gather rows256", off, cr, 8192, std::min(16, 24 / cps), 2, blocked, sms * cps, cps == 4 ? 160 : 288);
    }
    CK(cudaMemcpy(d_off, off.data(), off.size() * 8, cudaMemcpyHostToDevice));
#define RUN_GLDG(U, UPC) { uint32_t n = (uint32_t)off.size(); int grid = (n + UPC - 1) / UPC; \
    for (int w = 0; w < 3; ++w) k_gather_ldg<U><<<grid, 128, 0, st>>>(d + (w % NBUF) * BUF, d_off, n, UPC, (uint32_t)ROWB, sink); \
    CK(cudaEventRecord(e0, st)); for (int i = 0; i < ITERS; ++i) k_gather_ldg<U><<<grid, 128, 0, st>>>(d + (i % NBUF) * BUF, d_off, n, UPC, (uint32_t)ROWB, sink); \
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); char nm[96]; snprintf(nm, 96, "gather LDG U=%d units/cta=%d grid=%d", U, UPC, grid); report(nm, (size_t)n * 8192, ms); }
    RUN_GLDG(8, 2) RUN_GLDG(8, 4) RUN_GLDG(16, 4) RUN_GLDG(16, 8) RUN_GLDG(8, 8)
  }
  // ---- second batch: ring with single-lane polling, deeper rings
  g_poll1 = 1;
  { std::vector<uint64_t> off; for (size_t o = 0; o + 8192 <= META; o += 8192) off.push_back(o);
    run_ring("contig bulk1d", off, nocr, 8192, 16, 0, 0, sms, 288);
    run_ring("contig bulk1d", off, nocr, 8192, 24, 0, 0, sms, 288);
    run_ring("contig bulk1d", off, nocr, 8192, 24, 0, 1, sms, 288);
    run_ring("contig bulk1d", off, nocr, 8192, 12, 0, 0, sms * 2, 288);
    run_ring("contig bulk1d", off, nocr, 8192, 12, 0, 0, sms * 2, 160);
    run_ring("contig bulk1d", off, nocr, 8192, 6, 0, 0, sms * 4, 160); }
  { std::vector<uint64_t> off; for (size_t o = 0; o + 16384 <= META; o += 16384) off.push_back(o);
    run_ring("contig bulk1d", off, nocr, 16384, 12, 0, 0, sms, 288);
    run_ring("contig bulk1d", off, nocr, 16384, 6, 0, 0, sms * 2, 288); }
  g_poll1 = 0;
  // ---- persistent LDG with big CTAs; stream launches, then CUDA graph chains with / without PDL
#define RUN_P2(U, T, GRID) { size_t nvec = META / 16; \
    for (int w = 0; w < 3; ++w) k_ldg_persist2<U, T, 0><<<GRID, T, 0, st>>>((const uint4*)(d + (w % NBUF) * BUF), nvec, sink); \
    CK(cudaEventRecord(e0, st)); for (int i = 0; i < ITERS; ++i) k_ldg_persist2<U, T, 0><<<GRID, T, 0, st>>>((const uint4*)(d + (i % NBUF) * BUF), nvec, sink); \
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); char nm[96]; snprintf(nm, 96, "contig LDG persistent2 U=%d thr=%d grid=%d", U, T, GRID); report(nm, META, ms); }
  RUN_P2(8, 1024, sms) RUN_P2(4, 1024, sms) RUN_P2(8, 512, sms * 2) RUN_P2(4, 512, sms * 2) RUN_P2(4, 512, sms * 4) RUN_P2(8, 256, sms * 4) RUN_P2(4, 256, sms * 8)
  // graph chains: 32 kernels per graph, each streaming a different buffer
#define RUN_GRAPH(PDL, NAME) { size_t nvec = META / 16; cudaGraph_t g; cudaGraphExec_t ge; \
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal)); \
    for (int i = 0; i < 32; ++i) { cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(sms * 4); cfg.blockDim = dim3(256); cfg.stream = st; \
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1; \
      cfg.attrs = at; cfg.numAttrs = PDL ? 1 : 0; const uint4* sp = (const uint4*)(d + (i % NBUF) * BUF); \
      CK(cudaLaunchKernelEx(&cfg, k_ldg_persist2<8, 256, PDL>, sp, nvec, sink)); } \
    CK(cudaStreamEndCapture(st, &g)); CK(cudaGraphInstantiate(&ge, g, 0)); \
    for (int w = 0; w < 3; ++w) CK(cudaGraphLaunch(ge, st)); \
    CK(cudaEventRecord(e0, st)); for (int i = 0; i < 10; ++i) CK(cudaGraphLaunch(ge, st)); \
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); \
    printf("%-64s %8.2f us  %8.1f GB/s\n", NAME, ms * 1e3 / 320, META / (ms / 320 * 1e-3) / 1e9); }
  RUN_GRAPH(0, "graph chain x32, plain edges, LDG persistent U=8 4 ctas/sm")
  RUN_GRAPH(1, "graph chain x32, PDL edges (wait at top)")
  RUN_GRAPH(2, "graph chain x32, PDL edges (wait + early trigger)")
#define RUN_GRAPH_TINY(PDL, NAME) { cudaGraph_t g; cudaGraphExec_t ge; \
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal)); \
    for (int i = 0; i < 32; ++i) { cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(32); cfg.blockDim = dim3(256); cfg.stream = st; \
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1; \
      cfg.attrs = at; cfg.numAttrs = PDL ? 1 : 0; const uint4* sp = (const uint4*)(d + (i % NBUF) * BUF); \
      CK(cudaLaunchKernelEx(&cfg, k_ldg_persist2<1, 256, PDL>, sp, (size_t)8192, sink)); } \
    CK(cudaStreamEndCapture(st, &g)); CK(cudaGraphInstantiate(&ge, g, 0)); \
    for (int w = 0; w < 3; ++w) CK(cudaGraphLaunch(ge, st)); \
    CK(cudaEventRecord(e0, st)); for (int i = 0; i < 10; ++i) CK(cudaGraphLaunch(ge, st)); \
    CK(cudaEventRecord(e1, st)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); \
    printf("%-64s %8.2f us per kernel\n", NAME, ms * 1e3 / 320); }
  RUN_GRAPH_TINY(0, "graph chain x32 of tiny kernels (128 KB read), plain edges")
  RUN_GRAPH_TINY(1, "graph chain x32 of tiny kernels, PDL wait at top")
  RUN_GRAPH_TINY(2, "graph chain x32 of tiny kernels, PDL wait + trigger")
  // launch floor: empty-ish kernel
  { std::vector<uint64_t> off(1, 0); run_ring("launch floor (1 unit)", off, nocr, 8192, 2, 0, 0, 1, 288); }
  printf("done\n");
  return 0;
}
