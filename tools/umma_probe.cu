// Known-answer test of the tcgen05 building blocks in quest_b200/csrc/umma.cuh, one CTA, against a host GEMM:
//   T1  S[128x64]  = Q[128xD] * K[64xD]^T      (A and B K-major, 128B swizzle; D = 128 -> two 64-wide blocks)
//   T2  O[128xD]   = P[128x64] * V[64xD]       (A K-major, B MN-major), issued twice with accumulate -> 2x
//   T3  O *= 0.5 through tcgen05.ld / tcgen05.st, read back
//   T4  O[128xD]   = P[128x64] * V[64xD] with P in TENSOR MEMORY (fp16 pairs stored with tcgen05.st)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o tools/umma_probe tools/umma_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../quest_b200/csrc/umma.cuh"

using namespace quest::umma;

constexpr int D = 128, BQ = 128, BK = 64;

__global__ void __launch_bounds__(128) probe(const __half* q, const __half* k, const __half* p, const __half* v, float* s_out,
                                             float* o_out, float* o_half, float* o_ts, uint32_t* status) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* q_s = smem;                    // 2 blocks x [128][128 B]
  uint8_t* k_s = q_s + 2 * BQ * 128;      // 2 blocks x [64][128 B]
  uint8_t* v_s = k_s + 2 * BK * 128;      // 2 blocks x [64][128 B]
  uint8_t* p_s = v_s + 2 * BK * 128;      // 1 block  x [128][128 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(p_s + BQ * 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const uint32_t t = threadIdx.x, warp = t / 32;
  if (t == 0) status[0] = smem_u32(smem) & 1023u;
  if (warp == 0) tmem_alloc<256>(smem_u32(tmem_slot));
  if (t == 0) {
    mbar_init(smem_u32(bars), 1);
    mbar_init(smem_u32(bars + 1), 1);
    mbar_init_fence();
  }
  // stage the tiles (row-major fp16 in global) into the swizzled blocks
  for (uint32_t i = t; i < BQ * (D / 8); i += 128) {
    const uint32_t r = i / (D / 8), c = i % (D / 8);
    *reinterpret_cast<uint4*>(q_s + (c / 8) * BQ * 128 + sw128_off(r, c % 8)) = *reinterpret_cast<const uint4*>(q + r * D + c * 8);
  }
  for (uint32_t i = t; i < BK * (D / 8); i += 128) {
    const uint32_t r = i / (D / 8), c = i % (D / 8);
    *reinterpret_cast<uint4*>(k_s + (c / 8) * BK * 128 + sw128_off(r, c % 8)) = *reinterpret_cast<const uint4*>(k + r * D + c * 8);
    *reinterpret_cast<uint4*>(v_s + (c / 8) * BK * 128 + sw128_off(r, c % 8)) = *reinterpret_cast<const uint4*>(v + r * D + c * 8);
  }
  for (uint32_t i = t; i < BQ * (BK / 8); i += 128) {
    const uint32_t r = i / (BK / 8), c = i % (BK / 8);
    *reinterpret_cast<uint4*>(p_s + sw128_off(r, c)) = *reinterpret_cast<const uint4*>(p + r * BK + c * 8);
  }
  fence_smem_to_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tbase = *tmem_slot;
  if (t == 0) status[1] = tbase;
  const uint32_t t_s = tbase, t_o = tbase + 64;
  const uint32_t lane_addr = (warp * 32u) << 16;
  uint32_t ok = 1;

  // T1
  if (t == 0) {
    for (uint32_t kk = 0; kk < D / 16; ++kk) {
      const uint64_t a = desc_k_major(smem_u32(q_s) + (kk / 4) * BQ * 128 + (kk % 4) * 32);
      const uint64_t b = desc_k_major(smem_u32(k_s) + (kk / 4) * BK * 128 + (kk % 4) * 32);
      mma_f16_ss(t_s, a, b, idesc_f16(BQ, BK, 0), kk > 0);
    }
    mma_commit(smem_u32(bars));
  }
  ok &= mbar_wait(smem_u32(bars), 0);
  fence_after_sync();
  uint32_t r[32];
  for (uint32_t c0 = 0; c0 < BK; c0 += 32) {
    tmem_ld32(lane_addr + t_s + c0, r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) s_out[t * BK + c0 + i] = __uint_as_float(r[i]);
  }

  // T2: twice, second time accumulating
  if (t == 0) {
    for (uint32_t rep = 0; rep < 2; ++rep)
      for (uint32_t kk = 0; kk < BK / 16; ++kk) {
        const uint64_t a = desc_k_major(smem_u32(p_s) + kk * 32);
        const uint64_t b = desc_mn_major(smem_u32(v_s) + kk * 2048, BK * 128);
        mma_f16_ss(t_o, a, b, idesc_f16(BQ, D, 1), (rep | kk) > 0);
      }
    mma_commit(smem_u32(bars + 1));
  }
  ok &= mbar_wait(smem_u32(bars + 1), 0);
  fence_after_sync();
  for (uint32_t c0 = 0; c0 < D; c0 += 32) {
    tmem_ld32(lane_addr + t_o + c0, r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) o_out[t * D + c0 + i] = __uint_as_float(r[i]);
  }
  // T3
  for (uint32_t c0 = 0; c0 < D; c0 += 32) {
    tmem_ld32(lane_addr + t_o + c0, r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * 0.5f);
    tmem_st32(lane_addr + t_o + c0, r);
  }
  tmem_st_wait();
  for (uint32_t c0 = 0; c0 < D; c0 += 32) {
    tmem_ld32(lane_addr + t_o + c0, r);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) o_half[t * D + c0 + i] = __uint_as_float(r[i]);
  }
  // T4: this thread's row of P as packed fp16 pairs -> TMEM columns [0,32) of the S region, then the TS form
  {
    uint32_t pw[32];
    for (int i = 0; i < 32; ++i) pw[i] = *reinterpret_cast<const uint32_t*>(p + t * BK + 2 * i);
    tmem_st32(lane_addr + t_s, pw);
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
      fence_after_sync();
      for (uint32_t kk = 0; kk < BK / 16; ++kk) {
        const uint64_t b = desc_mn_major(smem_u32(v_s) + kk * 2048, BK * 128);
        mma_f16_ts(t_o, t_s + kk * 8, b, idesc_f16(BQ, D, 1), kk > 0);
      }
      mma_commit(smem_u32(bars));
    }
    ok &= mbar_wait(smem_u32(bars), 1);
    fence_after_sync();
    for (uint32_t c0 = 0; c0 < D; c0 += 32) {
      tmem_ld32(lane_addr + t_o + c0, r);
      tmem_ld_wait();
      for (int i = 0; i < 32; ++i) o_ts[t * D + c0 + i] = __uint_as_float(r[i]);
    }
  }
  if (!ok) atomicOr(status + 2, 1u);
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_free<256>(tbase);
}

#define CK(x)                                                                    \
  do {                                                                           \
    cudaError_t e = (x);                                                         \
    if (e != cudaSuccess) {                                                      \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      return 2;                                                                  \
    }                                                                            \
  } while (0)

int main() {
  std::vector<__half> q(BQ * D), k(BK * D), p(BQ * BK), v(BK * D);
  srand(7);
  auto rnd = [] { return (float)(rand() % 2001 - 1000) / 1000.f; };
  for (auto& x : q) x = __float2half(rnd());
  for (auto& x : k) x = __float2half(rnd());
  for (auto& x : p) x = __float2half(rnd());
  for (auto& x : v) x = __float2half(rnd());
  __half *dq, *dk, *dp, *dv;
  float *ds, *dout, *dhalf, *dts;
  uint32_t* dst;
  CK(cudaMalloc(&dq, q.size() * 2));
  CK(cudaMalloc(&dk, k.size() * 2));
  CK(cudaMalloc(&dp, p.size() * 2));
  CK(cudaMalloc(&dv, v.size() * 2));
  CK(cudaMalloc(&ds, BQ * BK * 4));
  CK(cudaMalloc(&dout, BQ * D * 4));
  CK(cudaMalloc(&dhalf, BQ * D * 4));
  CK(cudaMalloc(&dts, BQ * D * 4));
  CK(cudaMalloc(&dst, 16));
  CK(cudaMemset(dst, 0, 16));
  CK(cudaMemcpy(dq, q.data(), q.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dk, k.data(), k.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dp, p.data(), p.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dv, v.data(), v.size() * 2, cudaMemcpyHostToDevice));
  const int smem = 2 * BQ * 128 + 4 * BK * 128 + BQ * 128 + 64;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  probe<<<1, 128, smem>>>(dq, dk, dp, dv, ds, dout, dhalf, dts, dst);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  std::vector<float> s(BQ * BK), o(BQ * D), oh(BQ * D), ots(BQ * D);
  uint32_t st[4];
  CK(cudaMemcpy(s.data(), ds, s.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(oh.data(), dhalf, oh.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(ots.data(), dts, ots.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(st, dst, 16, cudaMemcpyDeviceToHost));
  printf("smem base & 1023 = %u, tmem base = 0x%08x, barrier timeouts = %u\n", st[0], st[1], st[2]);
  double e1 = 0, e2 = 0, e3 = 0, e4 = 0;
  for (int i = 0; i < BQ; ++i)
    for (int j = 0; j < BK; ++j) {
      double acc = 0;
      for (int d = 0; d < D; ++d) acc += (double)__half2float(q[i * D + d]) * __half2float(k[j * D + d]);
      e1 = fmax(e1, fabs(acc - s[i * BK + j]));
    }
  for (int i = 0; i < BQ; ++i)
    for (int d = 0; d < D; ++d) {
      double acc = 0;
      for (int j = 0; j < BK; ++j) acc += (double)__half2float(p[i * BK + j]) * __half2float(v[j * D + d]);
      e2 = fmax(e2, fabs(2 * acc - o[i * D + d]));
      e3 = fmax(e3, fabs(acc - oh[i * D + d]));
      e4 = fmax(e4, fabs(acc - ots[i * D + d]));
    }
  printf("T1 S=QK^T  max|err| = %.3e\nT2 O=2PV   max|err| = %.3e\nT3 O*0.5   max|err| = %.3e\nT4 O=PV(P in TMEM) max|err| = %.3e\n", e1,
         e2, e3, e4);
  const bool pass = e1 < 1e-3 && e2 < 1e-3 && e3 < 1e-3 && e4 < 1e-3 && st[0] == 0 && st[2] == 0;
  printf("%s\n", pass ? "UMMA PROBE PASS" : "UMMA PROBE FAIL");
  if (!pass) {  // a few entries help tell a layout error from a scaling one
    printf("S[0][0..3] = %f %f %f %f\n", s[0], s[1], s[2], s[3]);
    printf("O[0][0..3] = %f %f %f %f\n", o[0], o[1], o[2], o[3]);
  }
  return pass ? 0 : 1;
}
