// Issue-rate probe: mixed-precision FMA (fma.rn.f32.f16 -> FHFMA) against FFMA and against cvt + FFMA, per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fhfma_probe tools/fhfma_probe.cu && tools/fhfma_probe
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdint.h>

__device__ __forceinline__ float fhfma_lo(uint32_t a, uint32_t b, float c) {
  float d;
  asm volatile("{\n.reg .b16 al, ah, bl, bh;\nmov.b32 {al, ah}, %1;\nmov.b32 {bl, bh}, %2;\nfma.rn.f32.f16 %0, al, bl, %3;\n}" : "=f"(d) : "r"(a), "r"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float fhfma_hi(uint32_t a, uint32_t b, float c) {
  float d;
  asm volatile("{\n.reg .b16 al, ah, bl, bh;\nmov.b32 {al, ah}, %1;\nmov.b32 {bl, bh}, %2;\nfma.rn.f32.f16 %0, ah, bh, %3;\n}" : "=f"(d) : "r"(a), "r"(b), "f"(c));
  return d;
}

template <int MODE>
__global__ void __launch_bounds__(1024, 1) probe(const uint32_t* in, float* out, int iters) {
  uint32_t a = in[threadIdx.x], b = in[threadIdx.x + 1024];
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = (float)i;
  float fa = __uint_as_float(a), fb = __uint_as_float(b);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) {  // FFMA
        asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(fa), "f"(fb));
      } else if (MODE == 1) {  // FHFMA, alternating halves
        acc[i] = (i & 1) ? fhfma_hi(a, b, acc[i]) : fhfma_lo(a, b, acc[i]);
      } else {  // cvt + FFMA
        float x;
        asm volatile("{\n.reg .b16 l, h;\nmov.b32 {l, h}, %1;\ncvt.f32.f16 %0, l;\n}" : "=f"(x) : "r"(b ^ (uint32_t)i));
        asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(fa), "f"(x));
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  out[blockIdx.x * 1024 + threadIdx.x] = s;
}

template <int MODE>
static void run(const char* name, const uint32_t* in, float* out) {
  const int iters = 4096, blocks = 148;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  probe<MODE><<<blocks, 1024>>>(in, out, iters);
  cudaEventRecord(e0);
  probe<MODE><<<blocks, 1024>>>(in, out, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double fmas = (double)blocks * 1024 * iters * 8;
  printf("%-12s %.3f ms  %.1f FMA/clk/SM at 1.965 GHz\n", name, ms, fmas / (ms * 1e-3) / 148 / 1.965e9);
}

int main() {
  uint32_t* in;
  float* out;
  cudaMalloc(&in, 2048 * 4);
  cudaMemset(in, 0x3c, 2048 * 4);
  cudaMalloc(&out, 148 * 1024 * 4);
  run<0>("FFMA", in, out);
  run<1>("FHFMA", in, out);
  run<2>("cvt+FFMA", in, out);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
