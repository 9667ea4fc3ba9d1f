// Experiment harness (not product): cycle stamps inside topk_select (quest_b200/csrc/topk.cuh).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DQUEST_TOPK_CLOCKS -o kbench_topk kbench_topk.cu
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "../quest_b200/csrc/topk.cuh"
namespace quest { unsigned long long g_launch_count = 0; int pdl_enabled() { return 0; } }
using namespace quest;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

template <int W, int KPC>
__global__ void __launch_bounds__(32 * W) k(const uint16_t* scores, const int32_t* in_idx, int32_t* out_idx, uint32_t len, uint32_t kk) {
  extern __shared__ __align__(16) uint16_t keys[];
  __shared__ TopkScratch<W> sc;
  const uint32_t tid = threadIdx.x, head = blockIdx.x, T = 32 * W, len8 = (len + 7) & ~7u;
  const uint16_t* s = scores + (size_t)head * len;
  for (uint32_t i = tid; i < len8; i += T) keys[i] = i < len ? key_of(s[i]) : (uint16_t)0;
  __syncthreads();
  topk_select<W, KPC>(keys, len, kk, tid, 1, &sc, in_idx, out_idx + (size_t)head * kk, nullptr);
}

template <int W, int KPC>
void run(const char* name, const uint16_t* sc, const int32_t* idx, int32_t* out, uint32_t len, uint32_t kk) {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  size_t smem = ((len + 7) & ~7u) * 2;
  for (int i = 0; i < 3; ++i) k<W, KPC><<<32, 32 * W, smem>>>(sc, idx, out, len, kk);
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 50; ++i) k<W, KPC><<<32, 32 * W, smem>>>(sc, idx, out, len, kk);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  long long clk[64]; CK(cudaMemcpyFromSymbol(clk, g_topk_clk, sizeof(clk)));
  printf("%s: %.2f us/launch (stream launches)\n", name, ms * 1e3 / 50);
  for (int p = 0; p < 4; ++p) {
    printf("  pass %d: count %lld  redux %lld  sts+bar %lld  gather %lld  scan %lld", p, clk[p * 8 + 1] - clk[p * 8], clk[p * 8 + 2] - clk[p * 8 + 1],
           clk[p * 8 + 3] - clk[p * 8 + 2], clk[p * 8 + 4] - clk[p * 8 + 3], clk[p * 8 + 5] - clk[p * 8 + 4]);
    if (p < 3) printf("  -> next %lld", clk[(p + 1) * 8] - clk[p * 8 + 5]);
    printf("\n");
  }
}

int main() {
  const uint32_t len = 2047, kk = 127, H = 32;
  std::vector<uint16_t> h(H * len);
  srand(1);
  for (auto& x : h) { float f = ((rand() % 20001) - 10000) / 50.f; __half hv = __float2half(f); x = *reinterpret_cast<uint16_t*>(&hv); }
  std::vector<int32_t> hi(len); for (uint32_t i = 0; i < len; ++i) hi[i] = (int)i * 3;
  uint16_t* sc; int32_t *idx, *out;
  CK(cudaMalloc(&sc, h.size() * 2)); CK(cudaMemcpy(sc, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&idx, len * 4)); CK(cudaMemcpy(idx, hi.data(), len * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&out, H * kk * 4));
  run<16, 4>("W=16 KPC=4", sc, idx, out, len, kk);
  run<8, 8>("W=8 KPC=8", sc, idx, out, len, kk);
  run<8, 8>("W=8 KPC=8 no payload", sc, nullptr, out, len, kk);
  run<16, 4>("W=16 KPC=4 no payload", sc, nullptr, out, len, kk);
  run<32, 4>("W=32 KPC=4", sc, idx, out, len, kk);
  run<4, 8>("W=4 KPC=8", sc, idx, out, len, kk);
  return 0;
}
