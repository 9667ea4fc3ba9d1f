// How many thread-block clusters of a given shape can be co-resident on this GPU (B200 GPC packing).
// nvcc -gencode arch=compute_100a,code=sm_100a -o tools/cluster_probe tools/cluster_probe.cu
#include <cuda_runtime.h>
#include <stdio.h>
__global__ void __launch_bounds__(512, 2) k512(int* p) { if (p && threadIdx.x == 9999) *p = 1; }
__global__ void __launch_bounds__(1024, 1) k1024(int* p) { if (p && threadIdx.x == 9999) *p = 1; }
template <typename K>
static void probe(const char* name, K kern, int threads, int smem) {
  for (int cs : {1, 2, 4, 8, 16}) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(cs * 64);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs, at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
    cfg.attrs = at, cfg.numAttrs = 1;
    if (cs > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int n = -1;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
    printf("%s threads=%d smem=%d cluster=%d -> max active clusters %d (%d CTAs) %s\n", name, threads, smem, cs, n, n * cs,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
}
int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  printf("%s SMs=%d\n", p.name, p.multiProcessorCount);
  probe("k512", k512, 512, 48 * 1024);
  probe("k512", k512, 512, 100 * 1024);
  probe("k1024", k1024, 1024, 100 * 1024);
  return 0;
}
