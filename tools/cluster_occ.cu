// Bring-up helper: how many clusters of size S (1 CTA/SM-sized kernels) can be co-resident on this GPU?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(896, 1) big(int* p) { extern __shared__ int s[]; if (p) p[0] = s[0]; }
__global__ void __launch_bounds__(256, 1) small(int* p) { extern __shared__ int s[]; if (p) p[0] = s[0]; }
template <typename K> void probe(K kern, int threads, int smem, const char* name) {
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  for (int s = 1; s <= 16; ++s) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(s * 64); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = s; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = -1;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
    printf("%s smem=%d cluster=%2d: max active clusters %d (%d CTAs) %s\n", name, smem, s, n, n * s, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("%s SMs=%d\n", p.name, p.multiProcessorCount);
  probe(big, 896, 215 * 1024, "big");
  probe(small, 256, 100 * 1024, "half");
  return 0;
}
