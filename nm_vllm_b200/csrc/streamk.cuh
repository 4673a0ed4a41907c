// Stream-K work partition + per-device split-K scratch shared by the tcgen05 GEMM kernels (Marlin W4A16, W8A8 scaled_mm).
#pragma once
#include "common.cuh"

namespace b200 {

// `total` scheduling units dealt to `g` CTAs as contiguous ranges whose sizes differ by at most one
struct Partition {
  int q, r;
  __host__ __device__ __forceinline__ Partition(int total, int g) : q(total / g), r(total % g) {}
  __host__ __device__ __forceinline__ int start(int b) const { return b * q + (b < r ? b : r); }
  __host__ __device__ __forceinline__ int owner(int u) const {
    const int big = r * (q + 1);
    return u < big ? u / (q + 1) : r + (u - big) / q;
  }
};

struct GemmScratch {
  float* partials;
  int* counters;
};
// grows (outside CUDA-graph capture only) the per-device scratch; defined in marlin_gemm.cu
int gemm_ensure_scratch(int dev, size_t partial_floats, int counter_ints, cudaStream_t st, GemmScratch* out);

}  // namespace b200
