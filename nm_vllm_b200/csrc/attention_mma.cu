// GQA tensor-core paged-attention decode (placeholder until the mma path lands: reports "not covered").
#include "common.cuh"
namespace b200 {
int launch_mma(float*, float*, void*, void*, const void*, const void*, const void*, int, int, int, int, float,
               const int*, const int*, int, int, int, const float*, int64_t, int64_t, int64_t, int, int, int, int,
               float, cudaStream_t) {
  return 0;
}
}  // namespace b200
