// Decode-shaped Marlin GEMM kernel instantiations for (f16, 4-bit): see marlin_decode_kernel.cuh
#include "marlin_decode_kernel.cuh"
namespace b200 {
int dispatch_decode_f16_4(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  return dispatch_decode<__half, 4>(a, A, lda, st);
}
}  // namespace b200
