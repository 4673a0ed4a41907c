// Decode-shaped Marlin GEMM kernel instantiations for (bf16, 8-bit): see marlin_decode_kernel.cuh
#include "marlin_decode_kernel.cuh"
namespace b200 {
int dispatch_decode_bf16_8(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  return dispatch_decode<__nv_bfloat16, 8>(a, A, lda, st);
}
}  // namespace b200
