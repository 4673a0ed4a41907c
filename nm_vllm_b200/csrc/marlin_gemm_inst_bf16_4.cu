// Marlin GEMM kernel instantiations for (bf16, 4-bit): see marlin_gemm_kernel.cuh
#include "marlin_gemm_kernel.cuh"
namespace b200 {
int dispatch_gemm_bf16_4(const GemmArgs& a, const void* A, int64_t lda, cudaStream_t st) {
  return dispatch_gemm<__nv_bfloat16, 4>(a, A, lda, st);
}
}  // namespace b200
