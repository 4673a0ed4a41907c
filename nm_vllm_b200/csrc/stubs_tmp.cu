// TEMPORARY stubs so the library links while the tcgen05 kernels are being brought up.
#include "common.cuh"
extern "C" int b200_gptq_marlin_24_gemm(void*, const void*, const int32_t*, const int16_t*, const void*, int32_t*, int64_t, int, int64_t, int64_t, int64_t, int64_t, int, int, void*) { return b200::set_error(B200_ENOTSUP, "gptq_marlin_24_gemm: not built yet"); }
