// Shared helpers for the sm_100a kernels: error plumbing for the C ABI, scalar conversions, fp8 codecs.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/b200_ops.h"

namespace b200 {

// ---- error plumbing ---------------------------------------------------------------------------------
int set_error(int code, const char* fmt, ...);  // defined in capi.cu; returns `code`

#define B200_CHECK(cond, ...)                                   \
  do {                                                          \
    if (!(cond)) return ::b200::set_error(B200_EINVAL, __VA_ARGS__); \
  } while (0)

#define B200_LAUNCH_CHECK(name)                                                      \
  do {                                                                               \
    cudaError_t e__ = cudaGetLastError();                                            \
    if (e__ != cudaSuccess)                                                          \
      return ::b200::set_error(B200_ECUDA, "%s: %s", name, cudaGetErrorString(e__)); \
  } while (0)

int num_sms();  // cached SM count of the current device

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------------------------
// Every kernel of the decode step is launched with the programmatic-stream-serialization attribute: it may start
// while its predecessor is still running, calls pdl_wait() before touching anything the predecessor could still
// read or write, and calls pdl_launch_dependents() as early as possible so that ITS successor can do the same.  Rule:
// a kernel launched through launch_pdl() MUST execute pdl_wait() on every path that reaches global memory shared
// with other kernels (the wait also orders it after the predecessor's predecessor).  B200_NO_PDL=1 disables it.
bool pdl_enabled();  // defined in capi.cu
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// ---- scalar conversions -----------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// ---- fp8 codecs (semantics of csrc/quantization/fp8/nvidia/quant_utils.cuh:287-497) -----------------
// quant:   fp8 = cvt_satfinite(float(x) / scale)
// dequant: x   = T(float(half(fp8)) * scale)
template <int KV> struct Fp8Kind;
template <> struct Fp8Kind<B200_KV_FP8_E4M3> { static constexpr __nv_fp8_interpretation_t v = __NV_E4M3; };
template <> struct Fp8Kind<B200_KV_FP8_E5M2> { static constexpr __nv_fp8_interpretation_t v = __NV_E5M2; };

template <int KV>
__device__ __forceinline__ uint8_t fp8_quant(float x, float scale) {
  return (uint8_t)__nv_cvt_float_to_fp8(x / scale, __NV_SATFINITE, Fp8Kind<KV>::v);
}
template <int KV>
__device__ __forceinline__ float fp8_dequant(uint8_t v, float scale) {
  __half_raw h = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)v, Fp8Kind<KV>::v);
  return __half2float(*reinterpret_cast<__half*>(&h)) * scale;
}

// ---- warp helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// dtype dispatch for host launchers: calls FN<T>() with the element type
#define B200_DISPATCH_FLOAT(dtype, NAME, ...)                                        \
  switch (dtype) {                                                                   \
    case B200_F16: { using scalar_t = __half; __VA_ARGS__; break; }                  \
    case B200_BF16: { using scalar_t = __nv_bfloat16; __VA_ARGS__; break; }          \
    case B200_F32: { using scalar_t = float; __VA_ARGS__; break; }                   \
    default: return ::b200::set_error(B200_EINVAL, "%s: bad dtype %d", NAME, dtype); \
  }

}  // namespace b200
