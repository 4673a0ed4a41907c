// Activation quantizers feeding the W8A8 GEMM.
//   static/dynamic_scaled_fp8_quant : reference csrc/quantization/fp8/common.cu:24-165
//   static/dynamic_scaled_int8_quant: reference csrc/quantization/compressed_tensors/int8_quant_kernels.cu:8-115
// Arithmetic restated (same rounding points): fp8: e4m3(clamp(float(x) * (1/scale), +-448)); dynamic scale =
// max|x|/448 accumulated with an integer atomicMax on the (non-negative) float bit pattern; int8:
// cvt.rni.sat.s8.f32(float(x)/scale) (static) or (float(x) * (127/absmax)) (dynamic per token).
// B200 notes: decode-sized inputs (<= 64 x 8192) -> a few KB..1 MB, launch-latency bound; 16-byte vector
// loads, grid sized to the SM count.
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ uint8_t to_e4m3_sat(float x, float inv_scale) {
  float r = fmaxf(-448.0f, fminf(x * inv_scale, 448.0f));
  return (uint8_t)__nv_cvt_float_to_fp8(r, __NV_SATFINITE, __NV_E4M3);
}

template <typename scalar_t>
__global__ void fp8_absmax_kernel(float* __restrict__ scale, const scalar_t* __restrict__ input, int64_t numel) {
  __shared__ float red[32];
  float m = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += stride)
    m = fmaxf(m, fabsf(to_f32<scalar_t>(input[i])));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    m = warp_max(m);
    // non-negative floats order like their bit patterns (reference atomicMaxFloat, common.cu:12-20)
    if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(scale), __float_as_int(m / 448.0f));
  }
}

template <typename scalar_t>
__global__ void fp8_quant_kernel(uint8_t* __restrict__ out, const scalar_t* __restrict__ input,
                                 const float* __restrict__ scale, int64_t numel) {
  const float inv = 1.0f / (*scale);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  constexpr int V = 16 / sizeof(scalar_t);  // elements per 16-byte load
  const bool aligned = ((reinterpret_cast<uintptr_t>(input) & 15) == 0) && ((reinterpret_cast<uintptr_t>(out) & (V - 1)) == 0);
  int64_t nvec = aligned ? numel / V : 0;
  for (int64_t i = tid; i < nvec; i += stride) {
    int4 raw = reinterpret_cast<const int4*>(input)[i];
    const scalar_t* e = reinterpret_cast<const scalar_t*>(&raw);
    uint8_t o[V];
#pragma unroll
    for (int j = 0; j < V; ++j) o[j] = to_e4m3_sat(to_f32<scalar_t>(e[j]), inv);
    if constexpr (V == 8) reinterpret_cast<uint2*>(out)[i] = *reinterpret_cast<uint2*>(o);
    else reinterpret_cast<uint32_t*>(out)[i] = *reinterpret_cast<uint32_t*>(o);
  }
  for (int64_t i = nvec * V + tid; i < numel; i += stride) out[i] = to_e4m3_sat(to_f32<scalar_t>(input[i]), inv);
}

__device__ __forceinline__ int8_t f2i8_rn_sat(float x) {
  uint32_t d;
  asm volatile("cvt.rni.sat.s8.f32 %0, %1;" : "=r"(d) : "f"(x));
  return (int8_t)(d & 0xff);
}

template <typename scalar_t>
__global__ void static_int8_quant_kernel(int8_t* __restrict__ out, const scalar_t* __restrict__ input,
                                         const float* __restrict__ scale_ptr, int hidden) {
  const float scale = *scale_ptr;
  const int64_t base = (int64_t)blockIdx.x * hidden;
  for (int i = threadIdx.x; i < hidden; i += blockDim.x)
    out[base + i] = f2i8_rn_sat(to_f32<scalar_t>(input[base + i]) / scale);
}

template <typename scalar_t>
__global__ void dynamic_int8_quant_kernel(int8_t* __restrict__ out, const scalar_t* __restrict__ input,
                                          float* __restrict__ scales, int hidden) {
  __shared__ float red[32];
  __shared__ float s_absmax;
  const int64_t base = (int64_t)blockIdx.x * hidden;
  float m = 0.f;
  for (int i = threadIdx.x; i < hidden; i += blockDim.x) m = fmaxf(m, fabsf(to_f32<scalar_t>(input[base + i])));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < ((blockDim.x + 31) >> 5) ? red[threadIdx.x] : 0.f;
    m = warp_max(m);
    if (threadIdx.x == 0) {
      s_absmax = m;
      scales[blockIdx.x] = m / 127.0f;
    }
  }
  __syncthreads();
  const float mul = 127.0f / s_absmax;
  for (int i = threadIdx.x; i < hidden; i += blockDim.x)
    out[base + i] = f2i8_rn_sat(to_f32<scalar_t>(input[base + i]) * mul);
}

static inline int grid_for(int64_t numel, int threads, int per_thread) {
  int64_t b = (numel + (int64_t)threads * per_thread - 1) / ((int64_t)threads * per_thread);
  int64_t cap = (int64_t)num_sms() * 8;
  if (b < 1) b = 1;
  return (int)(b < cap ? b : cap);
}

}  // namespace b200

extern "C" int b200_static_scaled_fp8_quant(void* out, const void* input, const float* scale, int64_t numel,
                                            int dtype, void* stream) {
  using namespace b200;
  if (numel == 0) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_DISPATCH_FLOAT(dtype, "static_scaled_fp8_quant", {
    fp8_quant_kernel<scalar_t><<<grid_for(numel, 256, 8), 256, 0, st>>>((uint8_t*)out, (const scalar_t*)input, scale, numel);
  });
  B200_LAUNCH_CHECK("static_scaled_fp8_quant");
  return B200_OK;
}

extern "C" int b200_dynamic_scaled_fp8_quant(void* out, const void* input, float* scale, int64_t numel, int dtype,
                                             void* stream) {
  using namespace b200;
  if (numel == 0) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_DISPATCH_FLOAT(dtype, "dynamic_scaled_fp8_quant", {
    fp8_absmax_kernel<scalar_t><<<grid_for(numel, 256, 8), 256, 0, st>>>(scale, (const scalar_t*)input, numel);
    fp8_quant_kernel<scalar_t><<<grid_for(numel, 256, 8), 256, 0, st>>>((uint8_t*)out, (const scalar_t*)input, scale, numel);
  });
  B200_LAUNCH_CHECK("dynamic_scaled_fp8_quant");
  return B200_OK;
}

extern "C" int b200_static_scaled_int8_quant(int8_t* out, const void* input, const float* scale, int64_t num_tokens,
                                             int hidden_size, int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0 || hidden_size == 0) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int threads = hidden_size < 1024 ? ((hidden_size + 31) / 32) * 32 : 1024;
  B200_DISPATCH_FLOAT(dtype, "static_scaled_int8_quant", {
    static_int8_quant_kernel<scalar_t><<<(unsigned)num_tokens, threads, 0, st>>>(out, (const scalar_t*)input, scale, hidden_size);
  });
  B200_LAUNCH_CHECK("static_scaled_int8_quant");
  return B200_OK;
}

extern "C" int b200_dynamic_scaled_int8_quant(int8_t* out, const void* input, float* scales, int64_t num_tokens,
                                              int hidden_size, int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0 || hidden_size == 0) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int threads = hidden_size < 1024 ? ((hidden_size + 31) / 32) * 32 : 1024;
  B200_DISPATCH_FLOAT(dtype, "dynamic_scaled_int8_quant", {
    dynamic_int8_quant_kernel<scalar_t><<<(unsigned)num_tokens, threads, 0, st>>>(out, (const scalar_t*)input, scales, hidden_size);
  });
  B200_LAUNCH_CHECK("dynamic_scaled_int8_quant");
  return B200_OK;
}
