// Decode-step glue fused with the activation quantiser of the W8A8 linear that follows (second half):
//   rms_norm / fused_add_rms_norm  ->  fp8 (static per-tensor scale) | int8 (static per-tensor | dynamic per-token)
//   silu_and_mul                   ->  the same three
// Reference ops being fused: csrc/layernorm_kernels.cu:22-352, csrc/activation_kernels.cu:14-162 with
// csrc/quantization/fp8/common.cu:24-125 (static_scaled_fp8_quant) and
// csrc/quantization/compressed_tensors/int8_quant_kernels.cu:30-71 (static / dynamic per-token int8).  The rounding points of the
// unfused chain are kept: the 16-bit norm / activation result is formed first (glue.cu), then quantised exactly like quant.cu
// does -- so the fused output is bit-identical to norm -> quant.  The per-TENSOR dynamic fp8 scale of Fp8LinearMethod's
// "dynamic" scheme needs a grid-wide maximum and is not fused (it stays two kernels); static scales (fp8 checkpoints with
// activation_scheme "static") and the per-token int8 scheme are local to a token row.  One CTA per token.
#include "common.cuh"

namespace b200 {
namespace gq {

enum { kFp8Static = 1, kInt8Static = 2, kInt8Token = 3, kFp8Dynamic = 4 };
constexpr int64_t kDynMaxTokens = 512;   // kFp8Dynamic: every CTA of the grid must be co-resident (grid-wide maximum in one launch)

// kFp8Dynamic (Fp8LinearMethod activation_scheme "dynamic": ONE per-tensor scale = max|x| / 448, fp8/common.cu:39-72): the
// grid-wide maximum is taken inside the producing kernel -- every CTA holds its token row, publishes its row maximum with
// an integer atomicMax, meets the others at a counter, then quantises with the common scale.  No allocation, no pre-zeroed
// scale tensor: the three words below live for the life of the library (one set per GPU), are left zero by the last CTA of
// every call, and are shared by every call -- calls that overlap on one GPU would race (decode steps run on one stream).
// Bit-identical to norm / silu_and_mul -> dynamic_scaled_fp8_quant (same max|x| / 448 and 1 / scale).
static __device__ unsigned int g_dyn_ws[4];   // [0] max bits of (row max / 448), [1] arrivals, [2] departures

__device__ __forceinline__ float dyn_scale_exchange(float row_amax, float* __restrict__ scale_out, unsigned n_ctas) {
  __shared__ float s_scale;
  if (threadIdx.x == 0) {
    atomicMax(reinterpret_cast<int*>(&g_dyn_ws[0]), __float_as_int(row_amax / 448.0f));
    __threadfence();
    atomicAdd(&g_dyn_ws[1], 1u);
    while (*reinterpret_cast<volatile unsigned int*>(&g_dyn_ws[1]) < n_ctas) {}
    __threadfence();
    const float sc = __int_as_float(*reinterpret_cast<volatile int*>(&g_dyn_ws[0]));
    s_scale = sc;
    if (blockIdx.x == 0 && blockIdx.y == 0) *scale_out = sc;
    if (atomicAdd(&g_dyn_ws[2], 1u) == n_ctas - 1) {   // everyone has read the maximum: leave the words zero for the next call
      g_dyn_ws[0] = 0u;
      g_dyn_ws[2] = 0u;
      __threadfence();
      g_dyn_ws[1] = 0u;
    }
  }
  __syncthreads();
  return s_scale;
}

template <typename T> __device__ __forceinline__ T mul_t(T a, T b) { return from_f32<T>(to_f32<T>(a) * to_f32<T>(b)); }
template <typename T> __device__ __forceinline__ T add_t(T a, T b) { return from_f32<T>(to_f32<T>(a) + to_f32<T>(b)); }

__device__ __forceinline__ uint8_t q_fp8(float x, float inv_scale) {
  const float r = fmaxf(-448.0f, fminf(x * inv_scale, 448.0f));
  return (uint8_t)__nv_cvt_float_to_fp8(r, __NV_SATFINITE, __NV_E4M3);
}
__device__ __forceinline__ uint8_t q_i8(float x) {
  uint32_t d;
  asm volatile("cvt.rni.sat.s8.f32 %0, %1;" : "=r"(d) : "f"(x));
  return (uint8_t)(d & 0xff);
}

__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max) {
  v = is_max ? warp_max(v) : warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    float r = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    r = is_max ? warp_max(r) : warp_sum(r);
    if (threadIdx.x == 0) red[32] = r;
  }
  __syncthreads();
  const float out = red[32];
  __syncthreads();
  return out;
}

// 8 values of T (one 16-byte vector) -> 8 quantised bytes
template <typename T, int QK>
__device__ __forceinline__ uint2 quant8(const T* z, float qmul) {
  uint8_t o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float x = to_f32<T>(z[e]);
    o[e] = (QK == kFp8Static || QK == kFp8Dynamic) ? q_fp8(x, qmul) : (QK == kInt8Static ? q_i8(x / qmul) : q_i8(x * qmul));
  }
  return *reinterpret_cast<uint2*>(o);
}

// hidden <= 8192, hidden % 8 == 0, 256 threads: a token row stays in registers (VPT vectors of 8 per thread)
template <typename T, int QK, bool ADD, int VPT>
__global__ void __launch_bounds__(256)
rms_norm_quant_kernel(uint8_t* __restrict__ out_q, float* __restrict__ scales, const T* __restrict__ input,
                      T* __restrict__ residual, const T* __restrict__ weight, float eps, int hidden) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[33];
  const int64_t base = (int64_t)blockIdx.x * hidden;
  uint4 v[VPT];
  float var = 0.f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = (threadIdx.x + j * 256) * 8;
    if (i < hidden) {
      v[j] = *reinterpret_cast<const uint4*>(input + base + i);
      T* z = reinterpret_cast<T*>(&v[j]);
      if (ADD) {
        const uint4 r = *reinterpret_cast<const uint4*>(residual + base + i);
        const T* rr = reinterpret_cast<const T*>(&r);
#pragma unroll
        for (int e = 0; e < 8; ++e) z[e] = add_t<T>(z[e], rr[e]);
        *reinterpret_cast<uint4*>(residual + base + i) = v[j];
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float x = to_f32<T>(z[e]); var += x * x; }
    }
  }
  var = block_reduce(var, red, false);
  const float inv = rsqrtf(var / hidden + eps);
  float amax = 0.f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = (threadIdx.x + j * 256) * 8;
    if (i < hidden) {
      const uint4 w = *reinterpret_cast<const uint4*>(weight + i);
      const T* ww = reinterpret_cast<const T*>(&w);
      T* z = reinterpret_cast<T*>(&v[j]);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        z[e] = mul_t<T>(from_f32<T>(to_f32<T>(z[e]) * inv), ww[e]);
        if (QK == kInt8Token || QK == kFp8Dynamic) amax = fmaxf(amax, fabsf(to_f32<T>(z[e])));
      }
    }
  }
  float qmul;
  if (QK == kInt8Token) {
    amax = block_reduce(amax, red, true);
    if (threadIdx.x == 0) scales[blockIdx.x] = amax / 127.0f;
    qmul = 127.0f / amax;
  } else if (QK == kFp8Dynamic) {
    amax = block_reduce(amax, red, true);
    qmul = 1.0f / dyn_scale_exchange(amax, scales, gridDim.x);
  } else {
    qmul = QK == kFp8Static ? 1.0f / (*scales) : *scales;
  }
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = (threadIdx.x + j * 256) * 8;
    if (i < hidden) *reinterpret_cast<uint2*>(out_q + base + i) = quant8<T, QK>(reinterpret_cast<const T*>(&v[j]), qmul);
  }
}

// silu(x[:d]) * x[d:] -> quantised [T, d]; any d % 8 == 0.  Per-token int8 needs the row maximum first: the activation is
// computed twice (cheaper than staging up to 57 KB per row)
template <typename T, int QK>
__global__ void __launch_bounds__(256)
silu_and_mul_quant_kernel(uint8_t* __restrict__ out_q, float* __restrict__ scales, const T* __restrict__ input, int d) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[33];
  const int64_t tok = blockIdx.x;
  const T* xr = input + tok * 2 * d;
  auto act8 = [&](int i, T* z) {
    const uint4 a = *reinterpret_cast<const uint4*>(xr + i);
    const uint4 b = *reinterpret_cast<const uint4*>(xr + d + i);
    const T* x = reinterpret_cast<const T*>(&a);
    const T* y = reinterpret_cast<const T*>(&b);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float f = to_f32<T>(x[e]);
      z[e] = mul_t<T>(from_f32<T>(f / (1.0f + expf(-f))), y[e]);
    }
  };
  float qmul;
  if (QK == kInt8Token || QK == kFp8Dynamic) {
    float amax = 0.f;
    for (int i = threadIdx.x * 8; i < d; i += 256 * 8) {
      T z[8];
      act8(i, z);
#pragma unroll
      for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(to_f32<T>(z[e])));
    }
    amax = block_reduce(amax, red, true);
    if (QK == kFp8Dynamic) {
      qmul = 1.0f / dyn_scale_exchange(amax, scales, gridDim.x);
    } else {
      if (threadIdx.x == 0) scales[tok] = amax / 127.0f;
      qmul = 127.0f / amax;
    }
  } else {
    qmul = QK == kFp8Static ? 1.0f / (*scales) : *scales;
  }
  // static scales: the row is cut over gridDim.y CTAs (no row-wide reduction needed)
  for (int i = (blockIdx.y * 256 + threadIdx.x) * 8; i < d; i += gridDim.y * 256 * 8) {
    T z[8];
    act8(i, z);
    *reinterpret_cast<uint2*>(out_q + tok * d + i) = quant8<T, QK>(z, qmul);
  }
}

template <typename T, int QK, bool ADD>
static int launch_norm(uint8_t* out_q, float* scales, const T* in, T* res, const T* w, float eps, int hidden, int64_t tokens,
                       cudaStream_t st) {
  cudaError_t e;
  if (hidden <= 2048) e = launch_pdl(rms_norm_quant_kernel<T, QK, ADD, 1>, dim3((unsigned)tokens), dim3(256), 0, st, out_q, scales, in, res, w, eps, hidden);
  else if (hidden <= 4096) e = launch_pdl(rms_norm_quant_kernel<T, QK, ADD, 2>, dim3((unsigned)tokens), dim3(256), 0, st, out_q, scales, in, res, w, eps, hidden);
  else e = launch_pdl(rms_norm_quant_kernel<T, QK, ADD, 4>, dim3((unsigned)tokens), dim3(256), 0, st, out_q, scales, in, res, w, eps, hidden);
  return e == cudaSuccess ? B200_OK : set_error(B200_ECUDA, "rms_norm_quant launch: %s", cudaGetErrorString(e));
}

}  // namespace gq
}  // namespace b200

// qkind: 1 = fp8 e4m3 with the static per-tensor scale *scales; 2 = int8 with the static per-tensor scale *scales; 3 = int8 with
// dynamic per-token scales written to scales[num_tokens]; 4 = fp8 e4m3 with the DYNAMIC per-tensor scale max|x| / 448 computed
// in the same launch and written to *scales (num_tokens <= 512, B200_ENOTSUP beyond).  residual == NULL: rms_norm; otherwise fused_add_rms_norm (residual
// updated in place, `input` left untouched).
extern "C" int b200_rms_norm_quant(void* out_q, float* scales, const void* input, void* residual, const void* weight,
                                   int64_t num_tokens, int hidden_size, float epsilon, int dtype, int qkind, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(dtype == B200_F16 || dtype == B200_BF16, "rms_norm_quant: fp16 / bf16 activations only");
  B200_CHECK(qkind >= 1 && qkind <= 4, "rms_norm_quant: qkind %d", qkind);
  if (qkind == 4 && num_tokens > gq::kDynMaxTokens)
    return set_error(B200_ENOTSUP, "rms_norm_quant: the fused dynamic per-tensor fp8 scale needs all %lld token CTAs co-resident (limit %lld): "
                     "run the norm and dynamic_scaled_fp8_quant separately", (long long)num_tokens, (long long)gq::kDynMaxTokens);
  B200_CHECK(hidden_size % 8 == 0 && hidden_size <= 8192, "rms_norm_quant: hidden_size %d must be a multiple of 8 and <= 8192", hidden_size);
  B200_CHECK(((reinterpret_cast<uintptr_t>(input) | reinterpret_cast<uintptr_t>(residual) | reinterpret_cast<uintptr_t>(weight)) & 15) == 0 &&
             (reinterpret_cast<uintptr_t>(out_q) & 7) == 0, "rms_norm_quant: 16-byte aligned rows required");
  cudaStream_t st = (cudaStream_t)stream;
  int rc = B200_OK;
#define GQ_NORM(TT)                                                                                                            \
  do {                                                                                                                         \
    uint8_t* o = (uint8_t*)out_q; const TT* in = (const TT*)input; TT* rs = (TT*)residual; const TT* w = (const TT*)weight;    \
    if (residual) {                                                                                                            \
      if (qkind == 1) rc = gq::launch_norm<TT, 1, true>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);           \
      else if (qkind == 2) rc = gq::launch_norm<TT, 2, true>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);      \
      else if (qkind == 4) rc = gq::launch_norm<TT, 4, true>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);      \
      else rc = gq::launch_norm<TT, 3, true>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);                      \
    } else {                                                                                                                   \
      if (qkind == 1) rc = gq::launch_norm<TT, 1, false>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);          \
      else if (qkind == 2) rc = gq::launch_norm<TT, 2, false>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);     \
      else if (qkind == 4) rc = gq::launch_norm<TT, 4, false>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);     \
      else rc = gq::launch_norm<TT, 3, false>(o, scales, in, rs, w, epsilon, hidden_size, num_tokens, st);                     \
    }                                                                                                                          \
  } while (0)
  if (dtype == B200_F16) GQ_NORM(__half); else GQ_NORM(__nv_bfloat16);
#undef GQ_NORM
  if (rc) return rc;
  B200_LAUNCH_CHECK("rms_norm_quant");
  return B200_OK;
}

extern "C" int b200_silu_and_mul_quant(void* out_q, float* scales, const void* input, int64_t num_tokens, int d, int dtype,
                                       int qkind, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(dtype == B200_F16 || dtype == B200_BF16, "silu_and_mul_quant: fp16 / bf16 activations only");
  B200_CHECK(qkind >= 1 && qkind <= 4, "silu_and_mul_quant: qkind %d", qkind);
  if (qkind == 4 && num_tokens > gq::kDynMaxTokens)
    return set_error(B200_ENOTSUP, "silu_and_mul_quant: the fused dynamic per-tensor fp8 scale needs all %lld token CTAs co-resident (limit %lld)",
                     (long long)num_tokens, (long long)gq::kDynMaxTokens);
  B200_CHECK(d % 8 == 0 && (reinterpret_cast<uintptr_t>(input) & 15) == 0 && (reinterpret_cast<uintptr_t>(out_q) & 7) == 0,
             "silu_and_mul_quant: d %% 8 == 0 and aligned rows required");
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaSuccess;
  const unsigned chunks = (qkind == 3 || qkind == 4) ? 1u : (unsigned)((d + 2047) / 2048);
#define GQ_SILU(TT, QQ) e = launch_pdl(gq::silu_and_mul_quant_kernel<TT, QQ>, dim3((unsigned)num_tokens, chunks), dim3(256), 0, st, (uint8_t*)out_q, scales, (const TT*)input, d)
  if (dtype == B200_F16) { if (qkind == 1) GQ_SILU(__half, 1); else if (qkind == 2) GQ_SILU(__half, 2); else if (qkind == 4) GQ_SILU(__half, 4); else GQ_SILU(__half, 3); }
  else { if (qkind == 1) GQ_SILU(__nv_bfloat16, 1); else if (qkind == 2) GQ_SILU(__nv_bfloat16, 2); else if (qkind == 4) GQ_SILU(__nv_bfloat16, 4); else GQ_SILU(__nv_bfloat16, 3); }
#undef GQ_SILU
  if (e != cudaSuccess) return set_error(B200_ECUDA, "silu_and_mul_quant launch: %s", cudaGetErrorString(e));
  B200_LAUNCH_CHECK("silu_and_mul_quant");
  return B200_OK;
}
