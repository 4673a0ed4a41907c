// Decode-step glue kernels: RMSNorm (+fused residual add), SiLU*mul, rotary embedding.
// Reference: csrc/layernorm_kernels.cu:22-47, :201-243; csrc/activation_kernels.cu:14-40;
// csrc/pos_encoding_kernels.cu:10-93.  The rounding points of the reference are kept (each scalar_t operator there
// rounds to the element type): norm = T(x * rsqrt(var+eps)) * w ; residual' = T(in + res) ; silu = T(x/(1+e^-x)) * y ;
// rope = T(T(x*cos) - T(y*sin)).
// These move <= 64 x 8192 x 2 B per call in decode: one CTA per token, 8-byte vector accesses, fp32 statistics.
#include "common.cuh"

namespace b200 {

template <typename T> __device__ __forceinline__ T mul_t(T a, T b) { return from_f32<T>(to_f32<T>(a) * to_f32<T>(b)); }
template <typename T> __device__ __forceinline__ T add_t(T a, T b) { return from_f32<T>(to_f32<T>(a) + to_f32<T>(b)); }
template <typename T> __device__ __forceinline__ T sub_t(T a, T b) { return from_f32<T>(to_f32<T>(a) - to_f32<T>(b)); }

__device__ __forceinline__ float block_sum_1024(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    float r = threadIdx.x < ((blockDim.x + 31) >> 5) ? red[threadIdx.x] : 0.f;
    r = warp_sum(r);
    if (threadIdx.x == 0) red[32] = r;
  }
  __syncthreads();
  return red[32];
}

template <typename T, bool FUSED_ADD>
__global__ void rms_norm_kernel(T* __restrict__ out, T* __restrict__ input, T* __restrict__ residual,
                                const T* __restrict__ weight, float eps, int hidden) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[33];
  const int64_t base = (int64_t)blockIdx.x * hidden;
  float var = 0.f;
  for (int i = threadIdx.x; i < hidden; i += blockDim.x) {
    T z = input[base + i];
    if (FUSED_ADD) {
      z = add_t<T>(z, residual[base + i]);
      residual[base + i] = z;
    }
    const float x = to_f32<T>(z);
    var += x * x;
  }
  var = block_sum_1024(var, red);
  const float inv = rsqrtf(var / hidden + eps);
  for (int i = threadIdx.x; i < hidden; i += blockDim.x) {
    const float x = to_f32<T>(FUSED_ADD ? residual[base + i] : input[base + i]);
    out[base + i] = mul_t<T>(from_f32<T>(x * inv), weight[i]);
  }
}

template <typename T>
__global__ void silu_and_mul_kernel(T* __restrict__ out, const T* __restrict__ input, int d) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t tok = blockIdx.x;
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    const float x = to_f32<T>(input[tok * 2 * d + i]);
    const T y = input[tok * 2 * d + d + i];
    out[tok * d + i] = mul_t<T>(from_f32<T>(x / (1.0f + expf(-x))), y);
  }
}

// 16-byte vectorised variants (hidden % 8 == 0, 16-byte aligned rows): same arithmetic, element order and roundings as
// the scalar kernels above; a token row stays in registers between the two passes of the norm
template <typename T, bool FUSED_ADD, int VPT>
__global__ void __launch_bounds__(256)
rms_norm_vec_kernel(T* __restrict__ out, T* __restrict__ input, T* __restrict__ residual, const T* __restrict__ weight,
                    float eps, int hidden) {
  // the norm weights do not depend on the previous kernel of the stream: fetch them BEFORE the programmatic-dependency wait,
  // so their L2 latency (and, first time, HBM latency) overlaps the predecessor's tail instead of sitting behind the reduction
  uint4 wv[VPT];
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = (threadIdx.x + j * 256) * 8;
    if (i < hidden) wv[j] = *reinterpret_cast<const uint4*>(weight + i);
  }
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
      if (FUSED_ADD) {
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
  var = block_sum_1024(var, red);
  const float inv = rsqrtf(var / hidden + eps);
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int i = (threadIdx.x + j * 256) * 8;
    if (i < hidden) {
      const T* ww = reinterpret_cast<const T*>(&wv[j]);
      T* z = reinterpret_cast<T*>(&v[j]);
#pragma unroll
      for (int e = 0; e < 8; ++e) z[e] = mul_t<T>(from_f32<T>(to_f32<T>(z[e]) * inv), ww[e]);
      *reinterpret_cast<uint4*>(out + base + i) = v[j];
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
silu_and_mul_vec_kernel(T* __restrict__ out, const T* __restrict__ input, int d) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t tok = blockIdx.y;
  const int i = (blockIdx.x * 256 + threadIdx.x) * 8;
  if (i >= d) return;
  uint4 a = *reinterpret_cast<const uint4*>(input + tok * 2 * d + i);
  const uint4 b = *reinterpret_cast<const uint4*>(input + tok * 2 * d + d + i);
  T* x = reinterpret_cast<T*>(&a);
  const T* y = reinterpret_cast<const T*>(&b);
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float f = to_f32<T>(x[e]);
    x[e] = mul_t<T>(from_f32<T>(f / (1.0f + expf(-f))), y[e]);
  }
  *reinterpret_cast<uint4*>(out + tok * d + i) = a;
}

template <typename T, bool NEOX>
__device__ __forceinline__ void rope_one(T* arr, const T* cos_ptr, const T* sin_ptr, int rot_offset, int embed_dim) {
  int xi, yi;
  T c, s;
  if (NEOX) { xi = rot_offset; yi = embed_dim + rot_offset; c = cos_ptr[xi]; s = sin_ptr[xi]; }
  else { xi = 2 * rot_offset; yi = xi + 1; c = cos_ptr[rot_offset]; s = sin_ptr[rot_offset]; }
  const T x = arr[xi], y = arr[yi];
  arr[xi] = sub_t<T>(mul_t<T>(x, c), mul_t<T>(y, s));
  arr[yi] = add_t<T>(mul_t<T>(y, c), mul_t<T>(x, s));
}

template <typename T, bool NEOX>
__global__ void rotary_kernel(const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
                              const T* __restrict__ cos_sin_cache, int rot_dim, int64_t q_stride, int64_t k_stride,
                              int num_heads, int num_kv_heads, int head_size) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t tok = blockIdx.x;
  const T* cache = cos_sin_cache + positions[tok] * rot_dim;
  const int embed = rot_dim / 2;
  for (int i = threadIdx.x; i < num_heads * embed; i += blockDim.x)
    rope_one<T, NEOX>(query + tok * q_stride + (int64_t)(i / embed) * head_size, cache, cache + embed, i % embed, embed);
  for (int i = threadIdx.x; i < num_kv_heads * embed; i += blockDim.x)
    rope_one<T, NEOX>(key + tok * k_stride + (int64_t)(i / embed) * head_size, cache, cache + embed, i % embed, embed);
}

static inline int tok_threads(int n, int cap) {
  int t = ((n + 31) / 32) * 32;
  return t < cap ? (t < 32 ? 32 : t) : cap;
}
}  // namespace b200


namespace b200 {
static inline bool vec_ok(int dtype, int n, const void* a, const void* b, const void* c) {
  return dtype != B200_F32 && n % 8 == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) |
                                               reinterpret_cast<uintptr_t>(c)) & 15) == 0;
}
template <typename T, bool FUSED>
static void launch_rms_vec(T* out, T* in, T* res, const T* w, float eps, int hidden, int64_t tokens, cudaStream_t st) {
  if (hidden <= 2048) launch_pdl(rms_norm_vec_kernel<T, FUSED, 1>, dim3((unsigned)tokens), dim3(256), 0, st, out, in, res, w, eps, hidden);
  else if (hidden <= 4096) launch_pdl(rms_norm_vec_kernel<T, FUSED, 2>, dim3((unsigned)tokens), dim3(256), 0, st, out, in, res, w, eps, hidden);
  else launch_pdl(rms_norm_vec_kernel<T, FUSED, 4>, dim3((unsigned)tokens), dim3(256), 0, st, out, in, res, w, eps, hidden);
}
}  // namespace b200

extern "C" int b200_rms_norm(void* out, const void* input, const void* weight, int64_t num_tokens, int hidden_size,
                             float epsilon, int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  if (vec_ok(dtype, hidden_size, out, input, weight) && hidden_size <= 8192) {
    if (dtype == B200_F16) launch_rms_vec<__half, false>((__half*)out, (__half*)input, (__half*)input, (const __half*)weight, epsilon, hidden_size, num_tokens, (cudaStream_t)stream);
    else launch_rms_vec<__nv_bfloat16, false>((__nv_bfloat16*)out, (__nv_bfloat16*)input, (__nv_bfloat16*)input, (const __nv_bfloat16*)weight, epsilon, hidden_size, num_tokens, (cudaStream_t)stream);
    B200_LAUNCH_CHECK("rms_norm");
    return B200_OK;
  }
  B200_DISPATCH_FLOAT(dtype, "rms_norm", {
    launch_pdl(rms_norm_kernel<scalar_t, false>, dim3((unsigned)num_tokens), dim3(tok_threads(hidden_size, 1024)), 0, (cudaStream_t)stream, 
        (scalar_t*)out, (scalar_t*)input, nullptr, (const scalar_t*)weight, epsilon, hidden_size);
  });
  B200_LAUNCH_CHECK("rms_norm");
  return B200_OK;
}

extern "C" int b200_fused_add_rms_norm(void* input, void* residual, const void* weight, int64_t num_tokens,
                                       int hidden_size, float epsilon, int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  if (vec_ok(dtype, hidden_size, input, residual, weight) && hidden_size <= 8192) {
    if (dtype == B200_F16) launch_rms_vec<__half, true>((__half*)input, (__half*)input, (__half*)residual, (const __half*)weight, epsilon, hidden_size, num_tokens, (cudaStream_t)stream);
    else launch_rms_vec<__nv_bfloat16, true>((__nv_bfloat16*)input, (__nv_bfloat16*)input, (__nv_bfloat16*)residual, (const __nv_bfloat16*)weight, epsilon, hidden_size, num_tokens, (cudaStream_t)stream);
    B200_LAUNCH_CHECK("fused_add_rms_norm");
    return B200_OK;
  }
  B200_DISPATCH_FLOAT(dtype, "fused_add_rms_norm", {
    launch_pdl(rms_norm_kernel<scalar_t, true>, dim3((unsigned)num_tokens), dim3(tok_threads(hidden_size, 1024)), 0, (cudaStream_t)stream, 
        (scalar_t*)input, (scalar_t*)input, (scalar_t*)residual, (const scalar_t*)weight, epsilon, hidden_size);
  });
  B200_LAUNCH_CHECK("fused_add_rms_norm");
  return B200_OK;
}

extern "C" int b200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d, int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  if (vec_ok(dtype, d, out, input, input)) {
    const dim3 grid((unsigned)((d + 2047) / 2048), (unsigned)num_tokens);
    if (dtype == B200_F16) launch_pdl(silu_and_mul_vec_kernel<__half>, grid, dim3(256), 0, (cudaStream_t)stream, (__half*)out, (const __half*)input, d);
    else launch_pdl(silu_and_mul_vec_kernel<__nv_bfloat16>, grid, dim3(256), 0, (cudaStream_t)stream, (__nv_bfloat16*)out, (const __nv_bfloat16*)input, d);
    B200_LAUNCH_CHECK("silu_and_mul");
    return B200_OK;
  }
  B200_DISPATCH_FLOAT(dtype, "silu_and_mul", {
    launch_pdl(silu_and_mul_kernel<scalar_t>, dim3((unsigned)num_tokens), dim3(tok_threads(d, 1024)), 0, (cudaStream_t)stream, 
        (scalar_t*)out, (const scalar_t*)input, d);
  });
  B200_LAUNCH_CHECK("silu_and_mul");
  return B200_OK;
}

extern "C" int b200_rotary_embedding(const int64_t* positions, void* query, void* key, int head_size,
                                     const void* cos_sin_cache, int rot_dim, int is_neox, int64_t num_tokens,
                                     int num_heads, int num_kv_heads, int64_t query_stride, int64_t key_stride,
                                     int dtype, void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  const int threads = tok_threads(num_heads * rot_dim / 2, 512);
  B200_DISPATCH_FLOAT(dtype, "rotary_embedding", {
    if (is_neox)
      launch_pdl(rotary_kernel<scalar_t, true>, dim3((unsigned)num_tokens), dim3(threads), 0, (cudaStream_t)stream, 
          positions, (scalar_t*)query, (scalar_t*)key, (const scalar_t*)cos_sin_cache, rot_dim, query_stride,
          key_stride, num_heads, num_kv_heads, head_size);
    else
      launch_pdl(rotary_kernel<scalar_t, false>, dim3((unsigned)num_tokens), dim3(threads), 0, (cudaStream_t)stream, 
          positions, (scalar_t*)query, (scalar_t*)key, (const scalar_t*)cos_sin_cache, rot_dim, query_stride,
          key_stride, num_heads, num_kv_heads, head_size);
  });
  B200_LAUNCH_CHECK("rotary_embedding");
  return B200_OK;
}
