// Greedy sampling (argmax over the vocabulary) for the decode step: out[t] = argmax_v logits[t, v], lowest index on ties
// (what torch.argmax's CUDA reduction does; the reference's Sampler takes this path for temperature 0,
// vllm/model_executor/layers/sampler.py _greedy_sample).  The logits were just written by the lm_head GEMM and sit in L2.
// One cluster of 8 CTAs per row: each CTA scans 1/8 of the vocabulary with 16-byte loads, the partial (value, index)
// pairs meet in distributed shared memory -- one launch, no scratch, no atomics.  NaNs are not treated as maxima.
#include "common.cuh"
#include "ptx.cuh"

namespace b200 {
using namespace ptx;

constexpr int kAmCluster = 8;
constexpr int kAmThreads = 256;

__device__ __forceinline__ void am_take(float& bv, int& bi, float v, int i) {
  if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
}

template <typename T>
__global__ void __cluster_dims__(kAmCluster, 1, 1) __launch_bounds__(kAmThreads)
greedy_argmax_kernel(int64_t* __restrict__ out, const T* __restrict__ logits, const int V, const int64_t stride) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float s_val[kAmThreads / 32 + 1];
  __shared__ int s_idx[kAmThreads / 32 + 1];
  const int row = blockIdx.y, r = blockIdx.x;   // r == rank in the cluster (cluster spans x)
  const T* p = logits + (int64_t)row * stride;
  constexpr int VEC = 16 / sizeof(T);
  const int chunk = ((V + kAmCluster - 1) / kAmCluster + VEC - 1) / VEC * VEC;
  const int lo = r * chunk, hi = min(V, lo + chunk);
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(p) & 15) == 0);
  int i = lo + threadIdx.x * VEC;
  if (vec_ok) {
    for (; i + VEC <= hi; i += kAmThreads * VEC) {
      const uint4 w = *reinterpret_cast<const uint4*>(p + i);
      const T* e = reinterpret_cast<const T*>(&w);
#pragma unroll
      for (int k = 0; k < VEC; ++k) am_take(bv, bi, to_f32<T>(e[k]), i + k);
    }
  }
  // leftovers (unaligned rows, or the last partial vector of the range): the same stride pattern, element-wise
  for (; i < hi; i += kAmThreads * VEC)
    for (int k = 0; k < VEC && i + k < hi; ++k) am_take(bv, bi, to_f32<T>(p[i + k]), i + k);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float v2 = __shfl_xor_sync(0xffffffffu, bv, o);
    const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
    am_take(bv, bi, v2, i2);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kAmThreads / 32; ++w) am_take(bv, bi, s_val[w], s_idx[w]);
    s_val[kAmThreads / 32] = bv;
    s_idx[kAmThreads / 32] = bi;
  }
  cluster_arrive();
  cluster_wait();
  if (r == 0 && threadIdx.x == 0) {
    const uint32_t av = smem_u32(&s_val[kAmThreads / 32]), ai = smem_u32(&s_idx[kAmThreads / 32]);
    for (int c = 1; c < kAmCluster; ++c) {
      const float v2 = ld_dsmem_f32(mapa(av, (uint32_t)c));
      const int i2 = __float_as_int(ld_dsmem_f32(mapa(ai, (uint32_t)c)));
      am_take(bv, bi, v2, i2);
    }
    out[row] = bi == 0x7fffffff ? 0 : bi;
  }
  // keep every CTA's shared memory alive until rank 0 has read it
  cluster_arrive();
  cluster_wait();
}

}  // namespace b200

using namespace b200;

extern "C" int b200_greedy_sample(int64_t* out, const void* logits, int64_t num_tokens, int64_t vocab, int64_t row_stride,
                                  int dtype, void* stream) {
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(vocab > 0 && vocab < (int64_t)0x7fffffff, "greedy_sample: bad vocabulary size %lld", (long long)vocab);
  B200_CHECK(num_tokens <= 65535, "greedy_sample: too many rows");
  dim3 grid(kAmCluster, (unsigned)num_tokens);
  cudaError_t e = cudaSuccess;
  B200_DISPATCH_FLOAT(dtype, "greedy_sample", {
    e = launch_pdl(greedy_argmax_kernel<scalar_t>, grid, dim3(kAmThreads), 0, (cudaStream_t)stream, out,
                   (const scalar_t*)logits, (int)vocab, row_stride);
  });
  if (e != cudaSuccess) return set_error(B200_ECUDA, "greedy_sample: %s", cudaGetErrorString(e));
  return B200_OK;
}
