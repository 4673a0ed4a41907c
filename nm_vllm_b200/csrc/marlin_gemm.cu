// C ABI of the Marlin-layout W4A16 / W8A16 GEMM (kernel: marlin_gemm_kernel.cuh; instantiations: marlin_gemm_inst_*.cu).
// Reference: csrc/quantization/gptq_marlin/gptq_marlin.cu:1735-1868 (op), csrc/ops.h:75-77, :86-90.
#include "marlin_decode_kernel.cuh"

namespace b200 {

// act-order: the reference permutes the activation columns into a temporary before the GEMM (permute_cols_kernel,
// gptq_marlin.cu:283-339); the repacked weight rows are sorted by group, so with is_k_full the groups are again
// uniform blocks of group_size rows and the regular kernel applies.
template <typename T>
__global__ void permute_cols_kernel(T* __restrict__ out, const T* __restrict__ a, const int* __restrict__ perm, int K,
                                    int64_t lda) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t m = blockIdx.y;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < K; k += gridDim.x * blockDim.x)
    out[m * K + k] = a[m * lda + perm[k]];
}

// ---- act-order with a partial K range (row-parallel shard of a desc_act checkpoint, is_k_full = false) ------------------
// The shard's rows are sorted by group (gptq_marlin.py:398-407) but a group's rows are spread over the ranks, so the runs of
// equal g_idx have arbitrary lengths and the [num_groups_of_the_FULL_K, N] scale table is indexed per ROW -- the reference
// does this with cache_scales_chunk (gptq_marlin.cu:1416-1417).  It only arises for desc_act checkpoints under tensor
// parallelism; this is a plain CUDA-core kernel (correct for every M, not tuned): one CTA = 64 output channels x 32 rows,
// walks the 16-row slabs of the Marlin layout, expands a slab into shared memory with the per-row scale, FMAs in fp32.
template <typename T, int BITS>
__global__ void __launch_bounds__(128) marlin_actorder_partial_kernel(T* __restrict__ C, const T* __restrict__ A,
                                                                      const uint32_t* __restrict__ B,
                                                                      const T* __restrict__ scales,
                                                                      const int* __restrict__ g_idx, int M, int N, int K) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float Wd[16][64 + 1];
  __shared__ float As[32][16 + 1];
  const int t64 = blockIdx.x, m0 = blockIdx.y * 32, tid = threadIdx.x;
  const int n_own = tid & 63, mh = tid >> 6;
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  constexpr int WPS = 64 * 16 * BITS / 32;            // packed words of one 16 x 64 tile: 128 (4-bit) / 256 (8-bit)
  const int words_per_row = N * 16 * BITS / 32;
  for (int slab = 0; slab < K / 16; ++slab) {
    const uint32_t* brow = B + (size_t)slab * words_per_row + (size_t)t64 * WPS;
    if (BITS == 4) {
      const uint32_t w = brow[tid];
      const int l = tid >> 2, j = tid & 3, c = l >> 2, q = l & 3;
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        // nibble p: k = 2q + (p >> 2) + 8 (p & 1), n = 16 j + c + 8 ((p >> 1) & 1)   (marlin_perms.py:9-43)
        const int k = 2 * q + (p >> 2) + 8 * (p & 1), n = 16 * j + c + 8 * ((p >> 1) & 1);
        const int ng = t64 * 64 + n;
        const float sc = to_f32<T>(scales[(size_t)g_idx[slab * 16 + k] * N + (ng & ~63) + 8 * (ng & 7) + ((ng & 63) >> 3)]);
        Wd[k][n] = (float)((int)((w >> (4 * p)) & 15u) - 8) * sc;
      }
    } else {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int wi = tid * 2 + h;                   // [lane (8 words)][n16-tile j (2 words)][b]
        const uint32_t w = brow[wi];
        const int l = wi >> 3, j = (wi >> 1) & 3, b = wi & 1, c = l >> 2, q = l & 3;
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          // byte p: k = 2q + (p >> 1) + 8 (p & 1), n = 16 j + c + 8 b
          const int k = 2 * q + (p >> 1) + 8 * (p & 1), n = 16 * j + c + 8 * b;
          const int ng = t64 * 64 + n;
          const float sc = to_f32<T>(scales[(size_t)g_idx[slab * 16 + k] * N + (ng & ~63) + 8 * (ng & 7) + ((ng & 63) >> 3)]);
          Wd[k][n] = (float)((int)((w >> (8 * p)) & 255u) - 128) * sc;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + 128 * i, mm = e >> 4, kk = e & 15;
      As[mm][kk] = (m0 + mm < M) ? to_f32<T>(A[(size_t)(m0 + mm) * K + slab * 16 + kk]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      const float wv = Wd[kk][n_own];
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fmaf(As[mh * 16 + i][kk], wv, acc[i]);
    }
    __syncthreads();
  }
  const int n = t64 * 64 + n_own;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int m = m0 + mh * 16 + i;
    if (m < M && n < N) C[(size_t)m * N + n] = from_f32<T>(acc[i]);
  }
}

// ---- per-device scratch: split-K partial slots, tile counters, act-order activation copy ---------------------------
struct DeviceScratch {
  float* partials = nullptr;
  int* counters = nullptr;
  void* a_tmp = nullptr;
  size_t partial_floats = 0;
  size_t a_tmp_bytes = 0;
  int counter_ints = 0;
  float* dec_partials = nullptr;
  int* dec_counters = nullptr;
  float* dec_asum = nullptr;
};
static DeviceScratch g_scratch[64];

static bool capturing(cudaStream_t st) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cap);
  return cap != cudaStreamCaptureStatusNone;
}

// Scratch policy (CUDA-graph safety): a buffer that a captured graph may point to is NEVER freed.  The first call sizes
// the scratch for the largest request any of the three GEMM kernels makes at decode / prefill shapes
// (b200_gemm_scratch_reserve can raise that before capture); a later, larger request outside capture allocates a new
// buffer and RETIRES the old one (kept alive until process exit), so graphs captured earlier keep replaying correctly;
// during capture a too-small scratch is an error, never a reallocation.
int gemm_ensure_scratch(int dev, size_t partial_floats, int counter_ints, cudaStream_t st, GemmScratch* out) {
  DeviceScratch& s = g_scratch[dev & 63];
  if (!(s.partials && s.partial_floats >= partial_floats && s.counter_ints >= counter_ints)) {
    if (capturing(st))
      return set_error(B200_EINVAL,
                       "gemm: split-K scratch must be reserved before CUDA-graph capture "
                       "(run one eager call of the largest shape or b200_gemm_scratch_reserve first)");
    // floor: scaled_mm at MT = 256 on a full grid (num_sms * 2 * 128 * 256 floats, 39 MB) covers every decode shape
    const size_t floor_floats = (size_t)num_sms() * 2 * kTileN * 256;
    if (partial_floats < floor_floats) partial_floats = floor_floats;
    if (partial_floats < s.partial_floats) partial_floats = s.partial_floats;
    if (counter_ints < 1 << 16) counter_ints = 1 << 16;
    if (counter_ints < s.counter_ints) counter_ints = s.counter_ints;
    float* np = nullptr;
    int* nc = nullptr;
    if (cudaMalloc(&np, partial_floats * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&nc, (size_t)counter_ints * sizeof(int)) != cudaSuccess) {
      cudaGetLastError();
      return set_error(B200_ECUDA, "gemm: scratch allocation failed (%zu floats)", partial_floats);
    }
    cudaMemsetAsync(nc, 0, (size_t)counter_ints * sizeof(int), st);
    // the retired buffers stay allocated on purpose (see above); counters of the old buffer are all zero between launches
    s.partials = np;
    s.counters = nc;
    s.partial_floats = partial_floats;
    s.counter_ints = counter_ints;
  }
  if (out) {
    out->partials = s.partials;
    out->counters = s.counters;
  }
  return B200_OK;
}

// decode kernel: fixed-size partial slots (2 per CTA) and ticket counters, allocated once
int dec_scratch(int dev, cudaStream_t st, float** partials, int** counters) {
  DeviceScratch& s = g_scratch[dev & 63];
  if (!s.dec_partials) {
    if (capturing(st))
      return set_error(B200_EINVAL, "marlin decode gemm: scratch must exist before CUDA-graph capture (run one eager call or "
                                    "b200_gemm_scratch_reserve first)");
    if (cudaMalloc(&s.dec_partials, (size_t)kDecMaxGrid * 2 * 64 * kTileN * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&s.dec_counters, (size_t)kDecMaxTiles * sizeof(int)) != cudaSuccess) {
      cudaGetLastError();
      s.dec_partials = nullptr;
      return set_error(B200_ECUDA, "marlin decode gemm: scratch allocation failed");
    }
    cudaMemsetAsync(s.dec_counters, 0, (size_t)kDecMaxTiles * sizeof(int), st);
  }
  *partials = s.dec_partials;
  *counters = s.dec_counters;
  return B200_OK;
}

static int ensure_a_tmp(int dev, size_t bytes, cudaStream_t st) {
  DeviceScratch& s = g_scratch[dev & 63];
  if (s.a_tmp && s.a_tmp_bytes >= bytes) return B200_OK;
  if (capturing(st))
    return set_error(B200_EINVAL, "marlin gemm: act-order scratch must be reserved before CUDA-graph capture (run one eager call first)");
  if (bytes < ((size_t)8 << 20)) bytes = (size_t)8 << 20;
  // the old buffer (if any) is retired, not freed: a captured graph may still point to it
  if (cudaMalloc(&s.a_tmp, bytes) != cudaSuccess) return set_error(B200_ECUDA, "marlin gemm: act-order scratch allocation failed");
  s.a_tmp_bytes = bytes;
  return B200_OK;
}

// decode kernel v2: activation group sums when the caller does not bring them
int dec_asum_scratch(int dev, cudaStream_t st, float** asum) {
  DeviceScratch& s = g_scratch[dev & 63];
  if (!s.dec_asum) {
    if (capturing(st))
      return set_error(B200_EINVAL, "marlin decode gemm: scratch must exist before CUDA-graph capture (run one eager call or "
                                    "b200_gemm_scratch_reserve first)");
    if (cudaMalloc(&s.dec_asum, (size_t)kDecMaxGroups * 64 * sizeof(float)) != cudaSuccess) {
      cudaGetLastError();
      s.dec_asum = nullptr;
      return set_error(B200_ECUDA, "marlin decode gemm: scratch allocation failed");
    }
  }
  *asum = s.dec_asum;
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_gemm_scratch_reserve(int64_t max_m, int64_t max_n) {
  using namespace b200;
  int dev = 0;
  cudaGetDevice(&dev);
  float* p;
  int* c;
  if (int e = dec_scratch(dev, 0, &p, &c)) return e;
  if (int e = dec_asum_scratch(dev, 0, &p)) return e;
  // largest request of the three kernels for a [max_m, max_n] output: marlin (grid x 128 x MPAD), scaled_mm
  // (grid x 2 x 128 x MT), marlin24 split-K (n_tiles x m_chunks x splits x 64 x 64)
  const size_t m = (size_t)(max_m > 0 ? max_m : 1), n = (size_t)(max_n > 0 ? max_n : 1);
  const size_t m_pad = (m + 63) / 64 * 64;
  size_t want = (size_t)num_sms() * 2 * kTileN * 256;
  const size_t m24 = ((n + 127) / 128) * m_pad * 64 * 8;
  if (m24 > want) want = m24;
  const size_t tiles = ((n + 63) / 64) * ((m + 15) / 16);
  return gemm_ensure_scratch(dev, want, (int)(tiles < (1u << 20) ? tiles : (1u << 20)), 0, nullptr);
}

// silu: 0 = plain GEMM into c [M, N]; 1 = gate_up layer, c [M, N / 2] = silu_and_mul(GEMM) (see the header)
static int marlin_gemm_impl(void* c, const void* a, const int32_t* b_q_weight, const void* b_scales, const int32_t* g_idx,
                            const int32_t* perm, int32_t* workspace, int64_t workspace_numel, int num_bits, int64_t size_m,
                            int64_t size_n, int64_t size_k, int64_t lda, int num_groups, int is_k_full, int dtype,
                            void* stream, int silu) {
  using namespace b200;
  (void)workspace;
  // a fused all-reduce requested for "the next GEMM of this thread" (b200_ar_fuse_next_gemm) is consumed HERE, before any
  // validation: a rejected or empty call must not leave the request armed for some later, unrelated GEMM
  long long ar_m = 0, ar_n = 0;
  ArFuse ar_req;
  const bool ar_pending = ar_take_pending(&ar_req, &ar_m, &ar_n);
  if (silu) {
    B200_CHECK(!ar_pending, "gptq_marlin_gemm_silu_and_mul: gate_up is column-parallel, a fused all-reduce does not apply");
    B200_CHECK(g_idx == nullptr && perm == nullptr, "gptq_marlin_gemm_silu_and_mul: act-order layers are not supported");
    B200_CHECK(size_n % 128 == 0, "gptq_marlin_gemm_silu_and_mul: size_n / 2 = %lld must be a multiple of 64", (long long)(size_n / 2));
  }
  B200_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = %d", num_bits);
  B200_CHECK(dtype == B200_F16 || dtype == B200_BF16, "gpt_marlin_gemm only supports bfloat16 and float16");
  B200_CHECK(size_n % 64 == 0, "size_n = %lld, is not divisible by min_thread_n = 64", (long long)size_n);
  B200_CHECK(size_k % 128 == 0, "size_k = %lld is not divisible by 128 (gptq_marlin.py:21 min_thread_k)", (long long)size_k);
  B200_CHECK(workspace_numel >= (size_n / 64) * 16, "workspace.numel = %lld is below min_workspace_size = %lld",
             (long long)workspace_numel, (long long)((size_n / 64) * 16));
  B200_CHECK(num_groups >= 1 && size_k % num_groups == 0, "size_k = %lld, is not divisible by b_scales.size(0) = %d",
             (long long)size_k, num_groups);
  B200_CHECK((reinterpret_cast<uintptr_t>(a) & 15) == 0 && (lda * 2) % 16 == 0, "A must be 16-byte aligned with a 16-byte multiple row pitch");
  const bool act_order = (g_idx != nullptr && perm != nullptr);
  B200_CHECK((g_idx == nullptr) == (perm == nullptr), "g_idx and perm must both be given (act-order) or both be empty");
  if (size_m == 0 || size_n == 0) return B200_OK;
  const int gs = (int)(size_k / num_groups);
  if (num_groups > 1) B200_CHECK(gs == 32 || gs == 64 || gs == 128, "Unexpected group_size = %d", gs);
  cudaStream_t st = (cudaStream_t)stream;
  if (act_order) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (int e = ensure_a_tmp(dev, (size_t)size_m * size_k * 2, st)) return e;
    void* tmp = g_scratch[dev & 63].a_tmp;
    dim3 grid((unsigned)((size_k + 255) / 256), (unsigned)size_m);
    if (dtype == B200_F16)
      launch_pdl(permute_cols_kernel<__half>, grid, dim3(256), 0, st, (__half*)tmp, (const __half*)a, perm, (int)size_k, lda);
    else
      launch_pdl(permute_cols_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, (__nv_bfloat16*)tmp, (const __nv_bfloat16*)a, perm, (int)size_k, lda);
    a = tmp;
    lda = size_k;
    if (!is_k_full) {
      // partial K: per-row scale lookup through g_idx (see marlin_actorder_partial_kernel)
      if (ar_pending)
        return set_error(B200_ENOTSUP, "gptq_marlin_gemm: the fused all-reduce is not available for act-order with is_k_full=False");
      dim3 g2((unsigned)(size_n / 64), (unsigned)((size_m + 31) / 32));
#define AOP_GO(TT, BB)                                                                                              \
  launch_pdl(marlin_actorder_partial_kernel<TT, BB>, g2, dim3(128), 0, st, (TT*)c, (const TT*)a, (const uint32_t*)b_q_weight, \
             (const TT*)b_scales, g_idx, (int)size_m, (int)size_n, (int)size_k)
      if (dtype == B200_F16) { if (num_bits == 4) AOP_GO(__half, 4); else AOP_GO(__half, 8); }
      else { if (num_bits == 4) AOP_GO(__nv_bfloat16, 4); else AOP_GO(__nv_bfloat16, 8); }
#undef AOP_GO
      B200_LAUNCH_CHECK("gptq_marlin_gemm (act-order, partial K)");
      return B200_OK;
    }
  }
  GemmArgs ga;
  ga.B = reinterpret_cast<const uint8_t*>(b_q_weight);
  ga.scales = b_scales;
  ga.C = c;
  ga.scratch = nullptr;
  ga.counters = nullptr;
  ga.M = (int)size_m; ga.N = (int)size_n; ga.K = (int)size_k;
  ga.num_groups = num_groups; ga.group_size = gs;
  ga.n_tiles = (int)((size_n + kTileN - 1) / kTileN);
  ga.m_chunks = 1;
  ga.KS = (int)((size_k + kStageK - 1) / kStageK);
  ga.total_units = 0;
  ga.groups_per_stage = num_groups > 1 ? kStageK / gs : 0;
  ga.prof = nullptr; ga.trace = nullptr; ga.trace_cta = 0; ga.exp = 0;
  ga.cluster_s = 0;
  ga.asum = nullptr; ga.asum_stride = 0;
  ga.silu_inter = 0;
  if (silu) {
    ga.silu_inter = (int)(size_n / 2);
    ga.n_tiles = (int)(size_n / 128);   // one tile = 64 gate + 64 up columns
  }
  ga.ar.world = 0;
  if (ar_pending) {   // b200_ar_fuse_next_gemm: all-reduce this GEMM's output in its epilogue
    if (ar_m != size_m || ar_n != size_n)
      return set_error(B200_EINVAL, "gptq_marlin_gemm: fused all-reduce was requested for [%lld, %lld], the GEMM is [%lld, %lld]",
                       ar_m, ar_n, (long long)size_m, (long long)size_n);
    ga.ar = ar_req;
  }
  int rc;
  // B200_GEMM_DECODE=1: the experimental small-CTA decode kernels (marlin_decode_kernel.cuh; parity-green, but measured
  // slower than the kernel below on every Llama shape -- DESIGN.md section 6b has the numbers and the ncu evidence)
  static const int use_decode = getenv("B200_GEMM_DECODE") ? atoi(getenv("B200_GEMM_DECODE")) : 0;
  if (size_m <= 64 && ga.ar.world <= 1 && use_decode && !silu && ga.n_tiles <= kDecMaxTiles) {
    if (dtype == B200_F16) rc = (num_bits == 4) ? dispatch_decode_f16_4(ga, a, lda, st) : dispatch_decode_f16_8(ga, a, lda, st);
    else rc = (num_bits == 4) ? dispatch_decode_bf16_4(ga, a, lda, st) : dispatch_decode_bf16_8(ga, a, lda, st);
    if (rc) return rc;
    B200_LAUNCH_CHECK("gptq_marlin_gemm");
    return B200_OK;
  }
  if (dtype == B200_F16) rc = (num_bits == 4) ? dispatch_gemm_f16_4(ga, a, lda, st) : dispatch_gemm_f16_8(ga, a, lda, st);
  else rc = (num_bits == 4) ? dispatch_gemm_bf16_4(ga, a, lda, st) : dispatch_gemm_bf16_8(ga, a, lda, st);
  if (rc) return rc;
  B200_LAUNCH_CHECK("gptq_marlin_gemm");
  return B200_OK;
}

extern "C" int b200_gptq_marlin_gemm(void* c, const void* a, const int32_t* b_q_weight, const void* b_scales,
                                     const int32_t* g_idx, const int32_t* perm, int32_t* workspace,
                                     int64_t workspace_numel, int num_bits, int64_t size_m, int64_t size_n,
                                     int64_t size_k, int64_t lda, int num_groups, int is_k_full, int dtype,
                                     void* stream) {
  return marlin_gemm_impl(c, a, b_q_weight, b_scales, g_idx, perm, workspace, workspace_numel, num_bits, size_m, size_n, size_k,
                          lda, num_groups, is_k_full, dtype, stream, 0);
}

extern "C" int b200_gptq_marlin_gemm_silu_and_mul(void* out, const void* a, const int32_t* b_q_weight, const void* b_scales,
                                                  int32_t* workspace, int64_t workspace_numel, int num_bits, int64_t size_m,
                                                  int64_t size_n, int64_t size_k, int64_t lda, int num_groups, int dtype,
                                                  void* stream) {
  return marlin_gemm_impl(out, a, b_q_weight, b_scales, nullptr, nullptr, workspace, workspace_numel, num_bits, size_m, size_n,
                          size_k, lda, num_groups, 1, dtype, stream, 1);
}
