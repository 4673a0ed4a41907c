// gptq_marlin_repack for sm_100a -- GPTQ row-packed int32 [K/pf, N] -> Marlin tile layout [K/16, N*16/pf].
// Reference: csrc/quantization/gptq_marlin/gptq_marlin_repack.cu:32-260 (kernel), :276-348 (op); Python twin
// marlin_weights + get_perms (vllm/model_executor/layers/quantization/utils/marlin_utils.py:25-57,
// marlin_perms.py:9-43).  Integer work: bit-exact.
//
// Layout restated (one 16(k) x 64(n) tile = 1024 values, stored as 1024/pf consecutive int32 of output row k/16):
//   "lane" i = 0..31 (c = i/4 column in an 8-wide half tile, q = i%4), n16-tile j = 0..3, half b = 0..1
//   value(i, j, b, r) sits at k = {2q, 2q+1, 2q+8, 2q+9}[r], n = 16 j + c + 8 b
//   4-bit: one int32 per (i, j) at index 4 i + j, nibble order [b0r0, b0r2, b1r0, b1r2, b0r1, b0r3, b1r1, b1r3]
//   8-bit: two int32 per (i, j) at index (4 i + j) * 2 + b, byte order [r0, r2, r1, r3]
// This is a one-off (first apply()), gather-bound op: one thread per output int32, the 4 q-lanes of a column
// read the same source words (broadcast), consecutive c read consecutive n (coalesced 32 B runs).
#include "common.cuh"

namespace b200 {

template <int BITS, bool HAS_PERM>
__global__ void marlin_repack_kernel(const uint32_t* __restrict__ in, const int32_t* __restrict__ perm,
                                     uint32_t* __restrict__ out, int64_t size_k, int64_t size_n) {
  constexpr int PF = 32 / BITS;
  constexpr uint32_t MASK = (1u << BITS) - 1u;
  const int64_t out_row_words = size_n * 16 / PF;
  const int64_t total = (size_k / 16) * out_row_words;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += stride) {
    const int64_t kt = o / out_row_words;
    const int64_t w = o % out_row_words;
    constexpr int WORDS_PER_TILE = 1024 / PF;  // 128 (4-bit) / 256 (8-bit)
    const int64_t t64 = w / WORDS_PER_TILE;
    int r = (int)(w % WORDS_PER_TILE);
    int bsel = 0;
    if (BITS == 8) { bsel = r & 1; r >>= 1; }
    const int i = r >> 2, j = r & 3;
    const int c = i >> 2, q = i & 3;
    uint32_t res = 0;
#pragma unroll
    for (int p = 0; p < PF; ++p) {
      int b, rr;
      if (BITS == 4) {
        // nibble p <- entry e = [0,2,4,6,1,3,5,7][p]; e = 4 b + rr
        const int e = (p < 4) ? 2 * p : 2 * (p - 4) + 1;
        b = e >> 2; rr = e & 3;
      } else {
        b = bsel; rr = (p == 0) ? 0 : (p == 1) ? 2 : (p == 2) ? 1 : 3;
      }
      const int krow = 2 * q + (rr & 1) + ((rr >> 1) << 3);
      int64_t k = kt * 16 + krow;
      if (HAS_PERM) k = perm[k];
      const int64_t n = t64 * 64 + 16 * j + c + 8 * b;
      const uint32_t word = in[(k / PF) * size_n + n];
      res |= ((word >> (BITS * (int)(k % PF))) & MASK) << (BITS * p);
    }
    out[o] = res;
  }
}

}  // namespace b200

extern "C" int b200_gptq_marlin_repack(const int32_t* b_q_weight, const int32_t* perm, int32_t* out, int64_t size_k,
                                       int64_t size_n, int num_bits, void* stream) {
  using namespace b200;
  // preconditions of gptq_marlin_repack.cu:280-304
  B200_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = %d", num_bits);
  B200_CHECK(size_k % 16 == 0, "size_k = %lld is not divisible by tile_k_size = 16", (long long)size_k);
  B200_CHECK(size_n % 64 == 0, "size_n = %lld is not divisible by tile_n_size = 64", (long long)size_n);
  if (size_k == 0 || size_n == 0) return B200_OK;
  const int pf = 32 / num_bits;
  const int64_t total = (size_k / 16) * (size_n * 16 / pf);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = (cudaStream_t)stream;
  const uint32_t* in = (const uint32_t*)b_q_weight;
  uint32_t* o = (uint32_t*)out;
  if (num_bits == 4) {
    if (perm) marlin_repack_kernel<4, true><<<(unsigned)blocks, 256, 0, st>>>(in, perm, o, size_k, size_n);
    else marlin_repack_kernel<4, false><<<(unsigned)blocks, 256, 0, st>>>(in, perm, o, size_k, size_n);
  } else {
    if (perm) marlin_repack_kernel<8, true><<<(unsigned)blocks, 256, 0, st>>>(in, perm, o, size_k, size_n);
    else marlin_repack_kernel<8, false><<<(unsigned)blocks, 256, 0, st>>>(in, perm, o, size_k, size_n);
  }
  B200_LAUNCH_CHECK("gptq_marlin_repack");
  return B200_OK;
}
