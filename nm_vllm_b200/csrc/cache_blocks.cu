// KV block maintenance ops: copy_blocks, swap_blocks, reshape_and_cache_flash.
//   copy_blocks             : reference csrc/cache_kernels.cu:64-148  (copy-on-write of shared blocks, all layers at once)
//   swap_blocks             : reference csrc/cache_kernels.cu:24-62   (preemption swap HBM <-> pinned host, or HBM <-> HBM)
//   reshape_and_cache_flash : reference csrc/cache_kernels.cu:207-236 (kernel), :280-314 (op): [NB, BS, H, D] layout
// All three are pure byte moves: results are bit-exact.
// B200 notes: a KV block of every layout is a multiple of 16 bytes, so copy_blocks moves int4 vectors regardless of
// the element type, with the grid cut into (layer, pair, chunk) so that a handful of pairs still fills 148 SMs;
// swap_blocks merges runs of consecutive (src, dst) block numbers into one cudaMemcpyAsync each (the scheduler hands
// out mostly contiguous ranges, and one DMA descriptor per 2 MB run beats 128 descriptors of 16 KB).
#include "common.cuh"

namespace b200 {

constexpr int kCopyThreads = 256;
constexpr int kCopyVecPerCta = 8 * kCopyThreads;   // 32 KB per CTA

__global__ void __launch_bounds__(kCopyThreads)
copy_blocks_kernel(const int64_t* __restrict__ key_cache_ptrs, const int64_t* __restrict__ value_cache_ptrs,
                   const int64_t* __restrict__ block_mapping, const int64_t vec_per_block, const int chunks) {
  pdl_wait();
  pdl_launch_dependents();
  const int layer = blockIdx.x, pair = blockIdx.y;
  const int chunk = blockIdx.z % chunks;
  const bool is_value = blockIdx.z >= chunks;
  int4* cache = reinterpret_cast<int4*>(is_value ? value_cache_ptrs[layer] : key_cache_ptrs[layer]);
  const int64_t src = block_mapping[2 * pair], dst = block_mapping[2 * pair + 1];
  const int4* s = cache + src * vec_per_block;
  int4* d = cache + dst * vec_per_block;
  const int64_t v0 = (int64_t)chunk * kCopyVecPerCta;
  const int64_t v1 = v0 + kCopyVecPerCta < vec_per_block ? v0 + kCopyVecPerCta : vec_per_block;
  for (int64_t v = v0 + threadIdx.x; v < v1; v += kCopyThreads) d[v] = s[v];
}

template <typename scalar_t>
__global__ void reshape_and_cache_flash_kernel(const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
                                               scalar_t* __restrict__ k_cache, scalar_t* __restrict__ v_cache,
                                               const int64_t* __restrict__ slot_mapping, const int64_t block_stride,
                                               const int64_t key_stride, const int64_t value_stride, const int n,
                                               const int block_size) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;   // padded token
  const int64_t tgt = (slot / block_size) * block_stride + (slot % block_size) * (int64_t)n;
  const scalar_t* ks = key + token * key_stride;
  const scalar_t* vs = value + token * value_stride;
  scalar_t* kd = k_cache + tgt;
  scalar_t* vd = v_cache + tgt;
  constexpr int VEC = 16 / sizeof(scalar_t);
  const bool vec_ok = n % VEC == 0 && ((reinterpret_cast<uintptr_t>(ks) | reinterpret_cast<uintptr_t>(vs) |
                                        reinterpret_cast<uintptr_t>(kd) | reinterpret_cast<uintptr_t>(vd)) & 15) == 0;
  if (vec_ok) {
    for (int i = threadIdx.x; i < n / VEC; i += blockDim.x) {
      reinterpret_cast<int4*>(kd)[i] = reinterpret_cast<const int4*>(ks)[i];
      reinterpret_cast<int4*>(vd)[i] = reinterpret_cast<const int4*>(vs)[i];
    }
  } else {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      kd[i] = ks[i];
      vd[i] = vs[i];
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs, const int64_t* block_mapping,
                     int num_layers, int64_t num_pairs, int64_t block_bytes, void* stream) {
  if (num_layers == 0 || num_pairs == 0) return B200_OK;
  B200_CHECK(block_bytes > 0 && block_bytes % 16 == 0, "copy_blocks: block of %lld bytes is not a multiple of 16",
             (long long)block_bytes);
  B200_CHECK(num_pairs <= 65535, "copy_blocks: %lld pairs exceed the grid limit", (long long)num_pairs);
  const int64_t vec = block_bytes / 16;
  const int chunks = (int)((vec + kCopyVecPerCta - 1) / kCopyVecPerCta);
  B200_CHECK(2 * chunks <= 65535, "copy_blocks: block too large");
  dim3 grid((unsigned)num_layers, (unsigned)num_pairs, (unsigned)(2 * chunks));
  cudaError_t e = launch_pdl(copy_blocks_kernel, grid, dim3(kCopyThreads), 0, (cudaStream_t)stream, key_cache_ptrs,
                             value_cache_ptrs, block_mapping, vec, chunks);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "copy_blocks: %s", cudaGetErrorString(e));
  return B200_OK;
}

/* block_mapping: HOST int64 [num_pairs, 2] (the reference requires a CPU tensor, cache_kernels.cu:43).
 * direction: 0 device->device, 1 device->host, 2 host->device */
int b200_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host, int64_t num_pairs,
                     int64_t block_bytes, int direction, void* stream) {
  B200_CHECK(direction >= 0 && direction <= 2, "swap_blocks: Invalid device combination");
  B200_CHECK(block_bytes > 0, "swap_blocks: bad block size");
  const cudaMemcpyKind kind = direction == 0 ? cudaMemcpyDeviceToDevice
                              : direction == 1 ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  const char* s = (const char*)src;
  char* d = (char*)dst;
  int64_t i = 0;
  while (i < num_pairs) {
    const int64_t s0 = block_mapping_host[2 * i], d0 = block_mapping_host[2 * i + 1];
    int64_t run = 1;
    while (i + run < num_pairs && block_mapping_host[2 * (i + run)] == s0 + run &&
           block_mapping_host[2 * (i + run) + 1] == d0 + run)
      ++run;
    cudaError_t e = cudaMemcpyAsync(d + d0 * block_bytes, s + s0 * block_bytes, (size_t)(run * block_bytes), kind,
                                    (cudaStream_t)stream);
    if (e != cudaSuccess) return set_error(B200_ECUDA, "swap_blocks: %s", cudaGetErrorString(e));
    i += run;
  }
  return B200_OK;
}

int b200_reshape_and_cache_flash(const void* key, const void* value, void* k_cache, void* v_cache,
                                 const int64_t* slot_mapping, int64_t num_tokens, int num_heads, int head_size,
                                 int block_size, int64_t block_stride, int64_t key_stride, int64_t value_stride,
                                 int dtype, int kv_dtype, void* stream) {
  if (kv_dtype != B200_KV_AUTO) return set_error(B200_ENOTSUP, "Unsupported data type of kv cache: %d", kv_dtype);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(block_size > 0 && num_heads > 0 && head_size > 0, "reshape_and_cache_flash: bad geometry");
  const int n = num_heads * head_size;
  dim3 grid((unsigned)num_tokens), block(n < 512 ? ((n + 31) / 32) * 32 : 512);
  cudaError_t e = cudaSuccess;
  B200_DISPATCH_FLOAT(dtype, "reshape_and_cache_flash", {
    e = launch_pdl(reshape_and_cache_flash_kernel<scalar_t>, grid, block, 0, (cudaStream_t)stream,
                   (const scalar_t*)key, (const scalar_t*)value, (scalar_t*)k_cache, (scalar_t*)v_cache, slot_mapping,
                   block_stride, key_stride, value_stride, n, block_size);
  });
  if (e != cudaSuccess) return set_error(B200_ECUDA, "reshape_and_cache_flash: %s", cudaGetErrorString(e));
  return B200_OK;
}

}  // extern "C"
