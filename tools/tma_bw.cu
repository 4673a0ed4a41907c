// Bring-up microbenchmark: how fast can one producer thread per SM stream HBM into shared memory with bulk/tensor TMA?
// modes: 0 = one contiguous bulk copy of STAGE bytes per stage; 1 = 8 bulk copies of STAGE/8 bytes, rows `pitch` apart;
//        2 = one 2-D tensor-map box (8 rows x STAGE/8 bytes), rows `pitch` apart.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include "../nm_vllm_b200/csrc/ptx.cuh"
#include "../nm_vllm_b200/csrc/tma_host.cuh"
using namespace b200::ptx;

__global__ void __launch_bounds__(128, 1)
bw_kernel(const __grid_constant__ CUtensorMap tmap, const uint8_t* base, int mode, int stage_bytes, int nstages,
          int units_per_cta, size_t pitch, int tiles_per_row, int hint, int* sink) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sb = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = sb + nstages * stage_bytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < nstages; ++s) { mbar_init(bar + 8 * s, 1); mbar_init(bar + 8 * (nstages + s), 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const int row_bytes = stage_bytes / 8;
  const uint64_t pol = policy_evict_first();
  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < units_per_cta; ++it) {
        const int s = it % nstages;
        mbar_wait(bar + 8 * (nstages + s), ((it / nstages) & 1) ^ 1);
        mbar_arrive_expect_tx(bar + 8 * s, stage_bytes);
        const long long u = (long long)blockIdx.x * units_per_cta + it;
        if (mode == 0) {
          if (hint) bulk_g2s_hint(sb + s * stage_bytes, base + (size_t)u * stage_bytes, stage_bytes, bar + 8 * s, pol);
          else bulk_g2s(sb + s * stage_bytes, base + (size_t)u * stage_bytes, stage_bytes, bar + 8 * s);
        } else {
          // unit u -> (tile column t, k block kb): marlin-like: consecutive units of a CTA walk down K within one tile column
          const int units_per_col = tiles_per_row;  // (re-used arg) = number of 8-slab k blocks per tile column
          const int t = (int)(u / units_per_col), kb = (int)(u % units_per_col);
          if (mode == 1) {
            for (int r = 0; r < 8; ++r) {
              const uint8_t* src = base + (size_t)(kb * 8 + r) * pitch + (size_t)t * row_bytes;
              if (hint) bulk_g2s_hint(sb + s * stage_bytes + r * row_bytes, src, row_bytes, bar + 8 * s, pol);
              else bulk_g2s(sb + s * stage_bytes + r * row_bytes, src, row_bytes, bar + 8 * s);
            }
          } else {
            if (hint) tma_2d_g2s_hint(sb + s * stage_bytes, &tmap, t * (row_bytes / 8), kb * 8, bar + 8 * s, pol);
            else tma_2d_g2s(sb + s * stage_bytes, &tmap, t * (row_bytes / 8), kb * 8, bar + 8 * s);
          }
        }
      }
    }
  } else if (warp == 1) {
    int acc = 0;
    for (int it = 0; it < units_per_cta; ++it) {
      const int s = it % nstages;
      mbar_wait(bar + 8 * s, (it / nstages) & 1);
      acc += *reinterpret_cast<volatile int*>(smem_raw + (sb - smem_u32(smem_raw)) + s * stage_bytes + lane * 4);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + 8 * (nstages + s));
    }
    if (acc == 0x12345678) sink[0] = acc;
  }
}

int main(int argc, char** argv) {
  const int grid = 148;
  const size_t K = 4096, N = 28672;           // gate_up
  const size_t pitch = N * 8;                 // bytes per k16 row of a 4-bit marlin tensor
  const size_t total = K * N / 2;
  const int ncopy = 8;
  uint8_t* buf; cudaMalloc(&buf, total * ncopy); cudaMemset(buf, 1, total * ncopy);
  int* sink; cudaMalloc(&sink, 4);
  cudaFuncSetAttribute(bw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int stage_bytes : {4096, 8192, 16384, 32768}) {
    for (int mode = 0; mode < 3; ++mode) {
      for (int hint = 0; hint < 2; ++hint) {
        for (int nst : {4, 8, 16, 24}) {
          if ((size_t)nst * stage_bytes > 200 * 1024) continue;
          const int row_bytes = stage_bytes / 8;
          const int tiles_per_row = (int)(K / 16 / 8);  // k blocks (8 slabs each) per tile column
          size_t units = total / stage_bytes;
          int upc = (int)(units / grid);
          if (upc <= 0) continue;
          float best = 1e9;
          for (int rep = 0; rep < 12; ++rep) {
            const uint8_t* b = buf + (size_t)(rep % ncopy) * total;
            CUtensorMap tm;
            b200::make_tmap_2d(&tm, b, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, K / 16, pitch / 8, pitch, 8, row_bytes / 8, CU_TENSOR_MAP_SWIZZLE_NONE);
            cudaEventRecord(e0);
            bw_kernel<<<grid, 128, nst * stage_bytes + 2 * nst * 8 + 2048>>>(tm, b, mode, stage_bytes, nst, upc, pitch, tiles_per_row, hint, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep >= 4 && ms < best) best = ms;
          }
          cudaError_t e = cudaGetLastError();
          double bytes = (double)grid * upc * stage_bytes;
          printf("stage=%5d mode=%d hint=%d nst=%2d upc=%3d: %7.2f us  %7.1f GB/s %s\n", stage_bytes, mode, hint, nst, upc, best * 1e3,
                 bytes / (best * 1e-3) / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
          fflush(stdout);
        }
      }
    }
  }
  return 0;
}
