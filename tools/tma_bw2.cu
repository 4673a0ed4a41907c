// Bring-up microbenchmark 2: streaming HBM -> smem with bulk TMA as a function of grid size, CTAs/SM, request size and
// total bytes; 10 launches back to back between the events (sustained), plus an empty-kernel baseline.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include "../nm_vllm_b200/csrc/ptx.cuh"
using namespace b200::ptx;

__global__ void empty_kernel(int* sink) { if (sink == nullptr) printf("x"); }

// units of `req` bytes are dealt round-robin to CTAs (unit u -> CTA u % grid) so every grid size reads the same bytes
__global__ void __launch_bounds__(160, 1)
bw2(const uint8_t* base, int req, int reqs_per_stage, int nstages, long long total_units, int* sink, int P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sb = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stage_bytes = req * reqs_per_stage;
  const uint32_t bar = sb + nstages * stage_bytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < nstages; ++s) { mbar_init(bar + 8 * s, 1); mbar_init(bar + 8 * (nstages + s), 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const long long stages_total = total_units / reqs_per_stage;
  const int my = (int)((stages_total - blockIdx.x + gridDim.x - 1) / gridDim.x);
  if (warp < P) {
    if (lane == 0) {
      const uint64_t pol = policy_evict_first();
      for (int it = warp; it < my; it += P) {
        const int s = it % nstages;
        mbar_wait(bar + 8 * (nstages + s), ((it / nstages) & 1) ^ 1);
        mbar_arrive_expect_tx(bar + 8 * s, stage_bytes);
        const long long st = (long long)it * gridDim.x + blockIdx.x;
        for (int r = 0; r < reqs_per_stage; ++r)
          bulk_g2s_hint(sb + s * stage_bytes + r * req, base + (size_t)(st * reqs_per_stage + r) * req, req, bar + 8 * s, pol);
      }
    }
  } else if (warp == 4) {
    int acc = 0;
    for (int it = 0; it < my; ++it) {
      const int s = it % nstages;
      mbar_wait(bar + 8 * s, (it / nstages) & 1);
      acc += *reinterpret_cast<volatile int*>(smem_raw + (sb - smem_u32(smem_raw)) + s * stage_bytes + lane * 4);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar + 8 * (nstages + s));
    }
    if (acc == 0x12345678) sink[0] = acc;
  }
}

int main() {
  const size_t big = 59ull << 20;
  const int ncopy = 6;
  uint8_t* buf; cudaMalloc(&buf, big * ncopy); cudaMemset(buf, 1, big * ncopy);
  int* sink; cudaMalloc(&sink, 4);
  cudaFuncSetAttribute(bw2, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int L = 10;
  {
    float best = 1e9;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(e0);
      for (int l = 0; l < L; ++l) empty_kernel<<<148, 64>>>(sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    printf("empty kernel x%d: %.2f us per launch\n", L, best * 1e3 / L);
  }
  for (size_t total : {big}) {
    for (int req : {1024, 2048, 4096, 8192, 16384}) {
      for (int rps : {1, 4}) {
        const int stage = req * rps;
        if (stage > 65536) continue;
        for (int smem_kb : {96}) { for (int P : {1, 2, 4}) {
          const int nst = (smem_kb * 1024 - 2048) / stage;
          if (nst < 2) continue;
          for (int grid : {148}) {
            
            const long long units = total / req;
            float best = 1e9;
            for (int rep = 0; rep < 6; ++rep) {
              cudaEventRecord(e0);
              for (int l = 0; l < L; ++l)
                bw2<<<grid, 160, smem_kb * 1024>>>(buf + (size_t)((rep * L + l) % ncopy) * big, req, rps, nst, units, sink, P);
              cudaEventRecord(e1); cudaEventSynchronize(e1);
              float ms; cudaEventElapsedTime(&ms, e0, e1);
              if (rep >= 2 && ms < best) best = ms;
            }
            cudaError_t e = cudaGetLastError();
            printf("P=%d total=%5.1fMB req=%5d x%d smem=%3dKB nst=%3d grid=%3d: %7.2f us/launch %7.1f GB/s %s\n", P, total / 1048576.0, req, rps,
                   smem_kb, nst, grid, best * 1e3 / L, (double)total / (best * 1e-3 / L) / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
            fflush(stdout);
          }
        } }
      }
    }
  }
  return 0;
}
