// Bring-up microbenchmark: issue/execute rate of small tcgen05.mma (M=128, N=NMMA, K=16, A from TMEM, B from smem),
// from 1..4 issuing threads (separate warps, separate accumulators).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include "../nm_vllm_b200/csrc/ptx.cuh"
using namespace b200::ptx;

template <int NMMA, bool A_TMEM>
__global__ void __launch_bounds__(256, 1) k(int iters, int nthreads, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sb = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tslot;
  __shared__ uint64_t bars[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bars[i]), 1); fence_barrier_init(); }
  if (warp == 7) tmem_alloc<512>(smem_u32(&tslot));
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tslot;
  constexpr uint32_t idesc = make_idesc(1, 0, 0, 128, NMMA);
  long long t0 = 0, t1 = 0;
  if (warp < nthreads && lane == 0) {
    const uint32_t d = tb + 256 + warp * 64;
    const uint64_t bdesc = make_kmajor_sw128_desc(sb + warp * 16384);
    const uint64_t adesc = make_kmajor_sw128_desc(sb + 65536);
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        if (A_TMEM) mma_f16_ts(d, tb + (i & 3) * 64 + r * 8, bdesc + (r & 3) * 2, idesc, 1);
        else mma_f16_ss(d, adesc + (r & 3) * 2, bdesc + (r & 3) * 2, idesc, 1);
      }
    }
    t1 = clock64();
    tc_commit(smem_u32(&bars[warp]));
    mbar_wait(smem_u32(&bars[warp]), 0);
    long long t2 = clock64();
    out[blockIdx.x * 8 + warp * 2] = t1 - t0;
    out[blockIdx.x * 8 + warp * 2 + 1] = t2 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 7) tmem_dealloc<512>(tb);
}

template <int NMMA, bool A_TMEM>
void run(const char* name) {
  long long* out; cudaMalloc(&out, 148 * 8 * 8);
  cudaFuncSetAttribute(k<NMMA, A_TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 140 * 1024);
  for (int nt = 1; nt <= 4; nt *= 2) {
    const int iters = 2000;
    k<NMMA, A_TMEM><<<148, 256, 140 * 1024>>>(iters, nt, out);
    cudaDeviceSynchronize();
    k<NMMA, A_TMEM><<<148, 256, 140 * 1024>>>(iters, nt, out);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[8]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%s N=%3d threads=%d: issue %.1f cyc/MMA, complete %.1f cyc/MMA per thread (aggregate %.1f cyc/MMA) %s\n", name, NMMA, nt,
           (double)h[0] / (iters * 8), (double)h[1] / (iters * 8), (double)h[1] / (iters * 8) / nt, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  cudaFree(out);
}

int main() {
  run<16, true>("A=tmem");
  run<32, true>("A=tmem");
  run<64, true>("A=tmem");
  run<128, true>("A=tmem");
  run<256, true>("A=tmem");
  run<16, false>("A=smem");
  run<64, false>("A=smem");
  return 0;
}
