// Bring-up helper: issue rate of the legacy warp-level tensor-core instructions on sm_100a (HMMA m16n8k16 bf16,
// HMMA.SP m16n8k32 f16), cycles per instruction per warp at 1..8 warps per SM sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
template <int SP>
__global__ void k(long long* out, int iters) {
  uint32_t a0 = threadIdx.x, a1 = 2, a2 = 3, a3 = 4, b0 = 5, b1 = 6, b2 = 7, b3 = 8, e = 0x44444444;
  float c[4][4] = {};
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (SP)
        asm volatile("mma.sp::ordered_metadata.sync.aligned.m16n8k32.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9,%10,%11}, {%0,%1,%2,%3}, %12, 0x0;"
                     : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(b2), "r"(b3), "r"(e));
      else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
  }
  long long t1 = clock64();
  float s = 0; for (int j = 0; j < 4; ++j) for (int q = 0; q < 4; ++q) s += c[j][q];
  if (threadIdx.x == 0) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = (long long)s; }
}
int main() {
  long long* d; cudaMalloc(&d, 4096);
  for (int sp = 0; sp < 2; ++sp)
    for (int warps = 4; warps <= 32; warps *= 2) {
      const int iters = 2000;
      if (sp) k<1><<<1, warps * 32>>>(d, iters); else k<0><<<1, warps * 32>>>(d, iters);
      cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      double cyc = (double)h[0] / (iters * 4);
      printf("%s warps/SM=%2d: %.1f cycles per instr per warp -> %.3f instr/clk/SM\n", sp ? "HMMA.SP.16832" : "HMMA.16816   ", warps, cyc, warps / cyc);
    }
  return 0;
}
