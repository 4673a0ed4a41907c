// One-shot all-reduce (sum) of a small [M, hidden] tensor over NVLink / NVSwitch peer memory, for the row-parallel
// GEMM outputs of tensor-parallel decode.
//
// Replaces the reference's custom all-reduce for this path (csrc/custom_all_reduce.cuh:180-260 one/two stage kernels,
// :404-460 dispatch; vllm/distributed/device_communicators/custom_all_reduce.py:35-270) -- same job, different machine:
//
//   * PUSH, not pull.  Every rank owns a "region" (cudaMalloc'ed, exported with cudaIpcGetMemHandle, mapped by all
//     peers).  Rank r stores its 16-byte vectors straight into slot [parity][r] of EVERY rank's region (posted NVLink
//     writes, NVSwitch gives each pair full bandwidth), so the reduction afterwards reads LOCAL memory only.
//   * no grid-wide barrier: CTA b of rank r only depends on CTA b of the peers.  After its pushes it publishes
//     flag[parity][r][b] = epoch in every region (st.release.sys after a system fence), then waits for the W flags of
//     its own region (ld.acquire.sys) and sums the W slots in rank order with fp32 accumulation -- every rank computes
//     bit-identical results, independent of timing.
//   * CUDA-graph safe: the epoch lives in device memory (one counter per CTA index, bumped by the CTA itself), so a
//     replayed launch needs no new kernel arguments; slots are double-buffered by epoch parity, which is enough because
//     a rank can only start call n+2 after every peer has finished call n (its call n+1 needs their n+1 flags, which
//     they publish after their own call n has completed in stream order).
//   * programmatic dependent launch like every other kernel of the step.
#include "common.cuh"
#include "ar_fuse.cuh"
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <mutex>
#include <vector>

namespace b200 {

constexpr int kArMaxBlocks = 512;
constexpr int kArThreads = 256;

constexpr size_t kArLLMaxBytes = 1u << 20;   // payloads up to 1 MiB take the flag-in-data (LL) kernels
constexpr size_t kArLLSlotPayload = 2u << 20;   // LL slot capacity (payload bytes): room for the GEMM-fused exchange layout

struct ArLayout {
  // region = [flags: 2 x W x kArMaxBlocks u32][call counter u32, finished-CTA ticket u32][pad to 4096]
  //          [LL area: 2 parities x 2 phases x W x ll_slot][data: 2 x W x slot_bytes]
  static __host__ __device__ size_t flags_bytes(int world) { return (size_t)2 * world * kArMaxBlocks * 4; }
  static __host__ __device__ size_t ll_offset(int world) { return (flags_bytes(world) + 64 + 4095) / 4096 * 4096; }
  static __host__ __device__ size_t ll_slot(size_t slot_bytes) {
    return 2 * (slot_bytes < kArLLSlotPayload ? slot_bytes : kArLLSlotPayload);   // 8 payload bytes per 16-byte store
  }
  static __host__ __device__ size_t data_offset(int world, size_t slot_bytes) {
    return ll_offset(world) + (size_t)2 * 2 * world * ll_slot(slot_bytes);   // [parity][phase][rank]
  }
  static __host__ __device__ size_t region_bytes(int world, size_t slot_bytes) {
    return data_offset(world, slot_bytes) + (size_t)2 * world * slot_bytes;
  }
};

struct ArPtrs {
  uint8_t* region[kArMaxWorld];   // region[p]: rank p's region as mapped in THIS process (region[rank] is local)
};

struct ArCtx {
  int rank = 0, world = 0, device = 0;
  size_t slot_bytes = 0;
  ArPtrs ptrs;
  bool live = false;
  size_t two_shot_bytes = 256 * 1024;   // W > 2: payload x (W-1) above this goes two-shot (B200_AR_TWO_SHOT_BYTES)
  size_t fuse_max_bytes = 2u << 20;     // GEMM-fused one-shot exchange: payload x (W-1) limit (B200_AR_FUSE_MAX_BYTES)
  bool fuse_two_shot_always = false;    // B200_AR_FUSE_TWO_SHOT=1: two-shot exchange in the fused GEMM at any size (tests)
  bool no_ll = false;   // B200_AR_NO_LL=1: always the fence + flag kernel (bring-up / comparison)
};

static std::mutex g_ar_mu;
static std::vector<ArCtx> g_ar;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <typename T> struct ArVec;
template <> struct ArVec<__half> {
  static __device__ __forceinline__ void add(float* acc, const uint4& v) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __half22float2(h[i]); acc[2 * i] += f.x; acc[2 * i + 1] += f.y; }
  }
  static __device__ __forceinline__ uint4 pack(const float* acc) {
    uint4 r; __half2* h = reinterpret_cast<__half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
    return r;
  }
  static constexpr int N = 8;
};
template <> struct ArVec<__nv_bfloat16> {
  static __device__ __forceinline__ void add(float* acc, const uint4& v) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { acc[2 * i] += __uint_as_float(w[i] << 16); acc[2 * i + 1] += __uint_as_float(w[i] & 0xffff0000u); }
  }
  static __device__ __forceinline__ uint4 pack(const float* acc) {
    uint4 r; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
    return r;
  }
  static constexpr int N = 8;
};
template <> struct ArVec<float> {
  static __device__ __forceinline__ void add(float* acc, const uint4& v) {
    const float* f = reinterpret_cast<const float*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] += f[i];
  }
  static __device__ __forceinline__ uint4 pack(const float* acc) {
    uint4 r; float* f = reinterpret_cast<float*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = acc[i];
    return r;
  }
  static constexpr int N = 4;
};

__device__ __forceinline__ void ar_finish_call(volatile uint32_t* ctr, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x == 0) ar_ticket(ctr, epoch);
}

// in/out: nvec 16-byte vectors (out may alias in).  grid <= kArMaxBlocks; CTA b owns vectors b*T + t + i*grid*T.
template <typename T, int W>
__global__ void __launch_bounds__(kArThreads)
ar_oneshot_kernel(const ArPtrs p, const int rank, const size_t slot_bytes, const uint4* in, uint4* out,
                  const int nvec) {
  pdl_wait();                // `in` is the output of the preceding GEMM
  pdl_launch_dependents();
  const int b = blockIdx.x, t = threadIdx.x;
  const size_t data_off = ArLayout::data_offset(W, slot_bytes);
  // static indexing only (a runtime index would move the pointer table to local memory)
  uint8_t* self = nullptr;
  uint8_t* peer_t = nullptr;   // region of rank t (threads t < W)
#pragma unroll
  for (int q = 0; q < W; ++q) {
    if (q == rank) self = p.region[q];
    if (q == t) peer_t = p.region[q];
  }
  // one epoch per CALL, shared by all CTAs (the parity picks the slot set, so it must not depend on the CTA index)
  volatile uint32_t* ctr = reinterpret_cast<volatile uint32_t*>(self + ArLayout::flags_bytes(W));
  const uint32_t epoch = ctr[0] + 1;
  const uint32_t par = epoch & 1u;
  const size_t my_slot = data_off + ((size_t)par * W + rank) * slot_bytes;
  const int stride = gridDim.x * kArThreads;
  // ---- push my vectors into slot [par][rank] of every region ------------------------------------------------------
  for (int v = b * kArThreads + t; v < nvec; v += stride) {
    const uint4 x = in[v];
#pragma unroll
    for (int q = 0; q < W; ++q) reinterpret_cast<uint4*>(p.region[q] + my_slot)[v] = x;
  }
  __syncthreads();
  // ---- publish / wait: thread q talks to rank q ------------------------------------------------------------------
  if (t < W) {
    __threadfence_system();
    uint32_t* theirs = reinterpret_cast<uint32_t*>(peer_t) + ((size_t)par * W + rank) * kArMaxBlocks + b;
    st_release_sys(theirs, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(self) + ((size_t)par * W + t) * kArMaxBlocks + b;
    while (ld_acquire_sys(mine) != epoch) { }
  }
  __syncthreads();
  // ---- reduce the W local slots in rank order --------------------------------------------------------------------
  const uint8_t* base = self + data_off + (size_t)par * W * slot_bytes;
  for (int v = b * kArThreads + t; v < nvec; v += stride) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    uint4 x[W];
#pragma unroll
    for (int q = 0; q < W; ++q) x[q] = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)q * slot_bytes) + v);
#pragma unroll
    for (int q = 0; q < W; ++q) ArVec<T>::add(acc, x[q]);
    out[v] = ArVec<T>::pack(acc);
  }
  ar_finish_call(ctr, epoch);
}

// ---- flag-in-data ("LL") variant for payloads <= kArLLMaxBytes ---------------------------------------------------
// Every 16-byte store carries 8 payload bytes and the call epoch twice: {d0, epoch, d1, epoch}.  The receiver polls the
// data itself until both epochs match (only 8-byte store atomicity is assumed), so there is no fence and no separate
// flag round trip: the latency of a call is ONE NVLink traversal.  A rank's own contribution stays in registers.
template <typename T, int W>
__global__ void __launch_bounds__(kArThreads)
ar_ll_kernel(const ArPtrs p, const int rank, const size_t slot_bytes, const uint4* in, uint4* out, const int nvec) {
  pdl_wait();
  pdl_launch_dependents();
  const int b = blockIdx.x, t = threadIdx.x;
  uint8_t* self = nullptr;
#pragma unroll
  for (int q = 0; q < W; ++q)
    if (q == rank) self = p.region[q];
  volatile uint32_t* ctr = reinterpret_cast<volatile uint32_t*>(self + ArLayout::flags_bytes(W));
  const uint32_t epoch = ctr[0] + 1;
  const uint32_t par = epoch & 1u;
  const size_t ll_slot = ArLayout::ll_slot(slot_bytes);
  const size_t area = ArLayout::ll_offset(W) + (size_t)par * 2 * W * ll_slot;
  const size_t my_off = area + (size_t)rank * ll_slot;
  const int stride = gridDim.x * kArThreads;
  for (int v = b * kArThreads + t; v < nvec; v += stride) {
    const uint4 x = in[v];
#pragma unroll
    for (int q = 0; q < W; ++q) {
      if (q != rank) {
        uint8_t* d = p.region[q] + my_off + (size_t)v * 32;
        st_ll(d, x.x, x.y, epoch);
        st_ll(d + 16, x.z, x.w, epoch);
      }
    }
  }
  for (int v = b * kArThreads + t; v < nvec; v += stride) {
    const uint4 x = in[v];
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < W; ++q) {   // rank order: bit-identical sums on every rank
      uint4 y = x;
      if (q != rank) {
        const uint8_t* src = self + area + (size_t)q * ll_slot + (size_t)v * 32;
        uint4 lo, hi;
        do {
          lo = ld_ll(src);
          hi = ld_ll(src + 16);
        } while (lo.y != epoch || lo.w != epoch || hi.y != epoch || hi.w != epoch);
        y = make_uint4(lo.x, lo.z, hi.x, hi.z);
      }
      ArVec<T>::add(acc, y);
    }
    out[v] = ArVec<T>::pack(acc);
  }
  ar_finish_call(ctr, epoch);
}

// ---- two-shot LL variant (W > 2, larger payloads): reduce-scatter + all-gather inside one kernel ----------------------
// One-shot moves (W-1) x payload per rank; here rank j owns the j-th 1/W of the vectors: everyone pushes its copy of
// chunk j to rank j (phase area 0), rank j sums the W copies in rank order and pushes the result to everyone (phase
// area 1).  2 x (W-1)/W x payload on the wire per rank and two NVLink traversals; results are still bit-identical on
// all ranks (one rank computes each element).
__device__ __forceinline__ uint4 poll_ll(const uint8_t* src, uint32_t epoch) {
  uint4 lo, hi;
  do {
    lo = ld_ll(src);
    hi = ld_ll(src + 16);
  } while (lo.y != epoch || lo.w != epoch || hi.y != epoch || hi.w != epoch);
  return make_uint4(lo.x, lo.z, hi.x, hi.z);
}

template <typename T, int W>
__global__ void __launch_bounds__(kArThreads)
ar_ll2_kernel(const ArPtrs p, const int rank, const size_t slot_bytes, const uint4* in, uint4* out, const int nvec) {
  pdl_wait();
  pdl_launch_dependents();
  const int b = blockIdx.x, t = threadIdx.x;
  uint8_t* self = nullptr;
#pragma unroll
  for (int q = 0; q < W; ++q)
    if (q == rank) self = p.region[q];
  volatile uint32_t* ctr = reinterpret_cast<volatile uint32_t*>(self + ArLayout::flags_bytes(W));
  const uint32_t epoch = ctr[0] + 1;
  const uint32_t par = epoch & 1u;
  const size_t ll_slot = ArLayout::ll_slot(slot_bytes);
  const size_t area1 = ArLayout::ll_offset(W) + (size_t)par * 2 * W * ll_slot;
  const size_t area2 = area1 + (size_t)W * ll_slot;
  const int c = (nvec + W - 1) / W;   // vectors per owner
  const int stride = gridDim.x * kArThreads;
  // phase 1: my copy of chunk q -> rank q
  for (int i = b * kArThreads + t; i < c; i += stride) {
#pragma unroll
    for (int q = 0; q < W; ++q) {
      const int v = q * c + i;
      if (q != rank && v < nvec) {
        const uint4 x = in[v];
        uint8_t* d = p.region[q] + area1 + (size_t)rank * ll_slot + (size_t)i * 32;
        st_ll(d, x.x, x.y, epoch);
        st_ll(d + 16, x.z, x.w, epoch);
      }
    }
  }
  // phase 2: reduce my chunk in rank order, publish the result
  for (int i = b * kArThreads + t; i < c; i += stride) {
    const int v = rank * c + i;
    if (v < nvec) {
      const uint4 x = in[v];
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int q = 0; q < W; ++q) {
        uint4 y = x;
        if (q != rank) y = poll_ll(self + area1 + (size_t)q * ll_slot + (size_t)i * 32, epoch);
        ArVec<T>::add(acc, y);
      }
      const uint4 r = ArVec<T>::pack(acc);
      out[v] = r;
#pragma unroll
      for (int q = 0; q < W; ++q) {
        if (q != rank) {
          uint8_t* d = p.region[q] + area2 + (size_t)rank * ll_slot + (size_t)i * 32;
          st_ll(d, r.x, r.y, epoch);
          st_ll(d + 16, r.z, r.w, epoch);
        }
      }
    }
  }
  // phase 3: collect the other owners' results
  for (int i = b * kArThreads + t; i < c; i += stride) {
#pragma unroll
    for (int q = 0; q < W; ++q) {
      const int v = q * c + i;
      if (q != rank && v < nvec) out[v] = poll_ll(self + area2 + (size_t)q * ll_slot + (size_t)i * 32, epoch);
    }
  }
  ar_finish_call(ctr, epoch);
}

template <typename T>
static cudaError_t ar_launch(const ArCtx& c, const void* in, void* out, int nvec, cudaStream_t st) {
  const bool ll = (size_t)nvec * 16 <= kArLLMaxBytes && !c.no_ll;
  const bool two_shot = ll && c.world > 2 && (size_t)nvec * 16 * (c.world - 1) > c.two_shot_bytes;
  const int work = two_shot ? (nvec + c.world - 1) / c.world : nvec;   // vectors per thread-grid pass
  int grid = (work + kArThreads - 1) / kArThreads;
  if (grid > kArMaxBlocks) grid = kArMaxBlocks;
  if (grid < 1) grid = 1;
#define AR_CASE(W)                                                                                              \
  case W:                                                                                                       \
    if (two_shot)                                                                                                \
      return launch_pdl(ar_ll2_kernel<T, W>, dim3(grid), dim3(kArThreads), 0, st, c.ptrs, c.rank, c.slot_bytes,  \
                        (const uint4*)in, (uint4*)out, nvec);                                                   \
    return ll ? launch_pdl(ar_ll_kernel<T, W>, dim3(grid), dim3(kArThreads), 0, st, c.ptrs, c.rank, c.slot_bytes, \
                           (const uint4*)in, (uint4*)out, nvec)                                                 \
              : launch_pdl(ar_oneshot_kernel<T, W>, dim3(grid), dim3(kArThreads), 0, st, c.ptrs, c.rank,          \
                           c.slot_bytes, (const uint4*)in, (uint4*)out, nvec);
  switch (c.world) {
    AR_CASE(2) AR_CASE(4) AR_CASE(6) AR_CASE(8)
    default: return cudaErrorInvalidValue;
  }
#undef AR_CASE
}


static thread_local long long g_pending_ctx = 0, g_pending_m = 0, g_pending_n = 0;
static thread_local bool g_pending_two_shot = false;

// rows of 16-byte LL cells per output channel the fused GEMM epilogue may use (marlin_gemm_kernel.cuh: MPAD/4 groups,
// plus up to one partly filled group per cluster rank)
static size_t fused_cells(long long m) {
  const long long mpad = m <= 16 ? 16 : m <= 32 ? 32 : m <= 64 ? 64 : 128;
  return (size_t)(mpad / 4 + 8);
}

bool ar_take_pending(ArFuse* out, long long* m_out, long long* n_out) {
  if (g_pending_ctx == 0) return false;
  std::lock_guard<std::mutex> g(g_ar_mu);
  const ArCtx& c = g_ar[g_pending_ctx - 1];
  for (int i = 0; i < kArMaxWorld; ++i) out->region[i] = c.ptrs.region[i];
  out->rank = c.rank;
  out->world = c.world;
  out->ll_off = ArLayout::ll_offset(c.world);
  out->ll_slot = ArLayout::ll_slot(c.slot_bytes);
  out->ctr_off = ArLayout::flags_bytes(c.world);
  out->two_shot = g_pending_two_shot ? 1 : 0;
  *m_out = g_pending_m;
  *n_out = g_pending_n;
  g_pending_ctx = 0;
  return true;
}

}  // namespace b200

using namespace b200;

extern "C" {

int b200_ar_region_bytes(int world, int64_t max_bytes, int64_t* bytes_out) {
  if (world < 2 || world > kArMaxWorld || max_bytes <= 0 || !bytes_out)
    return set_error(B200_EINVAL, "ar_region_bytes: world %d / max_bytes %lld", world, (long long)max_bytes);
  const size_t slot = ((size_t)max_bytes + 4095) / 4096 * 4096;
  *bytes_out = (int64_t)ArLayout::region_bytes(world, slot);
  return B200_OK;
}

int b200_ar_alloc(int64_t bytes, void** ptr_out, void* handle_out64) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "ar_alloc: cudaMalloc(%lld): %s", (long long)bytes, cudaGetErrorString(e));
  e = cudaMemset(p, 0, (size_t)bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return set_error(B200_ECUDA, "ar_alloc: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  memcpy(handle_out64, &h, 64);
  *ptr_out = p;
  return B200_OK;
}

int b200_ar_open(const void* handle64, void** ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "ar_open: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
  *ptr_out = p;
  return B200_OK;
}

int b200_ar_close(void* ptr) {
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  return e == cudaSuccess ? B200_OK : set_error(B200_ECUDA, "ar_close: %s", cudaGetErrorString(e));
}

int b200_ar_free(void* ptr) {
  cudaError_t e = cudaFree(ptr);
  return e == cudaSuccess ? B200_OK : set_error(B200_ECUDA, "ar_free: %s", cudaGetErrorString(e));
}

int b200_ar_init(int rank, int world, void* const* region_ptrs, int64_t max_bytes, int64_t* ctx_out) {
  if (!(world == 2 || world == 4 || world == 6 || world == 8) || rank < 0 || rank >= world || max_bytes <= 0)
    return set_error(B200_EINVAL, "ar_init: rank %d world %d max_bytes %lld (world must be 2, 4, 6 or 8)", rank, world,
                     (long long)max_bytes);
  ArCtx c;
  c.rank = rank; c.world = world;
  cudaGetDevice(&c.device);
  c.slot_bytes = ((size_t)max_bytes + 4095) / 4096 * 4096;
  for (int i = 0; i < kArMaxWorld; ++i) c.ptrs.region[i] = i < world ? (uint8_t*)region_ptrs[i] : nullptr;
  for (int i = 0; i < world; ++i)
    if (!c.ptrs.region[i]) return set_error(B200_EINVAL, "ar_init: region pointer %d is null", i);
  c.live = true;
  { const char* e = getenv("B200_AR_NO_LL"); c.no_ll = e && atoi(e) != 0; }
  { const char* e = getenv("B200_AR_FUSE_TWO_SHOT"); c.fuse_two_shot_always = e && atoi(e) != 0; }
  { const char* e = getenv("B200_AR_FUSE_MAX_BYTES"); if (e && atoll(e) >= 0) c.fuse_max_bytes = (size_t)atoll(e); }
  { const char* e = getenv("B200_AR_TWO_SHOT_BYTES"); if (e && atoll(e) >= 0) c.two_shot_bytes = (size_t)atoll(e); }
  std::lock_guard<std::mutex> g(g_ar_mu);
  g_ar.push_back(c);
  *ctx_out = (int64_t)g_ar.size();   // 1-based handle
  return B200_OK;
}

/* The NEXT b200_gptq_marlin_gemm issued by this host thread ([m, n] output) exchanges its output tiles with the peers
 * inside its epilogue and writes the all-reduced result: one kernel instead of GEMM + all-reduce.  B200_ENOTSUP when
 * the shape does not fit the exchange area (the caller then runs GEMM and b200_ar_all_reduce separately). */
int b200_ar_fuse_next_gemm(int64_t ctx, int64_t m, int64_t n) {
  std::lock_guard<std::mutex> g(g_ar_mu);
  if (ctx < 1 || ctx > (int64_t)g_ar.size() || !g_ar[ctx - 1].live) return set_error(B200_EINVAL, "ar_fuse_next_gemm: bad handle");
  const ArCtx& c = g_ar[ctx - 1];
  const size_t npad = (size_t)((n + 127) / 128) * 128;
  if (m < 1 || m > 128 || fused_cells(m) * npad * 16 > ArLayout::ll_slot(c.slot_bytes))
    return set_error(B200_ENOTSUP, "ar_fuse_next_gemm: [%lld, %lld] does not fit the exchange area", (long long)m, (long long)n);
  // The fused exchange is one-shot (every rank sends its whole output to every peer) and only taken while
  // payload x (W-1) stays small.  Measured on 8 x B200, Llama-3-8B batch 64 (profiles/r1_bench_tp8_*.json): above that
  // GEMM + the stand-alone two-shot kernel (2.37 ms/step) beats both a one-shot (2.95) and a two-shot (2.59) exchange
  // inside the epilogue -- the few epilogue warps of 32 tiles poll serially where the stand-alone kernel has 128 CTAs.
  // B200_AR_FUSE_TWO_SHOT=1 forces the two-shot epilogue exchange (kept for tests and further tuning).
  g_pending_two_shot = c.fuse_two_shot_always;
  if (!g_pending_two_shot && (size_t)m * n * 2 * (c.world - 1) > c.fuse_max_bytes)
    return set_error(B200_ENOTSUP, "ar_fuse_next_gemm: payload too large for the one-shot exchange at world %d", c.world);
  g_pending_ctx = ctx; g_pending_m = m; g_pending_n = n;
  return B200_OK;
}

int b200_ar_dispose(int64_t ctx) {
  std::lock_guard<std::mutex> g(g_ar_mu);
  if (ctx < 1 || ctx > (int64_t)g_ar.size()) return set_error(B200_EINVAL, "ar_dispose: bad handle");
  g_ar[ctx - 1].live = false;
  return B200_OK;
}

int b200_ar_all_reduce(int64_t ctx, const void* in, void* out, int64_t numel, int dtype, void* stream) {
  ArCtx c;
  {
    std::lock_guard<std::mutex> g(g_ar_mu);
    if (ctx < 1 || ctx > (int64_t)g_ar.size() || !g_ar[ctx - 1].live) return set_error(B200_EINVAL, "ar_all_reduce: bad handle");
    c = g_ar[ctx - 1];
  }
  const int es = dtype == B200_F32 ? 4 : 2;
  const int64_t bytes = numel * es;
  if (numel <= 0) return B200_OK;
  if (bytes % 16 != 0 || ((uintptr_t)in & 15) || ((uintptr_t)out & 15))
    return set_error(B200_EINVAL, "ar_all_reduce: byte size and pointers must be multiples of 16 (reference: custom_all_reduce.cu should_custom_ar)");
  if ((size_t)bytes > c.slot_bytes)
    return set_error(B200_ENOTSUP, "ar_all_reduce: %lld bytes exceed the registered maximum %lld", (long long)bytes, (long long)c.slot_bytes);
  const int nvec = (int)(bytes / 16);
  cudaError_t e;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_F16) e = ar_launch<__half>(c, in, out, nvec, st);
  else if (dtype == B200_BF16) e = ar_launch<__nv_bfloat16>(c, in, out, nvec, st);
  else if (dtype == B200_F32) e = ar_launch<float>(c, in, out, nvec, st);
  else return set_error(B200_EINVAL, "ar_all_reduce: dtype %d", dtype);
  if (e != cudaSuccess) return set_error(B200_ECUDA, "ar_all_reduce: %s", cudaGetErrorString(e));
  return B200_OK;
}

}  // extern "C"
