// Shared between all_reduce.cu and the GEMM epilogue: the flag-in-data ("LL") exchange over NVLink peer memory.
// A 16-byte store carries 8 payload bytes and the call epoch twice, {d0, epoch, d1, epoch}; the receiver polls the data
// itself until both epochs match (8-byte store atomicity is all that is assumed), so no fence and no separate flag.
#pragma once
#include <stdint.h>

namespace b200 {

constexpr int kArMaxWorld = 8;

// what a kernel needs to take part in an exchange (filled by all_reduce.cu from a context)
struct ArFuse {
  uint8_t* region[kArMaxWorld];   // rank p's region as mapped in this process
  int rank, world;                // world == 0: no exchange
  unsigned long long ll_off;      // offset of the LL area inside a region
  unsigned long long ll_slot;     // bytes of one [parity][phase][rank] LL slot
  unsigned long long ctr_off;     // offset of {call counter, finished-CTA ticket}
  int two_shot;                   // fused GEMM exchange: 0 = every rank sums everything, 1 = the cell's owner rank sums and broadcasts
};

__device__ __forceinline__ void st_ll(uint8_t* p, uint32_t d0, uint32_t d1, uint32_t e) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(d0), "r"(e), "r"(d1), "r"(e) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint8_t* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// the last CTA of a call to get here bumps the call counter: the next call (which cannot pass griddepcontrol.wait /
// stream order before this kernel has completed) reads epoch + 1.  ctr[0] = epoch of the last finished call,
// ctr[1] = finished-CTA ticket.  Call with ONE thread per CTA after the CTA's last exchange.
__device__ __forceinline__ void ar_ticket(volatile uint32_t* ctr, uint32_t epoch) {
  __threadfence();
  const uint32_t ticket = atomicAdd(const_cast<uint32_t*>(ctr) + 1, 1u);
  if (ticket == gridDim.x * gridDim.y * gridDim.z - 1) {
    ctr[1] = 0;
    ctr[0] = epoch;
  }
}

// host side (all_reduce.cu): consume the "fuse the all-reduce into the next GEMM" request of this thread, if any
bool ar_take_pending(ArFuse* out, long long* m_out, long long* n_out);

}  // namespace b200
