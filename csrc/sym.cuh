// Device-side helpers for peer-mapped ("symmetric") memory: system-scope flag loads/stores, the signal-pad
// layout, and the push ("LL") exchange of small vectors.  Shared by comm.cu (collectives) and bn.cu (the
// SyncBatchNorm exchange fused into the BatchNorm kernels).  See include/pmdist_comm.h for the protocol.
#pragma once

#include "common.cuh"
#include "../include/pmdist_comm.h"

namespace pmd {
namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float ld_peer_f(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float4* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_st_f2(float2* p, float2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint32_t* flag_ptr(const PmdSym& s, int heap_rank, int slot, int src) {
  return reinterpret_cast<uint32_t*>(s.d_bases[heap_rank]) + slot * PMD_MAX_RANKS + src;
}
__device__ __forceinline__ uint32_t* epoch_ptr(const PmdSym& s, int slot) {
  return reinterpret_cast<uint32_t*>(s.d_bases[s.rank] + PMD_FLAGS_BYTES) + slot;
}
template <typename T>
__device__ __forceinline__ T* heap_ptr(const PmdSym& s, int w, unsigned long long off) {
  return reinterpret_cast<T*>(s.d_bases[w] + off);
}

// ------------------------------------------------------------------ push ("LL") exchange for tiny vectors
// The flag-barrier + peer-load exchange above costs two NVLink round trips.  For the <= 16 KB SyncBatchNorm
// vectors a push protocol needs one one-way trip: every rank STORES its n floats into every peer's receive
// area as 8-byte words {value, epoch}; a word is valid exactly when its epoch half matches, so data and flag
// arrive in one transaction and receivers poll local memory only.  Receive area (per exchange, in every
// heap): uint2 [2 parities][world][n]; the parity alternates per use, so a fast rank that is already one use
// ahead never overwrites words a slow rank is still waiting for.
__device__ __forceinline__ void st_ll(uint2* p, float v, uint32_t epoch) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(epoch) : "memory");
}
// Polls the words of element i from every peer's row of the local receive area, all peers in flight at once (a
// dependent load-compare loop per peer would serialise W L2 round trips), and returns the rank-ordered sum.
__device__ __forceinline__ float ll_poll_sum_one(const PmdSym& s, const uint2* mine, int n, int i, float own,
                                                 uint32_t epoch, int slot) {
  uint32_t v[PMD_MAX_RANKS], f[PMD_MAX_RANKS];
  uint32_t pending = ((1u << s.world) - 1u) & ~(1u << s.rank);
  const unsigned long long t0 = globaltimer_ns();
  uint32_t spins = 0;
  while (pending != 0u) {
#pragma unroll
    for (int w = 0; w < PMD_MAX_RANKS; ++w)
      if ((pending >> w) & 1u)
        asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];"
                     : "=r"(v[w]), "=r"(f[w])
                     : "l"(mine + static_cast<long long>(w) * n + i)
                     : "memory");
#pragma unroll
    for (int w = 0; w < PMD_MAX_RANKS; ++w)
      if (((pending >> w) & 1u) && f[w] == epoch) pending &= ~(1u << w);
    if (pending != 0u && ((++spins) & 0xfff) == 0 && globaltimer_ns() - t0 > 30000000000ULL) {
      printf("pmdist: rank %d timed out waiting for pushed data (slot %d, epoch %u, missing ranks mask 0x%x)\n", s.rank,
             slot, epoch, pending);
      __trap();
    }
  }
  float acc = 0.f;
#pragma unroll
  for (int w = 0; w < PMD_MAX_RANKS; ++w)      // rank order: identical bits on every rank
    if (w < s.world) acc += (w == s.rank) ? own : __uint_as_float(v[w]);
  return acc;
}

// The three phases of a push exchange, for ONE block (all its threads call each phase):
//   ll_next_epoch  advances the exchange's device-resident epoch counter and returns the new value to every thread;
//   ll_push        stores this rank's vector into every peer's receive area as {value, epoch} words;
//   ll_poll_sum    waits for every peer's words in the LOCAL receive area and leaves the rank-ordered totals in sum_s.
// `local` is read with ld.cg: it may have been accumulated by red.global.add from other SMs just before.
__device__ __forceinline__ uint32_t ll_next_epoch(const PmdSym& s, int slot, uint32_t* bcast_s) {
  if (threadIdx.x == 0) {
    volatile uint32_t* ep = epoch_ptr(s, slot);
    const uint32_t e = *ep + 1;
    *ep = e;
    *bcast_s = e;
  }
  __syncthreads();
  return *bcast_s;
}
__device__ __forceinline__ void ll_push(const PmdSym& s, const float* local, int n, unsigned long long recv_off,
                                        uint32_t epoch) {
  const int parity = epoch & 1;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = __ldcg(local + i);
    for (int w = 0; w < s.world; ++w) {
      if (w == s.rank) continue;
      st_ll(heap_ptr<uint2>(s, w, recv_off) + (static_cast<long long>(parity) * s.world + s.rank) * n + i, v, epoch);
    }
  }
}
__device__ __forceinline__ void ll_poll_sum(const PmdSym& s, int slot, const float* local, int n,
                                            unsigned long long recv_off, uint32_t epoch, float* sum_s) {
  const uint2* mine = heap_ptr<const uint2>(s, s.rank, recv_off) + static_cast<long long>(epoch & 1) * s.world * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) sum_s[i] = ll_poll_sum_one(s, mine, n, i, __ldcg(local + i), epoch, slot);
}

// Every thread of the (single) block calls this; returns with sum_s[0..n) = rank-ordered total in shared memory.
__device__ __forceinline__ void ll_exchange_sum(const PmdSym& s, int slot, const float* local, int n,
                                                unsigned long long recv_off, float* sum_s) {
  __shared__ uint32_t epoch_s;
  const uint32_t epoch = ll_next_epoch(s, slot, &epoch_s);
  ll_push(s, local, n, recv_off, epoch);
  ll_poll_sum(s, slot, local, n, recv_off, epoch, sum_s);
  __syncthreads();
}

}  // namespace
}  // namespace pmd
