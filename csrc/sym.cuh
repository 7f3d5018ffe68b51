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
// grid-wide exchanges: arrival tickets and "totals are published" markers, both local to the rank
__device__ __forceinline__ uint32_t* ticket_ptr(const PmdSym& s, int slot) {
  return reinterpret_cast<uint32_t*>(s.d_bases[s.rank] + PMD_FLAGS_BYTES) + PMD_MAX_SLOTS + slot;
}
__device__ __forceinline__ uint32_t* done_ptr(const PmdSym& s, int slot) {
  return reinterpret_cast<uint32_t*>(s.d_bases[s.rank] + PMD_FLAGS_BYTES) + 2 * PMD_MAX_SLOTS + slot;
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
__device__ __forceinline__ float poll_ll(const uint2* p, uint32_t epoch, int rank, int src, int slot) {
  uint32_t v, f;
  const unsigned long long t0 = globaltimer_ns();
  uint32_t spins = 0;
  for (;;) {
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(f) : "l"(p) : "memory");
    if (f == epoch) break;
    if (((++spins) & 0xfff) == 0 && globaltimer_ns() - t0 > 30000000000ULL) {
      printf("pmdist: rank %d timed out waiting for rank %d's data (slot %d, epoch %u, saw %u)\n", rank, src, slot, epoch,
             f);
      __trap();
    }
  }
  return __uint_as_float(v);
}

// Every thread of the (single) block calls this; returns with sum_s[0..n) = rank-ordered total in shared memory.
__device__ void ll_exchange_sum(const PmdSym& s, int slot, const float* local, int n, unsigned long long recv_off,
                                float* sum_s) {
  __shared__ uint32_t epoch_s;
  if (threadIdx.x == 0) {
    uint32_t* ep = epoch_ptr(s, slot);
    const uint32_t e = *ep + 1;
    *ep = e;
    epoch_s = e;
  }
  __syncthreads();
  const uint32_t epoch = epoch_s;
  const int parity = epoch & 1;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = local[i];
    for (int w = 0; w < s.world; ++w) {
      if (w == s.rank) continue;
      uint2* dst = heap_ptr<uint2>(s, w, recv_off) + (static_cast<long long>(parity) * s.world + s.rank) * n + i;
      st_ll(dst, v, epoch);
    }
  }
  const uint2* mine = heap_ptr<const uint2>(s, s.rank, recv_off) + static_cast<long long>(parity) * s.world * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float acc = 0.f;
    for (int w = 0; w < s.world; ++w)      // rank order: identical bits on every rank
      acc += (w == s.rank) ? local[i] : poll_ll(mine + static_cast<long long>(w) * n + i, epoch, s.rank, w, slot);
    sum_s[i] = acc;
  }
  __syncthreads();
}


// Grid-wide form for kernels with many blocks (the fused BatchNorm kernels): every block needs the totals, but
// only block 0 talks to the peers.  All blocks agree on the epoch through an arrival ticket (each block reads the
// epoch BEFORE taking its ticket; the last arrival advances it, so a launch never sees its own increment);
// block 0 pushes, polls, writes the totals to `totals_gmem` and publishes done[slot] = epoch; the other blocks
// wait for that marker and read the totals from local memory.  Returns with sum_s[0..n) filled on every block.
__device__ void ll_exchange_sum_grid(const PmdSym& s, int slot, const float* local, int n, unsigned long long recv_off,
                                     float* totals_gmem, float* sum_s) {
  __shared__ uint32_t epoch_g;
  if (threadIdx.x == 0) {
    volatile uint32_t* ep = epoch_ptr(s, slot);
    const uint32_t e0 = *ep;
    uint32_t* tk = ticket_ptr(s, slot);
    const uint32_t t = atomicAdd(tk, 1u);
    if (t == gridDim.x - 1) {
      *tk = 0;
      __threadfence();
      *ep = e0 + 1;
    }
    epoch_g = e0 + 1;
  }
  __syncthreads();
  const uint32_t epoch = epoch_g;
  if (blockIdx.x == 0) {
    const int parity = epoch & 1;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float v = local[i];
      for (int w = 0; w < s.world; ++w) {
        if (w == s.rank) continue;
        st_ll(heap_ptr<uint2>(s, w, recv_off) + (static_cast<long long>(parity) * s.world + s.rank) * n + i, v, epoch);
      }
    }
    const uint2* mine = heap_ptr<const uint2>(s, s.rank, recv_off) + static_cast<long long>(parity) * s.world * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float acc = 0.f;
      for (int w = 0; w < s.world; ++w)
        acc += (w == s.rank) ? local[i] : poll_ll(mine + static_cast<long long>(w) * n + i, epoch, s.rank, w, slot);
      sum_s[i] = acc;
      totals_gmem[i] = acc;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(done_ptr(s, slot)), "r"(epoch) : "memory");
  } else {
    if (threadIdx.x == 0) {
      const uint32_t* dp = done_ptr(s, slot);
      const unsigned long long t0 = globaltimer_ns();
      uint32_t v, spins = 0;
      for (;;) {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(dp) : "memory");
        if (v == epoch) break;
        if (((++spins) & 0xfff) == 0 && globaltimer_ns() - t0 > 30000000000ULL) {
          printf("pmdist: rank %d block %d timed out waiting for the exchanged totals (slot %d)\n", s.rank, blockIdx.x, slot);
          __trap();
        }
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) sum_s[i] = __ldcg(totals_gmem + i);
  }
  __syncthreads();
}

}  // namespace
}  // namespace pmd
