// SyncBatchNorm exchange completed by the PRODUCER of the partial sums.
//
// torch's SyncBatchNorm puts a blocking NCCL collective between the kernel that reduces per-channel partials and
// the kernel that consumes the global statistics (ref:main.py:43; torch/nn/modules/_functions.py:65-79, 155-165;
// SURVEY.md C3/C4, K6, K8): 24 per training step for the reference's ResNet18.  Round 1 of this framework ran a
// single-block exchange kernel in that place.  Here no kernel sits there at all: the kernel that accumulates the
// per-channel vector (a convolution's epilogue in the forward pass, the BatchNorm-backward reduction or the dgrad
// that carries it in the backward pass) takes an arrival ticket per CTA, and the LAST CTA to arrive
//   1. reads the finished local vector, stores it into every peer's receive area ({value, epoch} words, one one-way
//      NVLink trip, see sym.cuh),
//   2. polls its own receive area for the peers' vectors and sums them in rank order (bitwise identical everywhere),
//   3. forward: derives scale/shift/mean/invstd and updates the running statistics; backward: writes the totals.
// The consumer is then the ordinary single-GPU kernel reading C-sized vectors from local memory; with programmatic
// dependent launch its prologue overlaps the NVLink trip.
#pragma once

#include "../include/pmdist_comm.h"

namespace pmd {

struct SyncTail {
  int mode;            // 0: off, 1: forward ([sum, sumsq] -> BatchNorm coefficients), 2: backward (sums -> totals)
  int slot;            // signal-pad slot owning this exchange's epoch counter and arrival ticket
  PmdSym sym;
  unsigned long long recv_off;   // receive area, PMD_LL_RECV_BYTES(world, 2C), same offset in every heap
  int C;
  float count_per_rank, momentum, eps;                  // mode 1
  const float *gamma, *beta;
  float *running_mean, *running_var;
  long long* nbt;
  float *scale, *shift, *mean, *invstd;
  float* totals;                                        // mode 2: [2C] sums over all ranks
};

}  // namespace pmd

#ifdef __CUDACC__
#include "sym.cuh"

namespace pmd {
namespace {

// What every BatchNorm finalize in this library computes for channel c from the global [sum, sumsq].
__device__ __forceinline__ void bn_finalize_channel(int c, float s1, float s2, float n, const float* gamma,
                                                    const float* beta, float* running_mean, float* running_var,
                                                    float momentum, float eps, float* scale, float* shift,
                                                    float* save_mean, float* save_invstd) {
  const float mean = s1 / n;
  const float var = fmaxf(s2 / n - mean * mean, 0.f);
  const float invstd = rsqrtf(var + eps);
  const float a = gamma[c] * invstd;
  scale[c] = a;
  shift[c] = beta[c] - mean * a;
  save_mean[c] = mean;
  save_invstd[c] = invstd;
  if (running_mean != nullptr) {
    const float unbiased = var * (n / fmaxf(n - 1.f, 1.f));
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// Called by ALL threads of a CTA after the CTA's last red.global.add into the vector(s) being exchanged.
// Returns true (to every thread) in exactly one CTA of the grid: the last one to arrive.  The ticket is re-armed
// for the next launch by that CTA (nobody else touches it until then), so CUDA-graph replays need no reset.
__device__ __forceinline__ bool grid_last_arrival(uint32_t* ticket, uint32_t total_ctas) {
  __shared__ uint32_t last_s;
  __threadfence();                 // this thread's reductions are ordered before the ticket
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t t = atomicAdd(ticket, 1u);
    const bool last = t == total_ctas - 1;
    if (last) *ticket = 0;
    last_s = last ? 1u : 0u;
    __threadfence();               // the last arrival reads the others' reductions after this point
  }
  __syncthreads();
  return last_s != 0u;
}

// One or two exchanges completed by the last CTA of the grid.  `vec0/vec1`: the local [2C] vectors; `scratch`:
// shared memory for (2 * t0.C + 2 * t1.C) floats that nothing else uses any more.  All threads of every CTA call it.
__device__ __noinline__ void sync_tail_run(const SyncTail& t0, const float* vec0, const SyncTail* t1, const float* vec1,
                                           float* scratch, uint32_t total_ctas) {
  if (!grid_last_arrival(ticket_ptr(t0.sym, t0.slot), total_ctas)) return;
  __shared__ uint32_t ep_s[2];
  const int n0 = 2 * t0.C, n1 = t1 != nullptr ? 2 * t1->C : 0;
  const uint32_t e0 = ll_next_epoch(t0.sym, t0.slot, &ep_s[0]);
  const uint32_t e1 = t1 != nullptr ? ll_next_epoch(t1->sym, t1->slot, &ep_s[1]) : 0u;
  ll_push(t0.sym, vec0, n0, t0.recv_off, e0);                       // both vectors leave before either is awaited:
  if (t1 != nullptr) ll_push(t1->sym, vec1, n1, t1->recv_off, e1);  // one NVLink trip for a projection block's pair
  ll_poll_sum(t0.sym, t0.slot, vec0, n0, t0.recv_off, e0, scratch);
  if (t1 != nullptr) ll_poll_sum(t1->sym, t1->slot, vec1, n1, t1->recv_off, e1, scratch + n0);
  __syncthreads();
#pragma unroll 1
  for (int b = 0; b < (t1 != nullptr ? 2 : 1); ++b) {
    const SyncTail& t = b == 0 ? t0 : *t1;
    const float* tot = b == 0 ? scratch : scratch + n0;
    if (t.mode == 1) {
      const float n = t.count_per_rank * t.sym.world;
      if (threadIdx.x == 0 && t.nbt != nullptr) *t.nbt += 1;
      for (int c = threadIdx.x; c < t.C; c += blockDim.x)
        bn_finalize_channel(c, tot[c], tot[t.C + c], n, t.gamma, t.beta, t.running_mean, t.running_var, t.momentum, t.eps,
                            t.scale, t.shift, t.mean, t.invstd);
    } else {
      for (int i = threadIdx.x; i < 2 * t.C; i += blockDim.x) t.totals[i] = tot[i];
    }
  }
}

}  // namespace
}  // namespace pmd
#endif
