// Collectives as kernels over peer-mapped memory (NVLink 5 / NVSwitch): cross-rank barrier, fused
// gradient all-reduce + SGD-Nesterov (one-shot, two-shot owner-computes, NVLS multimem), plain
// all-reduce, broadcast, and the tiny SyncBatchNorm / metric exchanges.
//
// Replaces NCCL under DDP's Reducer and SyncBatchNorm (ref:main.py:43-44, 109-110, 173-177;
// SURVEY.md C2-C6, N3-N7, W5-W8, W11-W12, W18).  Protocol: flags[slot][src] in every rank's signal
// pad; a rank announces epoch e by a release-store into each peer's pad and waits until its own pad
// shows >= e from every peer (acquire loads).  Epoch counters live on the device and only ever
// increase, so nothing is reset and a captured CUDA graph can replay the kernels.  Summation is
// always in rank order 0..W-1, so all ranks compute bitwise-identical results.
#include "common.cuh"
#include "sgd_math.cuh"
#include "bn_math.cuh"
#include "sym.cuh"

namespace pmd {
namespace {

// Block-wide cross-rank barrier; every thread of the block calls it.  `publish` = this block has
// written peer memory that the other side will read after the barrier.
__device__ void rank_barrier(const PmdSym& s, int slot, bool publish) {
  if (s.world == 1) return;
  __shared__ uint32_t epoch_s;
  if (publish) __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* ep = epoch_ptr(s, slot);
    const uint32_t e = *ep + 1;
    *ep = e;
    epoch_s = e;
  }
  __syncthreads();
  const uint32_t e = epoch_s;
  if (threadIdx.x < s.world) {
    st_release_sys(flag_ptr(s, threadIdx.x, slot, s.rank), e);
    const uint32_t* mine = flag_ptr(s, s.rank, slot, threadIdx.x);
    const unsigned long long t0 = globaltimer_ns();
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - e) < 0) {
      if (((++spins) & 0xfff) == 0 && globaltimer_ns() - t0 > 30000000000ULL) {   // 30 s: a peer is gone
        printf("pmdist: rank %d timed out waiting for rank %d (slot %d, epoch %u)\n", s.rank, threadIdx.x, slot, e);
        __trap();
      }
    }
  }
  __syncthreads();
}

__global__ void barrier_kernel(PmdSym s, int slot) { rank_barrier(s, slot, false); }

// ------------------------------------------------------------------ fused all-reduce + SGD
template <int ALGO>
__global__ void __launch_bounds__(512) allreduce_sgd_kernel(PmdSym s, PmdAllreduceSgd a, int slot0) {
  const int slot = slot0 + blockIdx.x;
  const SgdHyper h{a.d_lr != nullptr ? *a.d_lr : a.lr, a.momentum, a.weight_decay, 1.0f / s.world, a.nesterov};
  rank_barrier(s, slot, false);   // every rank's backward has produced this range
  const long long n4 = a.n / 4, first4 = a.first / 4;
  long long lo = 0, hi = n4;
  if (ALGO != PMD_ALGO_ONESHOT) {   // owner-computes: this rank reduces and updates one contiguous slice
    const long long per = (n4 + s.world - 1) / s.world;
    lo = per * s.rank;
    hi = lo + per < n4 ? lo + per : n4;
  }
  float4* p = heap_ptr<float4>(s, s.rank, a.param_off) + first4;
  float4* m = heap_ptr<float4>(s, s.rank, a.mom_off) + first4;
  for (long long i = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 g;
    if (ALGO == PMD_ALGO_NVLS) {
      g = mc_ld_reduce_f4(reinterpret_cast<const float4*>(s.mc_base + a.grad_off) + first4 + i);
    } else {
      g = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int w = 0; w < s.world; ++w) {   // fixed order -> bitwise identical on every rank
        const float4 v = ld_peer_f4(heap_ptr<const float4>(s, w, a.grad_off) + first4 + i);
        g.x += v.x;
        g.y += v.y;
        g.z += v.z;
        g.w += v.w;
      }
    }
    float4 pv = p[i], mv = m[i];
    sgd_update(pv.x, g.x, mv.x, h);
    sgd_update(pv.y, g.y, mv.y, h);
    sgd_update(pv.z, g.z, mv.z, h);
    sgd_update(pv.w, g.w, mv.w, h);
    m[i] = mv;
    const float2 pb = make_float2(__uint_as_float(pack_bf16x2(pv.x, pv.y)), __uint_as_float(pack_bf16x2(pv.z, pv.w)));
    if (ALGO == PMD_ALGO_ONESHOT) {
      p[i] = pv;
      (heap_ptr<float2>(s, s.rank, a.bf16_off) + first4)[i] = pb;
    } else if (ALGO == PMD_ALGO_NVLS) {
      mc_st_f4(reinterpret_cast<float4*>(s.mc_base + a.param_off) + first4 + i, pv);
      mc_st_f2(reinterpret_cast<float2*>(s.mc_base + a.bf16_off) + first4 + i, pb);
    } else {
      for (int w = 0; w < s.world; ++w) {
        (heap_ptr<float4>(s, w, a.param_off) + first4)[i] = pv;
        (heap_ptr<float2>(s, w, a.bf16_off) + first4)[i] = pb;
      }
    }
  }
  // one-shot: peers must be done READING my gradients before the next step overwrites them;
  // two-shot / NVLS: my stores into the peers' weight slabs must have landed before they continue.
  rank_barrier(s, slot, ALGO != PMD_ALGO_ONESHOT);
}

// Small ranges (the step's last bucket: nothing is left to overlap it, so its LATENCY is exposed): push protocol instead
// of barrier + peer loads + barrier.  Single block; every rank ends up with identical sums (rank order) and applies the
// identical update to the whole range, so weights stay bitwise equal and momentum is replicated.
__global__ void __launch_bounds__(512) allreduce_sgd_ll_kernel(PmdSym s, PmdAllreduceSgd a, int slot) {
  __shared__ uint32_t epoch_s;
  const SgdHyper h{a.d_lr != nullptr ? *a.d_lr : a.lr, a.momentum, a.weight_decay, 1.0f / s.world, a.nesterov};
  const uint32_t epoch = ll_next_epoch(s, slot, &epoch_s);
  const int n = static_cast<int>(a.n);
  const float* g = heap_ptr<const float>(s, s.rank, a.grad_off) + a.first;
  float* p = heap_ptr<float>(s, s.rank, a.param_off) + a.first;
  float* m = heap_ptr<float>(s, s.rank, a.mom_off) + a.first;
  __nv_bfloat16* pb = heap_ptr<__nv_bfloat16>(s, s.rank, a.bf16_off) + a.first;
  ll_push(s, g, n, a.recv_off, epoch);
  const uint2* mine = heap_ptr<const uint2>(s, s.rank, a.recv_off) + static_cast<long long>(epoch & 1) * s.world * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float sum = ll_poll_sum_one(s, mine, n, i, __ldcg(g + i), epoch, slot);
    float pv = p[i], mv = m[i];
    sgd_update(pv, sum, mv, h);
    p[i] = pv;
    m[i] = mv;
    pb[i] = __float2bfloat16_rn(pv);
  }
}

// ------------------------------------------------------------------ plain all-reduce (sweep primitive)
template <int ALGO>
__global__ void __launch_bounds__(512) allreduce_f32_kernel(PmdSym s, unsigned long long in_off,
                                                            unsigned long long out_off, long long n4, int slot0) {
  const int slot = slot0 + blockIdx.x;
  rank_barrier(s, slot, false);
  long long lo = 0, hi = n4;
  if (ALGO != PMD_ALGO_ONESHOT) {
    const long long per = (n4 + s.world - 1) / s.world;
    lo = per * s.rank;
    hi = lo + per < n4 ? lo + per : n4;
  }
  for (long long i = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 g;
    if (ALGO == PMD_ALGO_NVLS) {
      g = mc_ld_reduce_f4(reinterpret_cast<const float4*>(s.mc_base + in_off) + i);
      mc_st_f4(reinterpret_cast<float4*>(s.mc_base + out_off) + i, g);
    } else {
      g = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int w = 0; w < s.world; ++w) {
        const float4 v = ld_peer_f4(heap_ptr<const float4>(s, w, in_off) + i);
        g.x += v.x;
        g.y += v.y;
        g.z += v.z;
        g.w += v.w;
      }
      if (ALGO == PMD_ALGO_ONESHOT) {
        heap_ptr<float4>(s, s.rank, out_off)[i] = g;
      } else {
        for (int w = 0; w < s.world; ++w) heap_ptr<float4>(s, w, out_off)[i] = g;
      }
    }
  }
  rank_barrier(s, slot, ALGO != PMD_ALGO_ONESHOT);
}

__global__ void __launch_bounds__(512) broadcast_kernel(PmdSym s, unsigned long long off, long long n16, int src,
                                                        int slot0) {
  const int slot = slot0 + blockIdx.x;
  rank_barrier(s, slot, false);
  if (s.rank != src) {
    const float4* from = heap_ptr<const float4>(s, src, off);
    float4* to = heap_ptr<float4>(s, s.rank, off);
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n16;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
      to[i] = ld_peer_f4(from + i);
  }
  rank_barrier(s, slot, false);   // the source may not touch the range until everyone has pulled it
}

// ------------------------------------------------------------------ tiny exchanges
__global__ void bn_finalize_push_kernel(PmdSym s, int slot, const float* __restrict__ stats, unsigned long long recv_off,
                                        float count_per_rank, int C, const float* __restrict__ gamma,
                                        const float* __restrict__ beta, float* running_mean, float* running_var,
                                        long long* num_batches, float momentum, float eps, float* scale, float* shift,
                                        float* save_mean, float* save_invstd, const float* pivot, float* pivot_new) {
  extern __shared__ float tot[];     // [2C]
  __shared__ uint32_t epoch_s;
  // Everything that does not depend on the producer of `stats` happens BEFORE the grid-dependency wait, while that
  // producer is still running: the exchange's epoch counter (private to this slot), and gamma / beta / pivot /
  // running statistics of the first channel this thread finalises (written by earlier steps only).  What remains on
  // the critical path behind the producer is: read sums -> push -> poll -> finalize from registers.
  const uint32_t epoch = ll_next_epoch(s, slot, &epoch_s);
  const int c0 = threadIdx.x;
  float g0 = 0.f, b0 = 0.f, p0 = 0.f, rm0 = 0.f, rv0 = 0.f;
  if (c0 < C) {
    g0 = gamma[c0];
    b0 = beta[c0];
    p0 = pivot != nullptr ? pivot[c0] : 0.f;
    if (running_mean != nullptr) {
      rm0 = running_mean[c0];
      rv0 = running_var[c0];
    }
  }
  pdl_wait();
  ll_push(s, stats, 2 * C, recv_off, epoch);
  ll_poll_sum(s, slot, stats, 2 * C, recv_off, epoch, tot);
  __syncthreads();
  if (threadIdx.x == 0 && num_batches != nullptr) *num_batches += 1;
  const float n = count_per_rank * s.world;
  if (c0 < C) {        // same arithmetic as bn_finalize_channel, operands already in registers
    const float d = tot[c0] / n;
    const float mean = p0 + d;
    const float var = fmaxf(tot[C + c0] / n - d * d, 0.f);
    const float invstd = rsqrtf(var + eps);
    const float a = g0 * invstd;
    if (pivot_new != nullptr) pivot_new[c0] = mean;
    scale[c0] = a;
    shift[c0] = b0 - mean * a;
    save_mean[c0] = mean;
    save_invstd[c0] = invstd;
    if (running_mean != nullptr) {
      const float unbiased = var * (n / fmaxf(n - 1.f, 1.f));
      running_mean[c0] = (1.f - momentum) * rm0 + momentum * mean;
      running_var[c0] = (1.f - momentum) * rv0 + momentum * unbiased;
    }
  }
  for (int c = c0 + blockDim.x; c < C; c += blockDim.x)       // C > 512 (Bottleneck stages)
    bn_finalize_channel(c, tot[c], tot[C + c], n, pivot, pivot_new, gamma, beta, running_mean, running_var, momentum, eps,
                        scale, shift, save_mean, save_invstd);
}

// One or two vectors exchanged by ONE block: both are pushed before either is awaited, so a projection block's pair
// of BatchNorm-backward sums costs one NVLink trip.
struct SmallExchangeDev {
  const float* local;
  float* total;
  unsigned long long recv_off;
  int n, slot;
};
__global__ void small_allreduce_push_kernel(PmdSym s, SmallExchangeDev x0, SmallExchangeDev x1, int nvec) {
  extern __shared__ float tot[];     // [x0.n + x1.n]
  __shared__ uint32_t ep_s[2];
  // the epoch counters are private to these slots: advance them while the producer of the vectors is still running
  const uint32_t e0 = ll_next_epoch(s, x0.slot, &ep_s[0]);
  const uint32_t e1 = nvec > 1 ? ll_next_epoch(s, x1.slot, &ep_s[1]) : 0u;
  pdl_wait();
  ll_push(s, x0.local, x0.n, x0.recv_off, e0);
  if (nvec > 1) ll_push(s, x1.local, x1.n, x1.recv_off, e1);
  ll_poll_sum(s, x0.slot, x0.local, x0.n, x0.recv_off, e0, tot);
  if (nvec > 1) ll_poll_sum(s, x1.slot, x1.local, x1.n, x1.recv_off, e1, tot + x0.n);
  __syncthreads();
  for (int i = threadIdx.x; i < x0.n; i += blockDim.x) x0.total[i] = tot[i];
  if (nvec > 1)
    for (int i = threadIdx.x; i < x1.n; i += blockDim.x) x1.total[i] = tot[x0.n + i];
}

__global__ void metric_allreduce_kernel(PmdSym s, int slot, unsigned long long off, double* total) {
  rank_barrier(s, slot, false);
  if (threadIdx.x < 4) {
    double acc = 0.0;
    for (int w = 0; w < s.world; ++w) {
      const volatile double* src = heap_ptr<const double>(s, w, off);
      acc += src[threadIdx.x];
    }
    total[threadIdx.x] = acc;
  }
  rank_barrier(s, slot, false);   // peers are done reading before anyone zeroes its accumulators
}

int clamp_blocks(int blocks, long long work4) {
  if (blocks <= 0) blocks = 32;
  const long long need = ceil_div_ll(work4 > 0 ? work4 : 1, 512);
  if (blocks > need) blocks = static_cast<int>(need);
  if (blocks > 1024) blocks = 1024;
  return blocks < 1 ? 1 : blocks;
}

}  // namespace
}  // namespace pmd

using namespace pmd;

#define SYM_CHECK(sym) PMD_REQUIRE((sym) != nullptr && (sym)->world >= 1 && (sym)->world <= PMD_MAX_RANKS, "bad PmdSym")

extern "C" PMD_API int pmd_dev_barrier(const PmdSym* sym, int slot, void* stream) {
  SYM_CHECK(sym);
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(*sym, slot);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_allreduce_sgd(const PmdSym* sym, const PmdAllreduceSgd* a, int slot, void* stream) {
  SYM_CHECK(sym);
  PMD_REQUIRE(a->n % 4 == 0 && a->first % 4 == 0, "all-reduce ranges must be multiples of 4 elements");
  int algo = a->algo;
  if (sym->world == 1) algo = PMD_ALGO_ONESHOT;
  if (algo == PMD_ALGO_AUTO) algo = PMD_ALGO_TWOSHOT;
  if (algo == PMD_ALGO_NVLS && sym->mc_base == 0) algo = PMD_ALGO_TWOSHOT;
  if (algo == PMD_ALGO_LL) {
    PMD_REQUIRE(a->n <= PMD_LL_MAX_ELEMS && a->recv_off != 0, "the LL all-reduce needs n <= PMD_LL_MAX_ELEMS and a receive area");
    allreduce_sgd_ll_kernel<<<1, 512, 0, static_cast<cudaStream_t>(stream)>>>(*sym, *a, slot);
    count_launch();
    PMD_LAUNCH_CHECK();
    return 0;
  }
  const long long work4 = algo == PMD_ALGO_ONESHOT ? a->n / 4 : ceil_div_ll(a->n / 4, sym->world);
  PmdAllreduceSgd aa = *a;
  aa.blocks = clamp_blocks(a->blocks, work4);
  if (sym->world > 1 && aa.blocks > 128) aa.blocks = 128;   // one flag slot per CTA; the engine reserves 128 per bucket
  PMD_REQUIRE(slot + aa.blocks <= PMD_MAX_SLOTS, "flag slots exhausted");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (algo == PMD_ALGO_ONESHOT)
    allreduce_sgd_kernel<PMD_ALGO_ONESHOT><<<aa.blocks, 512, 0, st>>>(*sym, aa, slot);
  else if (algo == PMD_ALGO_NVLS)
    allreduce_sgd_kernel<PMD_ALGO_NVLS><<<aa.blocks, 512, 0, st>>>(*sym, aa, slot);
  else
    allreduce_sgd_kernel<PMD_ALGO_TWOSHOT><<<aa.blocks, 512, 0, st>>>(*sym, aa, slot);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_allreduce_f32(const PmdSym* sym, unsigned long long in_off, unsigned long long out_off,
                                         long long n, int algo, int blocks, int slot, void* stream) {
  SYM_CHECK(sym);
  PMD_REQUIRE(n % 4 == 0 && in_off % 16 == 0 && out_off % 16 == 0, "all-reduce needs 16-byte aligned, multiple-of-4 ranges");
  if (sym->world == 1 || algo == PMD_ALGO_AUTO) algo = sym->world == 1 ? PMD_ALGO_ONESHOT : PMD_ALGO_TWOSHOT;
  if (algo == PMD_ALGO_NVLS && sym->mc_base == 0) algo = PMD_ALGO_TWOSHOT;
  const long long work4 = algo == PMD_ALGO_ONESHOT ? n / 4 : ceil_div_ll(n / 4, sym->world);
  blocks = clamp_blocks(blocks, work4);
  PMD_REQUIRE(slot + blocks <= PMD_MAX_SLOTS, "flag slots exhausted");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (algo == PMD_ALGO_ONESHOT)
    allreduce_f32_kernel<PMD_ALGO_ONESHOT><<<blocks, 512, 0, st>>>(*sym, in_off, out_off, n / 4, slot);
  else if (algo == PMD_ALGO_NVLS)
    allreduce_f32_kernel<PMD_ALGO_NVLS><<<blocks, 512, 0, st>>>(*sym, in_off, out_off, n / 4, slot);
  else
    allreduce_f32_kernel<PMD_ALGO_TWOSHOT><<<blocks, 512, 0, st>>>(*sym, in_off, out_off, n / 4, slot);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_broadcast(const PmdSym* sym, unsigned long long off, unsigned long long bytes, int src,
                                     int blocks, int slot, void* stream) {
  SYM_CHECK(sym);
  PMD_REQUIRE(bytes % 16 == 0 && off % 16 == 0, "broadcast needs 16-byte aligned ranges");
  blocks = clamp_blocks(blocks, static_cast<long long>(bytes / 16));
  PMD_REQUIRE(slot + blocks <= PMD_MAX_SLOTS, "flag slots exhausted");
  broadcast_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(*sym, off, static_cast<long long>(bytes / 16),
                                                                         src, slot);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_metric_allreduce(const PmdSym* sym, int slot, unsigned long long off, double* d_total,
                                            void* stream) {
  SYM_CHECK(sym);
  metric_allreduce_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(*sym, slot, off, d_total);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_bn_finalize_push(const PmdSym* sym, int slot, const float* d_stats, unsigned long long recv_off,
                                            float count_per_rank, int C, const float* d_gamma, const float* d_beta,
                                            float* d_running_mean, float* d_running_var, long long* d_num_batches,
                                            float momentum, float eps, float* d_scale, float* d_shift, float* d_mean,
                                            float* d_invstd, const float* d_pivot, float* d_pivot_new, void* stream) {
  SYM_CHECK(sym);
  PMD_REQUIRE(2 * C * 4 <= 48 * 1024, "too many channels for the push exchange");
  PMD_CHECK_CUDA(launch_pdl(bn_finalize_push_kernel, dim3(1), dim3(512), 2 * C * 4, static_cast<cudaStream_t>(stream), *sym,
                            slot, d_stats, recv_off, count_per_rank, C, d_gamma, d_beta, d_running_mean, d_running_var,
                            d_num_batches, momentum, eps, d_scale, d_shift, d_mean, d_invstd, d_pivot, d_pivot_new));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_small_allreduce_push(const PmdSym* sym, const PmdSmallExchange* x, int nvec, void* stream) {
  SYM_CHECK(sym);
  PMD_REQUIRE(nvec == 1 || nvec == 2, "one or two vectors per exchange kernel");
  SmallExchangeDev d[2] = {};
  int total = 0;
  for (int i = 0; i < nvec; ++i) {
    d[i].local = x[i].local;
    d[i].total = x[i].total;
    d[i].recv_off = x[i].recv_off;
    d[i].n = x[i].n;
    d[i].slot = x[i].slot;
    total += x[i].n;
  }
  PMD_REQUIRE(total * 4 <= 48 * 1024, "vectors too long for the push exchange");
  PMD_CHECK_CUDA(launch_pdl(small_allreduce_push_kernel, dim3(1), dim3(512), total * 4, static_cast<cudaStream_t>(stream), *sym,
                            d[0], d[1], nvec));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}
