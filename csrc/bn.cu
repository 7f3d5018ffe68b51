// BatchNorm for NHWC bf16 activations: batch statistics, finalize (+ running stats), fused
// apply (+ residual + ReLU), and the two backward passes.
//
// Replaces torch's SyncBatchNorm autograd function and its ATen kernels (ref:main.py:43;
// torch/nn/modules/_functions.py:39-179).  Cross-GPU exchange of the
// per-channel partials is done by comm.cu (peer loads over NVLink); these kernels take pointer
// tables so that one code path serves 1..8 ranks and rank-order-fixed summation keeps every rank
// bitwise identical.
#include "common.cuh"
#include "bn_math.cuh"
#include "../include/pmdist.h"

namespace pmd {
namespace {

constexpr int kBlock = 256;
// The streaming kernels run with a block size that is a multiple of the number of 8-channel groups (C / 8), and
// their grid stride is a multiple of the block size, so a thread meets the SAME channel group in every iteration:
// its per-channel coefficients are loaded into registers once, and the loop carries no index arithmetic (the
// 64-bit `i % groups` and the per-iteration shared-memory loads used to cost more issue slots than the math).
inline int fixed_block(int C) {
  const int groups = C / 8;
  return groups >= kBlock ? groups : (kBlock / groups) * groups;
}

// Generic per-channel two-value reduction over rows of a [rows, C] bf16 tensor.  A thread owns a
// group of 8 consecutive channels (one 16-byte load) and walks rows with stride (threads / groups).
template <typename F>
__device__ __forceinline__ void column_reduce2(F f, long long rows, int C, float* out0, float* out1) {
  extern __shared__ float red[];          // [2][rows_per_iter][C] would be large; reduce in two steps
  const int groups = C / 8;
  const int g = threadIdx.x % groups;
  const int rl = threadIdx.x / groups;
  const int rows_per_iter = blockDim.x / groups;
  float a0[8], a1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a0[i] = a1[i] = 0.f;
  if (rl < rows_per_iter) {
    for (long long r = static_cast<long long>(blockIdx.x) * rows_per_iter + rl; r < rows;
         r += static_cast<long long>(gridDim.x) * rows_per_iter)
      f(r, g * 8, a0, a1);
  }
  // cross-thread reduction: red[rl][c] for each of the two values
  float* r0 = red;
  float* r1 = red + rows_per_iter * C;
  if (rl < rows_per_iter) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      r0[rl * C + g * 8 + i] = a0[i];
      r1[rl * C + g * 8 + i] = a1[i];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s0 = 0.f, s1 = 0.f;
    for (int j = 0; j < rows_per_iter; ++j) {
      s0 += r0[j * C + c];
      s1 += r1[j * C + c];
    }
    red_add_f32(out0 + c, s0);
    red_add_f32(out1 + c, s1);
  }
}

// Streaming read: every BatchNorm pass touches its inputs once, so they are loaded with an L2 evict-first policy -- the
// tensor the pass WRITES (read next by a convolution) then survives in the 126 MB L2 instead of being pushed out by them.
// Same box, same run: 1.4437 -> 1.4155 ms/step together with the convolutions' operand hints (conv_l2_hint_mask).
#ifndef PMD_BN_STREAM      // -DPMD_BN_STREAM=0 builds the plain-load variant (A/B: PMDIST_LIB=<that library>)
#define PMD_BN_STREAM 1
#endif
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 u;
  if (!PMD_BN_STREAM) return *reinterpret_cast<const uint4*>(p);
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
               : "l"(p), "l"(pol)
               : "memory");
  return u;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = ld_stream16(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = unpack_bf16x2(w[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                            pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}

__global__ void bn_stats_kernel(const __nv_bfloat16* __restrict__ y, long long rows, int C, float* stats,
                                const float* __restrict__ pivot) {
  column_reduce2(
      [&](long long r, int c0, float(&a0)[8], float(&a1)[8]) {
        float v[8];
        load8(y + r * C + c0, v);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float d = v[i] - (pivot != nullptr ? __ldg(pivot + c0 + i) : 0.f);     // sums around the pivot (bn_math.cuh)
          a0[i] += d;
          a1[i] += d * d;
        }
      },
      rows, C, stats, stats + C);
}

__global__ void bn_finalize_kernel(const float* const* __restrict__ stats_ranks, int world, float count_per_rank,
                                   int C, const float* __restrict__ gamma, const float* __restrict__ beta,
                                   float* running_mean, float* running_var, long long* num_batches, float momentum,
                                   float eps, float* scale, float* shift, float* save_mean, float* save_invstd,
                                   const float* pivot, float* pivot_new) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && num_batches != nullptr) *num_batches += 1;
  if (c >= C) return;
  float s1 = 0.f, s2 = 0.f;
  for (int w = 0; w < world; ++w) {            // fixed rank order -> identical result on every rank
    const volatile float* st = stats_ranks[w];
    s1 += st[c];
    s2 += st[C + c];
  }
  bn_finalize_channel(c, s1, s2, count_per_rank * world, pivot, pivot_new, gamma, beta, running_mean, running_var, momentum,
                      eps, scale, shift, save_mean, save_invstd);
}

__global__ void bn_eval_affine_kernel(int C, const float* gamma, const float* beta, const float* rm, const float* rv,
                                      float eps, float* scale, float* shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float a = gamma[c] * rsqrtf(rv[c] + eps);
  scale[c] = a;
  shift[c] = beta[c] - rm[c] * a;
}

__device__ __forceinline__ void lds8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

template <int RES>   // 0: none, 1: plain residual, 2: residual with its own affine
__global__ void bn_apply_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ scale,
                                const float* __restrict__ shift, const __nv_bfloat16* __restrict__ res,
                                const float* __restrict__ rscale, const float* __restrict__ rshift, int relu,
                                __nv_bfloat16* __restrict__ z, unsigned char* __restrict__ relu_bits, long long total8,
                                int C) {
  pdl_wait();
  extern __shared__ float aff[];     // [scale | shift | rscale | rshift], C floats each: 2-4 wide shared loads per
  for (int c = threadIdx.x; c < C; c += blockDim.x) {   // 8 elements instead of 16-32 scalar global loads
    aff[c] = scale[c];
    aff[C + c] = shift[c];
    if (RES == 2) {
      aff[2 * C + c] = rscale[c];
      aff[3 * C + c] = rshift[c] ;
    }
  }
  __syncthreads();
  const int groups = C / 8;
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % groups) * 8;      // blockDim.x % groups == 0 (fixed_block)
  float sc[8], sh[8], rsc[8], rsh[8];
  lds8(aff + c0, sc);
  lds8(aff + C + c0, sh);
  if (RES == 2) {
    lds8(aff + 2 * C + c0, rsc);
    lds8(aff + 3 * C + c0, rsh);
  }
  auto one = [&](long long i, const float (&v)[8], const float (&rv)[8]) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(v[j], sc[j], sh[j]);
    if (RES == 2) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += fmaf(rv[j], rsc[j], rsh[j]);
    } else if (RES == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += rv[j];
    }
    if (relu) {
      if (relu_bits != nullptr) {        // one byte per 8-channel group: which outputs passed the ReLU (read by backward
        unsigned bits = 0;               // instead of the 16x larger z tensor)
#pragma unroll
        for (int j = 0; j < 8; ++j) bits |= (o[j] > 0.f ? 1u : 0u) << j;
        relu_bits[i] = static_cast<unsigned char>(bits);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaxf(o[j], 0.f);
    }
    store8(z + i * 8, o);
  };
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  for (; i + T < total8; i += 2 * T) {
    float va[8], vb[8], ra[8], rb[8];
    load8(y + i * 8, va);
    load8(y + (i + T) * 8, vb);
    if (RES != 0) {
      load8(res + i * 8, ra);
      load8(res + (i + T) * 8, rb);
    }
    one(i, va, ra);
    one(i + T, vb, rb);
  }
  for (; i < total8; i += T) {
    float va[8], ra[8];
    load8(y + i * 8, va);
    if (RES != 0) load8(res + i * 8, ra);
    one(i, va, ra);
  }
}

// ---------------------------------------------------------------- finalize + apply in one kernel
// The statistics of a BatchNorm are complete when its convolution has finished, so the separate finalize launch
// (a single tiny block between two large kernels, i.e. pure dependency latency) can go: every block of the
// apply kernel derives scale/shift from the [sum, sumsq] vector itself (2C loads, C rsqrt -- noise next to the
// activation traffic), and block 0 also stores what backward and the checkpoint need (mean, invstd, scale, shift,
// running statistics, num_batches_tracked).  With RES == 2 the projection shortcut's BatchNorm is finalised here
// too.  Single-rank statistics only: under multi-rank SyncBatchNorm the push-exchange kernel (comm.cu) finalises them
// and the plain bn_apply_kernel runs instead.
struct BnFinalizeDev {
  const float* stats;
  const float* gamma;
  const float* beta;
  float* running_mean;
  float* running_var;
  long long* nbt;
  float *scale, *shift, *mean, *invstd;
  const float* pivot;      // statistics were taken around this (null = 0)
  float* pivot_new;        // receives the batch mean (null: not kept)
};

template <int RES>
__global__ void bn_finalize_apply_kernel(BnFinalizeDev f0, BnFinalizeDev f1, float count_per_rank,
                                         float momentum, float eps, const __nv_bfloat16* __restrict__ y,
                                         const __nv_bfloat16* __restrict__ res, int relu, __nv_bfloat16* __restrict__ z,
                                         unsigned char* __restrict__ relu_bits, long long total8, int C) {
  pdl_wait();
  extern __shared__ float aff[];     // [scale | shift | rscale | rshift] (C each) + totals scratch [2C]
  float* tot = aff + 4 * C;
  const float n = count_per_rank;
#pragma unroll
  for (int b = 0; b < (RES == 2 ? 2 : 1); ++b) {
    const BnFinalizeDev& f = b == 0 ? f0 : f1;
    for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) tot[c] = f.stats[c];
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float d = tot[c] / n;
      const float mean = (f.pivot != nullptr ? f.pivot[c] : 0.f) + d;        // sums around the pivot (bn_math.cuh)
      const float var = fmaxf(tot[C + c] / n - d * d, 0.f);
      const float invstd = rsqrtf(var + eps);
      const float a = f.gamma[c] * invstd;
      const float sh = f.beta[c] - mean * a;
      aff[b * 2 * C + c] = a;
      aff[b * 2 * C + C + c] = sh;
      if (blockIdx.x == 0) {
        f.scale[c] = a;
        f.shift[c] = sh;
        f.mean[c] = mean;
        f.invstd[c] = invstd;
        if (f.pivot_new != nullptr) f.pivot_new[c] = mean;   // a different array: the other blocks still read f.pivot
        if (f.running_mean != nullptr) {
          const float unbiased = var * (n / fmaxf(n - 1.f, 1.f));
          f.running_mean[c] = (1.f - momentum) * f.running_mean[c] + momentum * mean;
          f.running_var[c] = (1.f - momentum) * f.running_var[c] + momentum * unbiased;
        }
        if (c == 0 && f.nbt != nullptr) *f.nbt += 1;
      }
    }
    __syncthreads();
  }
  const int groups = C / 8;
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % groups) * 8;      // blockDim.x % groups == 0 (fixed_block)
  float sc[8], sh[8], rsc[8], rsh[8];
  lds8(aff + c0, sc);
  lds8(aff + C + c0, sh);
  if (RES == 2) {
    lds8(aff + 2 * C + c0, rsc);
    lds8(aff + 3 * C + c0, rsh);
  }
  auto one = [&](long long i, const float (&v)[8], const float (&rv)[8]) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(v[j], sc[j], sh[j]);
    if (RES == 2) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += fmaf(rv[j], rsc[j], rsh[j]);
    } else if (RES == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += rv[j];
    }
    if (relu) {
      if (relu_bits != nullptr) {        // one byte per 8-channel group: which outputs passed the ReLU (read by backward
        unsigned bits = 0;               // instead of the 16x larger z tensor)
#pragma unroll
        for (int j = 0; j < 8; ++j) bits |= (o[j] > 0.f ? 1u : 0u) << j;
        relu_bits[i] = static_cast<unsigned char>(bits);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaxf(o[j], 0.f);
    }
    store8(z + i * 8, o);
  };
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  for (; i + T < total8; i += 2 * T) {
    float va[8], vb[8], ra[8], rb[8];
    load8(y + i * 8, va);
    load8(y + (i + T) * 8, vb);
    if (RES != 0) {
      load8(res + i * 8, ra);
      load8(res + (i + T) * 8, rb);
    }
    one(i, va, ra);
    one(i + T, vb, rb);
  }
  for (; i < total8; i += T) {
    float va[8], ra[8];
    load8(y + i * 8, va);
    if (RES != 0) load8(res + i * 8, ra);
    one(i, va, ra);
  }
}

// ---------------------------------------------------------------- backward, 1 or 2 branches
// A residual block's output gradient dz feeds the last main-path BatchNorm AND (projection blocks) the
// shortcut's BatchNorm through the same ReLU mask; handling both in one pass reads dz and the mask
// source once instead of twice.  MASK: 0 none, 1 z > 0, 2 recomputed from branch 0's own forward
// affine (scale*y + shift > 0; valid when that BatchNorm had no residual), which skips reading z at all.
struct BnBranchDev {
  const __nv_bfloat16* y;
  const float* mean;
  const float* invstd;
  const float* gamma;
  float* sums;          // [2C] local sums: accumulated by reduce, read by apply
  const float* totals;  // [2C] sums over all ranks (== sums on one rank)
  __nv_bfloat16* dy;
  float* dgamma;
  float* dbeta;
  float* pivot;             // next step's statistics pivot := pivot_new (this step's batch mean); both may be null
  const float* pivot_new;
};

template <int NB, int MASK>
__global__ void __launch_bounds__(kBlock, 4)
bn_bwd_reduce_multi_kernel(const __nv_bfloat16* __restrict__ dz, const __nv_bfloat16* __restrict__ z,
                           const float* __restrict__ mscale, const float* __restrict__ mshift, BnBranchDev b0,
                           BnBranchDev b1, long long rows, int C) {
  // Accumulates S_g = sum(g) and S_gy = sum(g*y) per channel; sum(g*xhat) = (S_gy - mean*S_g)*invstd is formed
  // once per block at the end, so the streaming loop carries no per-channel statistics in registers (the
  // register budget decides how many of these DRAM-bound blocks fit on an SM).
  pdl_wait();
  extern __shared__ float red[];      // [mask consts 2C] + reduction scratch [(1+NB) * rpi * C]
  float* msm = red;
  float* scratch = red + 2 * C;
  const int groups = C / 8;
  const int g = threadIdx.x % groups;
  const int rl = threadIdx.x / groups;
  const int rpi = blockDim.x / groups;
  const int c0 = g * 8;
  if (MASK == 2) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      msm[c] = mscale[c];
      msm[C + c] = mshift[c];
    }
    __syncthreads();
  }
  float sg[8], sgy0[8], sgy1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) sg[i] = sgy0[i] = sgy1[i] = 0.f;
  if (rl < rpi) {
    const long long stride = static_cast<long long>(gridDim.x) * rpi;
    float ms[8], mh[8];
    if (MASK == 2) {
      lds8(msm + c0, ms);
      lds8(msm + C + c0, mh);
    }
    const unsigned char* zbits = reinterpret_cast<const unsigned char*>(z);     // MASK == 3: one byte per 8-channel group
    auto body = [&](const float (&gin)[8], const float (&y0)[8], const float (&zin)[8], const float (&y1)[8],
                    unsigned bits) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float gv = gin[i];
        if (MASK == 1) gv = zin[i] > 0.f ? gv : 0.f;
        if (MASK == 3) gv = ((bits >> i) & 1u) ? gv : 0.f;
        if (MASK == 2) gv = fmaf(y0[i], ms[i], mh[i]) > 0.f ? gv : 0.f;
        sg[i] += gv;
        sgy0[i] = fmaf(gv, y0[i], sgy0[i]);
        if (NB == 2) sgy1[i] = fmaf(gv, y1[i], sgy1[i]);
      }
    };
    long long r = static_cast<long long>(blockIdx.x) * rpi + rl;
    for (; r + stride < rows; r += 2 * stride) {   // two rows in flight per thread
      float ga[8], gb[8], ya[8], yb[8], za[8], zb[8], y1a[8], y1b[8];
      load8(dz + r * C + c0, ga);
      load8(dz + (r + stride) * C + c0, gb);
      load8(b0.y + r * C + c0, ya);
      load8(b0.y + (r + stride) * C + c0, yb);
      if (MASK == 1) {
        load8(z + r * C + c0, za);
        load8(z + (r + stride) * C + c0, zb);
      }
      if (NB == 2) {
        load8(b1.y + r * C + c0, y1a);
        load8(b1.y + (r + stride) * C + c0, y1b);
      }
      unsigned ba = 0, bb = 0;
      if (MASK == 3) {
        ba = zbits[r * groups + g];
        bb = zbits[(r + stride) * groups + g];
      }
      body(ga, ya, za, y1a, ba);
      body(gb, yb, zb, y1b, bb);
    }
    for (; r < rows; r += stride) {
      float ga[8], ya[8], za[8], y1a[8];
      load8(dz + r * C + c0, ga);
      load8(b0.y + r * C + c0, ya);
      if (MASK == 1) load8(z + r * C + c0, za);
      if (NB == 2) load8(b1.y + r * C + c0, y1a);
      body(ga, ya, za, y1a, MASK == 3 ? zbits[r * groups + g] : 0u);
    }
    // block reduction scratch: [v][rl][c], v = 0: S_g, 1: S_gy0, 2: S_gy1
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      scratch[(0 * rpi + rl) * C + c0 + i] = sg[i];
      scratch[(1 * rpi + rl) * C + c0 + i] = sgy0[i];
      if (NB == 2) scratch[(2 * rpi + rl) * C + c0 + i] = sgy1[i];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float tg = 0.f, t0 = 0.f, t1 = 0.f;
    for (int j = 0; j < rpi; ++j) {
      tg += scratch[(0 * rpi + j) * C + c];
      t0 += scratch[(1 * rpi + j) * C + c];
      if (NB == 2) t1 += scratch[(2 * rpi + j) * C + c];
    }
    red_add_f32(b0.sums + c, tg);
    red_add_f32(b0.sums + C + c, (t0 - b0.mean[c] * tg) * b0.invstd[c]);
    if (NB == 2) {
      red_add_f32(b1.sums + c, tg);
      red_add_f32(b1.sums + C + c, (t1 - b1.mean[c] * tg) * b1.invstd[c]);
    }
  }
}

template <int NB, int MASK>
__global__ void __launch_bounds__(kBlock, NB == 1 ? 3 : 2) bn_bwd_apply_multi_kernel(const __nv_bfloat16* __restrict__ dz,
                                          const __nv_bfloat16* __restrict__ z, const float* __restrict__ mscale,
                                          const float* __restrict__ mshift, BnBranchDev b0, BnBranchDev b1,
                                          float count_total, __nv_bfloat16* __restrict__ g_out, long long total8,
                                          int C) {
  // dy = gamma*invstd * (g - mean_g - xhat*mean_gx), xhat = (y - mean)*invstd, folded per channel into
  // dy = a*g + k1*y + k0.  Shared layout: per branch [a | k1 | k0] (C floats each), then [mscale | mshift].
  // br.totals = [sum_g, sum_gx] over ALL ranks (left there by the producer's exchange under SyncBatchNorm).
  pdl_wait();
  extern __shared__ float coef[];
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const BnBranchDev& br = b == 0 ? b0 : b1;
      float* cf = coef + b * 3 * C;
      const float is = br.invstd[c], mu = br.mean[c];
      const float a = br.gamma[c] * is;
      const float tg = br.totals[c];
      const float tgx = br.totals[C + c];
      const float mg = tg / count_total, mgx = tgx / count_total;
      cf[c] = a;
      cf[C + c] = -a * mgx * is;
      cf[2 * C + c] = a * (mgx * is * mu - mg);
      if (blockIdx.x == 0 && br.dgamma != nullptr) {
        br.dbeta[c] += br.sums[c];
        br.dgamma[c] += br.sums[C + c];
      }
      // forward of this step is over and the next one has not begun: the only moment nobody reads the pivot
      if (blockIdx.x == 0 && br.pivot != nullptr) br.pivot[c] = br.pivot_new[c];
    }
    if (MASK == 2) {
      coef[NB * 3 * C + c] = mscale[c];
      coef[NB * 3 * C + C + c] = mshift[c];
    }
  }
  __syncthreads();
  const int groups = C / 8;
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int c0 = static_cast<int>(threadIdx.x % groups) * 8;      // blockDim.x % groups == 0 (fixed_block)
  float ka[8], k1[8], k0[8], qa[8], q1[8], q0[8], ms[8], mh[8];   // this thread's channel group, for the whole loop
  lds8(coef + c0, ka);
  lds8(coef + C + c0, k1);
  lds8(coef + 2 * C + c0, k0);
  if (NB == 2) {
    lds8(coef + 3 * C + c0, qa);
    lds8(coef + 4 * C + c0, q1);
    lds8(coef + 5 * C + c0, q0);
  }
  if (MASK == 2) {
    lds8(coef + NB * 3 * C + c0, ms);
    lds8(coef + NB * 3 * C + C + c0, mh);
  }
  const unsigned char* zbits = reinterpret_cast<const unsigned char*>(z);     // MASK == 3: one byte per 8-channel group
  auto one = [&](long long i, float (&gv)[8], const float (&y0)[8], const float (&zv)[8], const float (&y1)[8]) {
    float o[8];
    if (MASK == 3) {
      const unsigned bits = zbits[i];
#pragma unroll
      for (int j = 0; j < 8; ++j) gv[j] = ((bits >> j) & 1u) ? gv[j] : 0.f;
    } else if (MASK == 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) gv[j] = zv[j] > 0.f ? gv[j] : 0.f;
    } else if (MASK == 2) {
#pragma unroll
      for (int j = 0; j < 8; ++j) gv[j] = fmaf(y0[j], ms[j], mh[j]) > 0.f ? gv[j] : 0.f;
    }
    if (g_out != nullptr) store8(g_out + i * 8, gv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(ka[j], gv[j], fmaf(k1[j], y0[j], k0[j]));
    store8(b0.dy + i * 8, o);
    if (NB == 2) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaf(qa[j], gv[j], fmaf(q1[j], y1[j], q0[j]));
      store8(b1.dy + i * 8, o);
    }
  };
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  for (; i + T < total8; i += 2 * T) {     // two 16-byte groups per thread in flight
    float ga[8], gb[8], ya[8], yb[8], za[8], zb[8], y1a[8], y1b[8];
    load8(dz + i * 8, ga);
    load8(dz + (i + T) * 8, gb);
    load8(b0.y + i * 8, ya);
    load8(b0.y + (i + T) * 8, yb);
    if (MASK == 1) {
      load8(z + i * 8, za);
      load8(z + (i + T) * 8, zb);
    }
    if (NB == 2) {
      load8(b1.y + i * 8, y1a);
      load8(b1.y + (i + T) * 8, y1b);
    }
    one(i, ga, ya, za, y1a);
    one(i + T, gb, yb, zb, y1b);
  }
  for (; i < total8; i += T) {
    float ga[8], ya[8], za[8], y1a[8];
    load8(dz + i * 8, ga);
    load8(b0.y + i * 8, ya);
    if (MASK == 1) load8(z + i * 8, za);
    if (NB == 2) load8(b1.y + i * 8, y1a);
    one(i, ga, ya, za, y1a);
  }
}

int grid_for(long long work_items, int per_block) {
  long long b = ceil_div_ll(work_items, per_block);
  const long long cap = 148LL * 8;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

// Grid-stride kernels: at most ONE resident wave (148 SMs x the blocks of this kernel an SM holds), so there is no
// partial last wave and every block streams for the whole kernel.
template <typename Kernel>
int wave_grid(Kernel kernel, int block, int smem, long long work_items, int per_block) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, smem) != cudaSuccess || per_sm < 1) {
    cudaGetLastError();
    per_sm = 1;
  }
  const long long b = ceil_div_ll(work_items, per_block), cap = 148LL * per_sm;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace
}  // namespace pmd

using namespace pmd;

#define BN_CHECK_C(C) PMD_REQUIRE((C) % 8 == 0 && (C) >= 8 && (C) <= 2048, "BatchNorm channel count must be a multiple of 8 in [8, 2048]")

static int bn_stats_impl(const void* d_y, long long rows, int C, float* d_stats, const float* d_pivot, void* stream);
extern "C" PMD_API int pmd_bn_stats(const void* d_y, long long rows, int C, float* d_stats, void* stream) {
  return bn_stats_impl(d_y, rows, C, d_stats, nullptr, stream);
}
extern "C" PMD_API int pmd_bn_stats_pivot(const void* d_y, long long rows, int C, float* d_stats, const float* d_pivot,
                                          void* stream) {
  return bn_stats_impl(d_y, rows, C, d_stats, d_pivot, stream);
}
static int bn_stats_impl(const void* d_y, long long rows, int C, float* d_stats, const float* d_pivot, void* stream) {
  BN_CHECK_C(C);
  const int groups = C / 8, rpi = kBlock / groups > 0 ? kBlock / groups : 1;
  PMD_REQUIRE(groups <= kBlock, "C too large for bn_stats");
  const int smem = 2 * rpi * C * 4;
  bn_stats_kernel<<<grid_for(rows, rpi * 4), kBlock, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(d_y), rows, C, d_stats, d_pivot);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_bn_finalize_pivot(const float* const* d_stats_ranks, int world, float count_per_rank, int C,
                                             const float* d_gamma, const float* d_beta, float* d_running_mean,
                                             float* d_running_var, long long* d_num_batches, float momentum, float eps,
                                             float* d_scale, float* d_shift, float* d_mean, float* d_invstd,
                                             const float* d_pivot, float* d_pivot_new, void* stream) {
  bn_finalize_kernel<<<ceil_div(C, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      d_stats_ranks, world, count_per_rank, C, d_gamma, d_beta, d_running_mean, d_running_var, d_num_batches, momentum,
      eps, d_scale, d_shift, d_mean, d_invstd, d_pivot, d_pivot_new);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}
extern "C" PMD_API int pmd_bn_finalize(const float* const* d_stats_ranks, int world, float count_per_rank, int C,
                               const float* d_gamma, const float* d_beta, float* d_running_mean, float* d_running_var,
                               long long* d_num_batches, float momentum, float eps, float* d_scale, float* d_shift,
                               float* d_mean, float* d_invstd, void* stream) {
  return pmd_bn_finalize_pivot(d_stats_ranks, world, count_per_rank, C, d_gamma, d_beta, d_running_mean, d_running_var,
                               d_num_batches, momentum, eps, d_scale, d_shift, d_mean, d_invstd, nullptr, nullptr, stream);
}

extern "C" PMD_API int pmd_bn_eval_affine(int C, const float* d_gamma, const float* d_beta, const float* d_running_mean,
                                  const float* d_running_var, float eps, float* d_scale, float* d_shift, void* stream) {
  bn_eval_affine_kernel<<<ceil_div(C, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      C, d_gamma, d_beta, d_running_mean, d_running_var, eps, d_scale, d_shift);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_bn_apply_bits(const void* d_y, const float* d_scale, const float* d_shift, const void* d_res,
                                         const float* d_res_scale, const float* d_res_shift, int relu, void* d_z,
                                         void* d_relu_bits, long long rows, int C, void* stream) {
  BN_CHECK_C(C);
  const long long total8 = rows * C / 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* y = static_cast<const __nv_bfloat16*>(d_y);
  const __nv_bfloat16* res = static_cast<const __nv_bfloat16*>(d_res);
  __nv_bfloat16* z = static_cast<__nv_bfloat16*>(d_z);
  unsigned char* bits = static_cast<unsigned char*>(d_relu_bits);
  const int blk = fixed_block(C);
  if (d_res == nullptr)
    PMD_CHECK_CUDA(launch_pdl(bn_apply_kernel<0>, dim3(wave_grid(bn_apply_kernel<0>, blk, 2 * C * 4, total8, blk * 2)), dim3(blk),
                              2 * C * 4, st, y, d_scale, d_shift, nullptr, nullptr, nullptr, relu, z, bits, total8, C));
  else if (d_res_scale == nullptr)
    PMD_CHECK_CUDA(launch_pdl(bn_apply_kernel<1>, dim3(wave_grid(bn_apply_kernel<1>, blk, 2 * C * 4, total8, blk * 2)), dim3(blk),
                              2 * C * 4, st, y, d_scale, d_shift, res, nullptr, nullptr, relu, z, bits, total8, C));
  else
    PMD_CHECK_CUDA(launch_pdl(bn_apply_kernel<2>, dim3(wave_grid(bn_apply_kernel<2>, blk, 4 * C * 4, total8, blk * 2)), dim3(blk),
                              4 * C * 4, st, y, d_scale, d_shift, res, d_res_scale, d_res_shift, relu, z, bits, total8, C));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_bn_apply(const void* d_y, const float* d_scale, const float* d_shift, const void* d_res,
                            const float* d_res_scale, const float* d_res_shift, int relu, void* d_z, long long rows,
                            int C, void* stream) {
  return pmd_bn_apply_bits(d_y, d_scale, d_shift, d_res, d_res_scale, d_res_shift, relu, d_z, nullptr, rows, C, stream);
}

static BnBranchDev to_dev(const PmdBnBranch& b) {
  BnBranchDev d;
  d.y = static_cast<const __nv_bfloat16*>(b.y);
  d.mean = b.mean;
  d.invstd = b.invstd;
  d.gamma = b.gamma;
  d.sums = b.sums;
  d.totals = b.totals != nullptr ? b.totals : b.sums;
  d.dy = static_cast<__nv_bfloat16*>(b.dy);
  d.dgamma = b.dgamma;
  d.dbeta = b.dbeta;
  d.pivot = b.pivot_new != nullptr ? b.pivot : nullptr;
  d.pivot_new = b.pivot_new;
  return d;
}

template <int NB>
static cudaError_t launch_reduce(int mask, int smem, cudaStream_t st, const __nv_bfloat16* dz, const __nv_bfloat16* z,
                                 const float* ms, const float* mh, BnBranchDev b0, BnBranchDev b1, long long rows, int C) {
  const int rpi8 = (kBlock / (C / 8) > 0 ? kBlock / (C / 8) : 1) * 8;
#define PMD_RED(M)                                                                                                  \
  return launch_pdl(bn_bwd_reduce_multi_kernel<NB, M>,                                                              \
                    dim3(wave_grid(bn_bwd_reduce_multi_kernel<NB, M>, kBlock, smem, rows, rpi8)), dim3(kBlock), smem, \
                    st, dz, z, ms, mh, b0, b1, rows, C)
  if (mask == 0) PMD_RED(0);
  if (mask == 1) PMD_RED(1);
  if (mask == 3) PMD_RED(3);
  PMD_RED(2);
#undef PMD_RED
}

extern "C" PMD_API int pmd_bn_bwd_reduce_multi(const void* d_dz, const void* d_z, const float* d_mask_scale,
                                               const float* d_mask_shift, int mask_mode, const PmdBnBranch* br, int nb,
                                               long long rows, int C, void* stream) {
  BN_CHECK_C(C);
  PMD_REQUIRE(nb == 1 || nb == 2, "1 or 2 BatchNorm branches");
  PMD_REQUIRE(mask_mode >= 0 && mask_mode <= 3, "mask mode must be 0, 1, 2 or 3");
  PMD_REQUIRE((mask_mode != 1 && mask_mode != 3) || d_z != nullptr, "mask modes 1 and 3 need z (mode 3: the ReLU bit array)");
  PMD_REQUIRE(mask_mode != 2 || (d_mask_scale != nullptr && d_mask_shift != nullptr), "mask mode 2 needs the forward affine");
  const int groups = C / 8, rpi = kBlock / groups > 0 ? kBlock / groups : 1;
  PMD_REQUIRE(groups <= kBlock, "C too large for bn_bwd_reduce");
  const int smem = (2 * C + (1 + nb) * rpi * C) * 4;
  const BnBranchDev b0 = to_dev(br[0]), b1 = to_dev(br[nb - 1]);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* dz = static_cast<const __nv_bfloat16*>(d_dz);
  const __nv_bfloat16* z = static_cast<const __nv_bfloat16*>(d_z);
  if (nb == 1)
    PMD_CHECK_CUDA(launch_reduce<1>(mask_mode, smem, st, dz, z, d_mask_scale, d_mask_shift, b0, b1, rows, C));
  else
    PMD_CHECK_CUDA(launch_reduce<2>(mask_mode, smem, st, dz, z, d_mask_scale, d_mask_shift, b0, b1, rows, C));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

template <int NB>
static cudaError_t launch_apply(int mask, int smem, cudaStream_t st, const __nv_bfloat16* dz, const __nv_bfloat16* z,
                                const float* ms, const float* mh, BnBranchDev b0, BnBranchDev b1, float count,
                                __nv_bfloat16* g_out, long long total8, int C) {
  static PerDevice<bool> ready;       // > 48 KB of dynamic shared memory must be allowed per kernel, per device
  bool* once = ready.get();
  if (once == nullptr) return cudaErrorInvalidDevice;
  if (!*once) {
    cudaFuncSetAttribute(bn_bwd_apply_multi_kernel<NB, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    cudaFuncSetAttribute(bn_bwd_apply_multi_kernel<NB, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    cudaFuncSetAttribute(bn_bwd_apply_multi_kernel<NB, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    cudaFuncSetAttribute(bn_bwd_apply_multi_kernel<NB, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    *once = true;
  }
  const int blk = fixed_block(C);
#define PMD_APP(M)                                                                                                     \
  return launch_pdl(bn_bwd_apply_multi_kernel<NB, M>,                                                                  \
                    dim3(wave_grid(bn_bwd_apply_multi_kernel<NB, M>, blk, smem, total8, blk * 2)), dim3(blk), smem, st, \
                    dz, z, ms, mh, b0, b1, count, g_out, total8, C)
  if (mask == 0) PMD_APP(0);
  if (mask == 1) PMD_APP(1);
  if (mask == 3) PMD_APP(3);
  PMD_APP(2);
#undef PMD_APP
}

extern "C" PMD_API int pmd_bn_bwd_apply_multi(const void* d_dz, const void* d_z, const float* d_mask_scale,
                                              const float* d_mask_shift, int mask_mode, const PmdBnBranch* br, int nb,
                                              float count_total, void* d_g_out, long long rows, int C, void* stream) {
  BN_CHECK_C(C);
  PMD_REQUIRE(nb == 1 || nb == 2, "1 or 2 BatchNorm branches");
  PMD_REQUIRE(mask_mode >= 0 && mask_mode <= 3, "mask mode must be 0, 1, 2 or 3");
  PMD_REQUIRE((mask_mode != 1 && mask_mode != 3) || d_z != nullptr, "mask modes 1 and 3 need z (mode 3: the ReLU bit array)");
  const long long total8 = rows * C / 8;
  const BnBranchDev b0 = to_dev(br[0]), b1 = to_dev(br[nb - 1]);
  const int smem = (nb * 3 + 2) * C * 4;
  PMD_REQUIRE(smem <= 128 * 1024, "too many channels for the fused BatchNorm backward");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* dz = static_cast<const __nv_bfloat16*>(d_dz);
  const __nv_bfloat16* z = static_cast<const __nv_bfloat16*>(d_z);
  __nv_bfloat16* g = static_cast<__nv_bfloat16*>(d_g_out);
  if (nb == 1)
    PMD_CHECK_CUDA(launch_apply<1>(mask_mode, smem, st, dz, z, d_mask_scale, d_mask_shift, b0, b1, count_total, g, total8, C));
  else
    PMD_CHECK_CUDA(launch_apply<2>(mask_mode, smem, st, dz, z, d_mask_scale, d_mask_shift, b0, b1, count_total, g, total8, C));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

// Forward: finalize + apply in one launch (single-rank statistics).
template <int RES>
static cudaError_t launch_finalize_apply(int smem, cudaStream_t st, const BnFinalizeDev& f0, const BnFinalizeDev& f1,
                                         float count, float momentum, float eps, const __nv_bfloat16* y,
                                         const __nv_bfloat16* res, int relu, __nv_bfloat16* z, unsigned char* bits,
                                         long long total8, int C) {
  static PerDevice<bool> ready;
  bool* once = ready.get();
  if (once == nullptr) return cudaErrorInvalidDevice;
  if (!*once) {
    cudaFuncSetAttribute(bn_finalize_apply_kernel<RES>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    *once = true;
  }
  const int blk = fixed_block(C);
  return launch_pdl(bn_finalize_apply_kernel<RES>, dim3(wave_grid(bn_finalize_apply_kernel<RES>, blk, smem, total8, blk * 2)),
                    dim3(blk), smem, st, f0, f1, count, momentum, eps, y, res, relu, z, bits, total8, C);
}

static BnFinalizeDev to_dev(const PmdBnFinalize& f) {
  BnFinalizeDev d;
  d.stats = f.stats;
  d.gamma = f.gamma;
  d.beta = f.beta;
  d.running_mean = f.running_mean;
  d.running_var = f.running_var;
  d.nbt = f.num_batches;
  d.scale = f.scale;
  d.shift = f.shift;
  d.mean = f.mean;
  d.invstd = f.invstd;
  d.pivot = f.pivot;
  d.pivot_new = f.pivot_new;
  return d;
}

extern "C" PMD_API int pmd_bn_finalize_apply(const PmdBnFinalize* f, const PmdBnFinalize* f_res, float count,
                                             float momentum, float eps, const void* d_y, const void* d_res, int relu,
                                             void* d_z, long long rows, int C, void* stream) {
  return pmd_bn_finalize_apply_bits(f, f_res, count, momentum, eps, d_y, d_res, relu, d_z, nullptr, rows, C, stream);
}

extern "C" PMD_API int pmd_bn_finalize_apply_bits(const PmdBnFinalize* f, const PmdBnFinalize* f_res, float count,
                                                  float momentum, float eps, const void* d_y, const void* d_res, int relu,
                                                  void* d_z, void* d_relu_bits, long long rows, int C, void* stream) {
  unsigned char* bits = static_cast<unsigned char*>(d_relu_bits);
  BN_CHECK_C(C);
  PMD_REQUIRE(f != nullptr, "missing BatchNorm description");
  PMD_REQUIRE(f_res == nullptr || d_res != nullptr, "a residual BatchNorm needs the residual tensor");
  const long long total8 = rows * C / 8;
  const int smem = 6 * C * 4;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* y = static_cast<const __nv_bfloat16*>(d_y);
  const __nv_bfloat16* res = static_cast<const __nv_bfloat16*>(d_res);
  __nv_bfloat16* z = static_cast<__nv_bfloat16*>(d_z);
  const BnFinalizeDev f0 = to_dev(*f), f1 = to_dev(f_res != nullptr ? *f_res : *f);
  const int resmode = f_res != nullptr ? 2 : (d_res != nullptr ? 1 : 0);
  if (resmode == 0)
    PMD_CHECK_CUDA(launch_finalize_apply<0>(smem, st, f0, f1, count, momentum, eps, y, res, relu, z, bits, total8, C));
  else if (resmode == 1)
    PMD_CHECK_CUDA(launch_finalize_apply<1>(smem, st, f0, f1, count, momentum, eps, y, res, relu, z, bits, total8, C));
  else
    PMD_CHECK_CUDA(launch_finalize_apply<2>(smem, st, f0, f1, count, momentum, eps, y, res, relu, z, bits, total8, C));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

// single-branch forms kept for callers that have one BatchNorm at a time
extern "C" PMD_API int pmd_bn_bwd_reduce(const void* d_dz, const void* d_z, const void* d_y, const float* d_mean,
                                         const float* d_invstd, int relu, float* d_sums, long long rows, int C,
                                         void* stream) {
  PmdBnBranch b{};
  b.y = d_y;
  b.mean = d_mean;
  b.invstd = d_invstd;
  b.sums = d_sums;
  return pmd_bn_bwd_reduce_multi(d_dz, d_z, nullptr, nullptr, relu ? 1 : 0, &b, 1, rows, C, stream);
}

// Totals over `world` ranks are summed on the fly from a pointer table (kept for the 1-GPU multi-table tests).
namespace pmd {
namespace {
__global__ void sum_tables_kernel(const float* const* tabs, int world, int n, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float acc = 0.f;
  for (int w = 0; w < world; ++w) acc += tabs[w][i];
  out[i] = acc;
}
}  // namespace
}  // namespace pmd

namespace {
struct BwdApplyScratch {
  pmd::PerDevice<float*> slots;
  ~BwdApplyScratch() {
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess) return;     // runtime already torn down at process exit: nothing to free
    for (int d = 0; d < 64; ++d)
      if (slots.slot[d] != nullptr && cudaSetDevice(d) == cudaSuccess) cudaFree(slots.slot[d]);
    cudaSetDevice(cur);
  }
} g_bwd_apply_scratch;
}  // namespace

extern "C" PMD_API int pmd_bn_bwd_apply(const void* d_dz, const void* d_z, const void* d_y, const float* d_mean,
                                        const float* d_invstd, const float* d_gamma, const float* const* d_sums_ranks,
                                        int world, const float* d_local_sums, float count_total, int relu, void* d_dy,
                                        float* d_dgamma, float* d_dbeta, void* d_g_out, long long rows, int C,
                                        void* stream) {
  BN_CHECK_C(C);
  // [2 * 2048] scratch for the table sum: one buffer per device, allocated on that device's first call and
  // released when the library is unloaded
  float** slot = g_bwd_apply_scratch.slots.get();
  if (slot == nullptr) return -1;
  if (*slot == nullptr) PMD_CHECK_CUDA(cudaMalloc(slot, 2 * 2048 * sizeof(float)));
  float* totals = *slot;
  sum_tables_kernel<<<ceil_div(2 * C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_sums_ranks, world, 2 * C, totals);
  count_launch();
  PmdBnBranch b{};
  b.y = d_y;
  b.mean = d_mean;
  b.invstd = d_invstd;
  b.gamma = d_gamma;
  b.sums = const_cast<float*>(d_local_sums);
  b.totals = totals;
  b.dy = d_dy;
  b.dgamma = d_dgamma;
  b.dbeta = d_dbeta;
  return pmd_bn_bwd_apply_multi(d_dz, d_z, nullptr, nullptr, relu ? 1 : 0, &b, 1, count_total, d_g_out, rows, C, stream);
}

// ---------------------------------------------------------------------------------------------
// pmd_bn_bwd_apply_static: the same computation as pmd_bn_bwd_apply for hosts that must not allocate
// device memory.  pmd_bn_bwd_apply sums the per-rank [2C] tables into a scratch vector that it obtains
// from cudaMalloc on first use.  Some programs that want this kernel cannot afford that call: an
// application that has handed the whole GPU to its own memory pool (there is nothing left for the CUDA
// allocator to give), or a control loop that may not take the device-wide synchronisation cudaMalloc
// implies.  For them the scratch lives in the library image instead: a file-scope __device__ array of
// 2 * kStaticMaxC floats (8 KB), present on every GPU that loads libpmdist.so until the process exits.
//
//   * Limit: C <= 1024 channels; larger C is refused with an error, nothing is launched.
//   * The device address of the array is looked up with cudaGetSymbolAddress the first time the entry
//     point runs on a given device (that device being current).  Make that first call during start-up,
//     before any stream capture is begun; later calls only launch kernels.
//   * No cudaMalloc / cudaFree / allocating library call is made on any path through this function.
//   * All calls on a device share the one array and it keeps whatever the previous call left in it.
//     Nothing in it has to start at zero: the first kernel overwrites entries [0, 2C) and the second
//     reads only those.  Two calls whose kernels can overlap on the GPU (different streams without an
//     ordering between them) would race on the array -- serialise them.
namespace pmd {
namespace {
constexpr int kStaticMaxC = 1024;
constexpr int kStaticMaxDevices = 64;
static __device__ __align__(256) float g_bn_totals_static[2 * kStaticMaxC];
float* g_bn_totals_static_addr[kStaticMaxDevices] = {};   // per device: where the array above lives
}  // namespace
}  // namespace pmd

extern "C" PMD_API int pmd_bn_bwd_apply_static(const void* d_dz, const void* d_z, const void* d_y, const float* d_mean,
                                               const float* d_invstd, const float* d_gamma,
                                               const float* const* d_sums_ranks, int world, const float* d_local_sums,
                                               float count_total, int relu, void* d_dy, float* d_dgamma, float* d_dbeta,
                                               void* d_g_out, long long rows, int C, void* stream) {
  if (C > kStaticMaxC) {
    set_error("pmd_bn_bwd_apply_static: C = " + std::to_string(C) + " exceeds the built-in scratch (at most " +
              std::to_string(kStaticMaxC) + " channels); use pmd_bn_bwd_apply, which allocates its scratch");
    return -1;
  }
  BN_CHECK_C(C);
  PMD_REQUIRE(world >= 1, "world must be >= 1");
  int dev = -1;
  PMD_CHECK_CUDA(cudaGetDevice(&dev));
  PMD_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "device ordinal outside the static table");
  if (g_bn_totals_static_addr[dev] == nullptr) {
    // In host code the array's name is only a handle for the runtime, not a device pointer: resolve it once
    // per device, with that device current, and hand the resolved address to the kernels.
    float* ptr = nullptr;
    PMD_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_bn_totals_static));
    PMD_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address");
    g_bn_totals_static_addr[dev] = ptr;
  }
  float* totals = g_bn_totals_static_addr[dev];
  sum_tables_kernel<<<ceil_div(2 * C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(d_sums_ranks, world, 2 * C, totals);
  count_launch();
  PMD_LAUNCH_CHECK();
  PmdBnBranch b{};
  b.y = d_y;
  b.mean = d_mean;
  b.invstd = d_invstd;
  b.gamma = d_gamma;
  b.sums = const_cast<float*>(d_local_sums);
  b.totals = totals;
  b.dy = d_dy;
  b.dgamma = d_dgamma;
  b.dbeta = d_dbeta;
  return pmd_bn_bwd_apply_multi(d_dz, d_z, nullptr, nullptr, relu ? 1 : 0, &b, 1, count_total, d_g_out, rows, C, stream);
}

