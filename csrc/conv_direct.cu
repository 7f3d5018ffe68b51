// Direct (SIMT) convolutions: the C=3 stem of the CIFAR ResNet (ref:model/resnet.py:79-80), whose
// K = 27 reduction is too thin for a 64-wide tensor-core k-block, and an in-library cross-check
// for the tcgen05 kernels on small shapes.  Layouts as in conv_tc.cu: x NHWC bf16 (with an optional
// channel stride for the RGB->4 padded input), w KRSC bf16, dw KRSC fp32 (accumulated).
#include "common.cuh"
#include "../include/pmdist.h"

namespace pmd {
namespace {

__global__ void fprop_direct_kernel(PmdConvShape s, int xcs, const __nv_bfloat16* __restrict__ x,
                                    const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y, int Ho, int Wo,
                                    int w_in_smem) {
  extern __shared__ float wsm[];   // [K][R*R*C] when it fits
  const int RSC = s.R * s.R * s.C, pad = s.R / 2;
  if (w_in_smem) {
    for (int i = threadIdx.x; i < s.K * RSC; i += blockDim.x) wsm[i] = __bfloat162float(w[i]);
    __syncthreads();
  }
  const long long total = static_cast<long long>(s.N) * Ho * Wo * s.K;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(idx % s.K);
    long long pix = idx / s.K;
    const int wo = static_cast<int>(pix % Wo);
    pix /= Wo;
    const int ho = static_cast<int>(pix % Ho);
    const int n = static_cast<int>(pix / Ho);
    float acc = 0.f;
    for (int r = 0; r < s.R; ++r) {
      const int h = ho * s.stride - pad + r;
      if (h < 0 || h >= s.H) continue;
      for (int q = 0; q < s.R; ++q) {
        const int wc = wo * s.stride - pad + q;
        if (wc < 0 || wc >= s.W) continue;
        const __nv_bfloat16* xp = x + ((static_cast<long long>(n) * s.H + h) * s.W + wc) * xcs;
        const int wbase = (k * s.R * s.R + r * s.R + q) * s.C;
        if (w_in_smem) {
          for (int c = 0; c < s.C; ++c) acc += __bfloat162float(xp[c]) * wsm[wbase + c];
        } else {
          for (int c = 0; c < s.C; ++c) acc += __bfloat162float(xp[c]) * __bfloat162float(w[wbase + c]);
        }
      }
    }
    y[idx] = __float2bfloat16_rn(acc);
  }
}

__global__ void dgrad_direct_kernel(PmdConvShape s, const __nv_bfloat16* __restrict__ dy,
                                    const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ dx, int Ho,
                                    int Wo) {
  const int pad = s.R / 2;
  const long long total = static_cast<long long>(s.N) * s.H * s.W * s.C;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(idx % s.C);
    long long pix = idx / s.C;
    const int wc = static_cast<int>(pix % s.W);
    pix /= s.W;
    const int h = static_cast<int>(pix % s.H);
    const int n = static_cast<int>(pix / s.H);
    float acc = 0.f;
    for (int r = 0; r < s.R; ++r) {
      const int nh = h + pad - r;
      if (nh < 0 || nh % s.stride != 0) continue;
      const int ho = nh / s.stride;
      if (ho >= Ho) continue;
      for (int q = 0; q < s.R; ++q) {
        const int nw = wc + pad - q;
        if (nw < 0 || nw % s.stride != 0) continue;
        const int wo = nw / s.stride;
        if (wo >= Wo) continue;
        const __nv_bfloat16* dyp = dy + ((static_cast<long long>(n) * Ho + ho) * Wo + wo) * s.K;
        for (int k = 0; k < s.K; ++k)
          acc += __bfloat162float(dyp[k]) *
                 __bfloat162float(w[((static_cast<long long>(k) * s.R + r) * s.R + q) * s.C + c]);
      }
    }
    dx[idx] = __float2bfloat16_rn(acc);
  }
}

// Small-filter wgrad (K * R*R*C <= 4096 outputs): a block owns 64 output pixels, stages dy and the
// input patches in shared memory and adds its partial dW with one atomic per output.
constexpr int kWgPix = 64;
__global__ void wgrad_small_kernel(PmdConvShape s, int xcs, const __nv_bfloat16* __restrict__ x,
                                   const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw, int Ho, int Wo) {
  extern __shared__ float sm[];
  const int RSC = s.R * s.R * s.C, pad = s.R / 2;
  float* dys = sm;                    // [kWgPix][K]
  float* xps = sm + kWgPix * s.K;     // [kWgPix][RSC]
  const long long npix = static_cast<long long>(s.N) * Ho * Wo;
  const long long p0 = static_cast<long long>(blockIdx.x) * kWgPix;
  for (int i = threadIdx.x; i < kWgPix * s.K; i += blockDim.x) {
    const long long pix = p0 + i / s.K;
    dys[i] = pix < npix ? __bfloat162float(dy[pix * s.K + i % s.K]) : 0.f;
  }
  for (int i = threadIdx.x; i < kWgPix * RSC; i += blockDim.x) {
    const long long pix = p0 + i / RSC;
    const int j = i % RSC;
    float v = 0.f;
    if (pix < npix) {
      const int c = j % s.C, q = (j / s.C) % s.R, r = j / (s.C * s.R);
      const int wo = static_cast<int>(pix % Wo), ho = static_cast<int>((pix / Wo) % Ho);
      const int n = static_cast<int>(pix / (static_cast<long long>(Wo) * Ho));
      const int h = ho * s.stride - pad + r, wc = wo * s.stride - pad + q;
      if (h >= 0 && h < s.H && wc >= 0 && wc < s.W)
        v = __bfloat162float(x[((static_cast<long long>(n) * s.H + h) * s.W + wc) * xcs + c]);
    }
    xps[i] = v;
  }
  __syncthreads();
  for (int o = threadIdx.x; o < s.K * RSC; o += blockDim.x) {
    const int k = o / RSC, j = o % RSC;
    float acc = 0.f;
#pragma unroll 8
    for (int p = 0; p < kWgPix; ++p) acc += dys[p * s.K + k] * xps[p * RSC + j];
    atomicAdd(dw + o, acc);
  }
}

// Generic wgrad: one warp per dW element (cross-check only; slow).
__global__ void wgrad_generic_kernel(PmdConvShape s, int xcs, const __nv_bfloat16* __restrict__ x,
                                     const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw, int Ho, int Wo) {
  const int RSC = s.R * s.R * s.C, pad = s.R / 2;
  const long long o = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
  const int lane = threadIdx.x & 31;
  if (o >= static_cast<long long>(s.K) * RSC) return;
  const int k = static_cast<int>(o / RSC), j = static_cast<int>(o % RSC);
  const int c = j % s.C, q = (j / s.C) % s.R, r = j / (s.C * s.R);
  const long long npix = static_cast<long long>(s.N) * Ho * Wo;
  float acc = 0.f;
  for (long long pix = lane; pix < npix; pix += 32) {
    const int wo = static_cast<int>(pix % Wo), ho = static_cast<int>((pix / Wo) % Ho);
    const int n = static_cast<int>(pix / (static_cast<long long>(Wo) * Ho));
    const int h = ho * s.stride - pad + r, wc = wo * s.stride - pad + q;
    if (h < 0 || h >= s.H || wc < 0 || wc >= s.W) continue;
    acc += __bfloat162float(dy[pix * s.K + k]) *
           __bfloat162float(x[((static_cast<long long>(n) * s.H + h) * s.W + wc) * xcs + c]);
  }
  acc = warp_sum(acc);
  if (lane == 0) atomicAdd(dw + o, acc);
}

}  // namespace
}  // namespace pmd

using namespace pmd;

static int out_dim(int in, int R, int stride) { return (in + 2 * (R / 2) - R) / stride + 1; }

extern "C" PMD_API int pmd_conv_fprop_direct(const PmdConvShape* s, int x_cstride, const void* d_x, const void* d_w_krsc,
                                     void* d_y, void* stream) {
  const int Ho = out_dim(s->H, s->R, s->stride), Wo = out_dim(s->W, s->R, s->stride);
  const long long total = static_cast<long long>(s->N) * Ho * Wo * s->K;
  const int wbytes = s->K * s->R * s->R * s->C * 4;
  const int in_smem = wbytes <= 40 * 1024;
  const int blocks = static_cast<int>(std::min<long long>(ceil_div_ll(total, 256), 148LL * 16));
  fprop_direct_kernel<<<blocks, 256, in_smem ? wbytes : 0, static_cast<cudaStream_t>(stream)>>>(
      *s, x_cstride, static_cast<const __nv_bfloat16*>(d_x), static_cast<const __nv_bfloat16*>(d_w_krsc),
      static_cast<__nv_bfloat16*>(d_y), Ho, Wo, in_smem);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_conv_dgrad_direct(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, void* d_dx,
                                     void* stream) {
  const int Ho = out_dim(s->H, s->R, s->stride), Wo = out_dim(s->W, s->R, s->stride);
  const long long total = static_cast<long long>(s->N) * s->H * s->W * s->C;
  const int blocks = static_cast<int>(std::min<long long>(ceil_div_ll(total, 256), 148LL * 16));
  dgrad_direct_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      *s, static_cast<const __nv_bfloat16*>(d_dy), static_cast<const __nv_bfloat16*>(d_w_krsc),
      static_cast<__nv_bfloat16*>(d_dx), Ho, Wo);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_conv_wgrad_direct(const PmdConvShape* s, int x_cstride, const void* d_x, const void* d_dy,
                                     float* d_dw_krsc, void* stream) {
  const int Ho = out_dim(s->H, s->R, s->stride), Wo = out_dim(s->W, s->R, s->stride);
  const int RSC = s->R * s->R * s->C;
  const long long npix = static_cast<long long>(s->N) * Ho * Wo;
  const long long outs = static_cast<long long>(s->K) * RSC;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (outs <= 4096 && (kWgPix * (s->K + RSC) * 4) <= 48 * 1024) {
    const int smem = kWgPix * (s->K + RSC) * 4;
    wgrad_small_kernel<<<static_cast<int>(ceil_div_ll(npix, kWgPix)), 256, smem, st>>>(
        *s, x_cstride, static_cast<const __nv_bfloat16*>(d_x), static_cast<const __nv_bfloat16*>(d_dy), d_dw_krsc, Ho,
        Wo);
  } else {
    wgrad_generic_kernel<<<static_cast<int>(ceil_div_ll(outs * 32, 256)), 256, 0, st>>>(
        *s, x_cstride, static_cast<const __nv_bfloat16*>(d_x), static_cast<const __nv_bfloat16*>(d_dy), d_dw_krsc, Ho,
        Wo);
  }
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}
