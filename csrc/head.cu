// Network head and loss: global average pool + Linear, cross-entropy with top-1 / top-5 counters,
// and their backward.  Replaces avg_pool2d + view + nn.Linear (cuBLAS addmm / mm),
// nn.CrossEntropyLoss and utils.accuracy (ref:model/resnet.py:102-104; ref:main.py:48,105,112-114;
// ref:utils.py:64-77).  Metrics accumulate on the device (double[4]) and are
// read once per print interval instead of two .item() syncs per step.
#include "common.cuh"
#include "../include/pmdist.h"

namespace pmd {
namespace {

constexpr int kFcSplit = 16;

__global__ void pool_fc_fwd_kernel(const __nv_bfloat16* __restrict__ z, int HW, int C, const float* __restrict__ w,
                                   const float* __restrict__ b, int classes, float* __restrict__ feat,
                                   float* __restrict__ logits) {
  pdl_wait();
  extern __shared__ float fs[];   // [C]
  const int n = blockIdx.x;
  const __nv_bfloat16* zn = z + static_cast<long long>(n) * HW * C;
  const float inv = 1.f / HW;
  for (int c2 = threadIdx.x; c2 < C / 2; c2 += blockDim.x) {
    float s0 = 0.f, s1 = 0.f;
    for (int p = 0; p < HW; ++p) {
      const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(zn + static_cast<long long>(p) * C + 2 * c2));
      s0 += f.x;
      s1 += f.y;
    }
    fs[2 * c2] = s0 * inv;
    fs[2 * c2 + 1] = s1 * inv;
    feat[static_cast<long long>(n) * C + 2 * c2] = s0 * inv;
    feat[static_cast<long long>(n) * C + 2 * c2 + 1] = s1 * inv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int j = warp; j < classes; j += nwarps) {
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc += fs[c] * __ldg(w + static_cast<long long>(j) * C + c);
    acc = warp_sum(acc);
    if (lane == 0) logits[static_cast<long long>(n) * classes + j] = acc + b[j];
  }
}

// Large final maps (224x224 inputs: 28x28x2048 per image): one block per image walking HW pixels serially leaves the
// GPU idle (32 blocks).  Split: pool_partial_kernel sums slices of the map with many blocks (atomics into the zeroed
// feature vector), fc_kernel finishes.
__global__ void pool_partial_kernel(const __nv_bfloat16* __restrict__ z, int HW, int C, int slices, float* __restrict__ feat) {
  pdl_wait();
  const int n = blockIdx.x / slices, sl = blockIdx.x % slices;
  const int p0 = static_cast<int>(static_cast<long long>(HW) * sl / slices), p1 = static_cast<int>(static_cast<long long>(HW) * (sl + 1) / slices);
  const __nv_bfloat16* zn = z + static_cast<long long>(n) * HW * C;
  const float inv = 1.f / HW;
  for (int c2 = threadIdx.x; c2 < C / 2; c2 += blockDim.x) {
    float s0 = 0.f, s1 = 0.f;
    for (int p = p0; p < p1; ++p) {
      const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(zn + static_cast<long long>(p) * C + 2 * c2));
      s0 += f.x;
      s1 += f.y;
    }
    atomicAdd(feat + static_cast<long long>(n) * C + 2 * c2, s0 * inv);
    atomicAdd(feat + static_cast<long long>(n) * C + 2 * c2 + 1, s1 * inv);
  }
}
__global__ void fc_kernel(const float* __restrict__ feat, int C, const float* __restrict__ w, const float* __restrict__ b,
                          int classes, float* __restrict__ logits) {
  pdl_wait();
  const int n = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int j = warp; j < classes; j += nwarps) {
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc += feat[static_cast<long long>(n) * C + c] * __ldg(w + static_cast<long long>(j) * C + c);
    acc = warp_sum(acc);
    if (lane == 0) logits[static_cast<long long>(n) * classes + j] = acc + b[j];
  }
}

// one warp per sample
__global__ void ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                          const unsigned char* __restrict__ valid, int N, int classes, float grad_scale,
                          float* __restrict__ dlogits, double* metrics) {
  pdl_wait();
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  __shared__ float acc_s[4];
  if (threadIdx.x < 4) acc_s[threadIdx.x] = 0.f;
  __syncthreads();
  if (warp_global < N) {
    const int n = warp_global;
    const float* l = logits + static_cast<long long>(n) * classes;
    const int label = static_cast<int>(labels[n]);
    const bool ok = valid == nullptr || valid[n] != 0;
    float mx = -INFINITY;
    for (int j = lane; j < classes; j += 32) mx = fmaxf(mx, l[j]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float se = 0.f;
    for (int j = lane; j < classes; j += 32) se += __expf(l[j] - mx);
    se = warp_sum(se);
    const float ll = l[label];
    float above = 0.f;       // logits that beat the label's (ties broken towards the lower index, like topk)
    for (int j = lane; j < classes; j += 32) above += (l[j] > ll || (l[j] == ll && j < label)) ? 1.f : 0.f;
    above = warp_sum(above);
    const float lse = mx + __logf(se);
    if (dlogits != nullptr) {
      for (int j = lane; j < classes; j += 32) {
        const float pj = __expf(l[j] - lse);
        dlogits[static_cast<long long>(n) * classes + j] = ok ? (pj - (j == label ? 1.f : 0.f)) * grad_scale : 0.f;
      }
    }
    if (lane == 0 && ok) {
      atomicAdd(&acc_s[0], lse - ll);
      atomicAdd(&acc_s[1], above < 0.5f ? 1.f : 0.f);
      atomicAdd(&acc_s[2], above < 4.5f ? 1.f : 0.f);
      atomicAdd(&acc_s[3], 1.f);
    }
  }
  __syncthreads();
  if (threadIdx.x < 4 && metrics != nullptr) atomicAdd(metrics + threadIdx.x, static_cast<double>(acc_s[threadIdx.x]));
}

__global__ void pool_fc_bwd_kernel(const float* __restrict__ dlogits, const float* __restrict__ feat,
                                   const float* __restrict__ w, int N, int HW, int C, int classes, float* dw,
                                   float* db, __nv_bfloat16* __restrict__ dz) {
  pdl_wait();
  if (static_cast<int>(blockIdx.x) < N) {
    // d(feature) -> broadcast over the pooled window
    extern __shared__ float dls[];   // [classes]
    const int n = blockIdx.x;
    for (int j = threadIdx.x; j < classes; j += blockDim.x) dls[j] = dlogits[static_cast<long long>(n) * classes + j];
    __syncthreads();
    const float inv = 1.f / HW;
    for (int c2 = threadIdx.x; c2 < C / 2; c2 += blockDim.x) {
      float d0 = 0.f, d1 = 0.f;
      for (int j = 0; j < classes; ++j) {
        d0 += dls[j] * __ldg(w + static_cast<long long>(j) * C + 2 * c2);
        d1 += dls[j] * __ldg(w + static_cast<long long>(j) * C + 2 * c2 + 1);
      }
      const uint32_t pk = pack_bf16x2(d0 * inv, d1 * inv);
      for (int p = 0; p < HW; ++p)
        *reinterpret_cast<uint32_t*>(dz + (static_cast<long long>(n) * HW + p) * C + 2 * c2) = pk;
    }
  } else {
    // dW[j][c] = sum_n dlogits[n][j] * feat[n][c]; the sample range is split kFcSplit ways for parallelism
    const int bid = blockIdx.x - N;
    const int chunk = bid / kFcSplit, part = bid % kFcSplit;
    const long long o = static_cast<long long>(chunk) * blockDim.x + threadIdx.x;
    if (o >= static_cast<long long>(classes) * C) return;
    const int j = static_cast<int>(o / C), c = static_cast<int>(o % C);
    const int per = (N + kFcSplit - 1) / kFcSplit;
    const int n0 = part * per, n1 = min(n0 + per, N);
    float acc = 0.f, accb = 0.f;
#pragma unroll 4
    for (int n = n0; n < n1; ++n) {
      const float d = dlogits[static_cast<long long>(n) * classes + j];
      acc += d * feat[static_cast<long long>(n) * C + c];
      accb += d;
    }
    atomicAdd(dw + o, acc);
    if (c == 0) atomicAdd(db + j, accb);
  }
}

}  // namespace
}  // namespace pmd

using namespace pmd;

extern "C" PMD_API int pmd_pool_fc_fwd(const void* d_z, int N, int HW, int C, const float* d_w, const float* d_b, int classes,
                               float* d_feat, float* d_logits, void* stream) {
  PMD_REQUIRE(C % 2 == 0, "feature width must be even");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long work = static_cast<long long>(HW) * C;
  if (HW >= 64 && N * 8 < 148 * 4 && work >= (1 << 18)) {
    // few images with large final maps: pool with (N x slices) blocks, then the tiny FC
    int slices = (148 * 4 + N - 1) / N;
    if (slices > HW / 8) slices = HW / 8;
    PMD_CHECK_CUDA(cudaMemsetAsync(d_feat, 0, static_cast<size_t>(N) * C * sizeof(float), st));
    PMD_CHECK_CUDA(launch_pdl(pool_partial_kernel, dim3(N * slices), dim3(256), 0, st, static_cast<const __nv_bfloat16*>(d_z), HW,
                              C, slices, d_feat));
    PMD_CHECK_CUDA(launch_pdl(fc_kernel, dim3(N), dim3(256), 0, st, static_cast<const float*>(d_feat), C, d_w, d_b, classes,
                              d_logits));
    count_launch(2);
    PMD_LAUNCH_CHECK();
    return 0;
  }
  PMD_CHECK_CUDA(launch_pdl(pool_fc_fwd_kernel, dim3(N), dim3(256), C * 4, st, static_cast<const __nv_bfloat16*>(d_z), HW, C, d_w,
                            d_b, classes, d_feat, d_logits));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_ce_fwd_bwd(const float* d_logits, const long long* d_labels, const unsigned char* d_valid, int N,
                              int classes, float grad_scale, float* d_dlogits, double* d_metrics, void* stream) {
  PMD_CHECK_CUDA(launch_pdl(ce_kernel, dim3(ceil_div(N * 32, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), d_logits,
                            d_labels, d_valid, N, classes, grad_scale, d_dlogits, d_metrics));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_pool_fc_bwd(const float* d_dlogits, const float* d_feat, const float* d_w, int N, int HW, int C,
                               int classes, float* d_dw, float* d_db, void* d_dz, void* stream) {
  PMD_REQUIRE(C % 2 == 0, "feature width must be even");
  const int extra = ceil_div(classes * C, 256) * kFcSplit;
  PMD_CHECK_CUDA(launch_pdl(pool_fc_bwd_kernel, dim3(N + extra), dim3(256), classes * 4, static_cast<cudaStream_t>(stream),
                            d_dlogits, d_feat, d_w, N, HW, C, classes, d_dw, d_db, static_cast<__nv_bfloat16*>(d_dz)));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}
