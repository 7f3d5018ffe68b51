// SGD (weight decay, momentum, Nesterov) on flat fp32 slabs with the bf16 working copy written in
// the same pass.  Replaces optimizer.zero_grad/step's _foreach
// kernels (ref:main.py:51-55, 108-110; math torch/optim/sgd.py:343-380).  The
// multi-GPU build folds this update into the gradient all-reduce kernel (comm.cu); this file is the
// single-GPU path and the shared device function.
#include "common.cuh"
#include "sgd_math.cuh"
#include "../include/pmdist.h"

namespace pmd {
namespace {

__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                           __nv_bfloat16* __restrict__ pb, long long n4, long long n, SgdHyper h) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    sgd_update(pv.x, gv.x, mv.x, h);
    sgd_update(pv.y, gv.y, mv.y, h);
    sgd_update(pv.z, gv.z, mv.z, h);
    sgd_update(pv.w, gv.w, mv.w, h);
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    if (pb != nullptr)
      reinterpret_cast<uint2*>(pb)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
  }
  // tail (n not a multiple of 4)
  const long long t = n4 * 4 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (t < n) {
    float pv = p[t], mv = m[t];
    sgd_update(pv, g[t], mv, h);
    p[t] = pv;
    m[t] = mv;
    if (pb != nullptr) pb[t] = __float2bfloat16_rn(pv);
  }
}


__global__ void cast_kernel(const float* __restrict__ s, __nv_bfloat16* __restrict__ d, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    d[i] = __float2bfloat16_rn(s[i]);
}

}  // namespace
}  // namespace pmd

using namespace pmd;

extern "C" PMD_API int pmd_sgd_step(float* d_param, const float* d_grad, float* d_mom, void* d_param_bf16, long long n,
                            float lr, float momentum, float weight_decay, int nesterov, float grad_scale, void* stream) {
  SgdHyper h{lr, momentum, weight_decay, grad_scale, nesterov};
  const long long n4 = n / 4;
  long long blocks = ceil_div_ll(n4 > 0 ? n4 : 1, 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  sgd_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_param, d_grad, d_mom, static_cast<__nv_bfloat16*>(d_param_bf16), n4, n, h);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}


extern "C" PMD_API int pmd_cast_f32_to_bf16(const float* d_src, void* d_dst, long long n, void* stream) {
  long long blocks = ceil_div_ll(n, 256 * 4);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  cast_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_src, static_cast<__nv_bfloat16*>(d_dst), n);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}
