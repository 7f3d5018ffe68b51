// BatchNorm finalize math shared by the single-GPU kernels (bn.cu) and the SyncBatchNorm exchange kernels (comm.cu).
#pragma once

namespace pmd {
namespace {

// What every BatchNorm finalize in this library computes for channel c from the global [sum, sumsq] over n samples
// (torch: batch_norm_gather_stats_with_counts, torch/nn/modules/_functions.py:106-115).
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

}  // namespace
}  // namespace pmd
