// BatchNorm finalize math shared by the single-GPU kernels (bn.cu) and the SyncBatchNorm exchange kernels (comm.cu).
#pragma once

namespace pmd {
namespace {

// What every BatchNorm finalize in this library computes for channel c from the global sums over n samples
// (torch: batch_norm_gather_stats_with_counts, torch/nn/modules/_functions.py:106-115).
// The sums are taken around a per-channel pivot P: s1 = sum(y - P), s2 = sum((y - P)^2), so
//   mean = P + s1/n,  var = s2/n - (s1/n)^2
// which, unlike E[y^2] - E[y]^2 in fp32, does not cancel when |mean| >> std as long as P is near the mean (torch
// avoids the same cancellation with Welford + count-weighted merges).  The producers use the PREVIOUS step's batch
// mean as P (0 on the first step); `pivot_new` receives this step's mean and becomes P after the backward pass.
__device__ __forceinline__ void bn_finalize_channel(int c, float s1, float s2, float n, const float* pivot,
                                                    float* pivot_new, const float* gamma, const float* beta,
                                                    float* running_mean, float* running_var, float momentum, float eps,
                                                    float* scale, float* shift, float* save_mean, float* save_invstd) {
  const float d = s1 / n;
  const float mean = (pivot != nullptr ? pivot[c] : 0.f) + d;
  const float var = fmaxf(s2 / n - d * d, 0.f);
  if (pivot_new != nullptr) pivot_new[c] = mean;
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
