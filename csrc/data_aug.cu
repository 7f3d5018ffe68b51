// Input pipeline on the GPU: zero-pad random crop + horizontal flip + ToTensor + Normalize(0.5, 0.5),
// uint8 NHWC RGB -> NHWC bf16 with the channel dim padded to `c_out`.  Replaces the PIL/torchvision
// transforms that run in 8 DataLoader worker processes per rank in the reference
// (ref:data.py:11-19, 41-53).  Crop/flip parameters are drawn on the host
// (pmdist/data/loader.py) so this kernel is bit-comparable with the CPU path.
#include "common.cuh"
#include "../include/pmdist.h"

namespace pmd {
namespace {

__global__ void augment_kernel(const unsigned char* __restrict__ u8, const int* __restrict__ aug, int N, int H, int W,
                               int pad, int c_out, __nv_bfloat16* __restrict__ out) {
  const long long total = static_cast<long long>(N) * H * W;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x = static_cast<int>(i % W);
    const int y = static_cast<int>((i / W) % H);
    const int n = static_cast<int>(i / (static_cast<long long>(W) * H));
    const int ox = aug[n * 3], oy = aug[n * 3 + 1], fl = aug[n * 3 + 2];
    const int xc = fl ? (W - 1 - x) : x;
    const int sx = xc + ox - pad, sy = y + oy - pad;
    float v[3] = {0.f, 0.f, 0.f};
    if (sx >= 0 && sx < W && sy >= 0 && sy < H) {
      const unsigned char* p = u8 + ((static_cast<long long>(n) * H + sy) * W + sx) * 3;
      v[0] = p[0];
      v[1] = p[1];
      v[2] = p[2];
    }
    __nv_bfloat16* o = out + i * c_out;
#pragma unroll
    for (int c = 0; c < 3; ++c) o[c] = __float2bfloat16_rn((v[c] / 255.0f - 0.5f) / 0.5f);
    for (int c = 3; c < c_out; ++c) o[c] = __float2bfloat16_rn(0.f);
  }
}

}  // namespace
}  // namespace pmd

using namespace pmd;

extern "C" PMD_API int pmd_augment(const unsigned char* d_u8, const int* d_aug, int N, int H, int W, int pad, int c_out,
                           void* d_out, void* stream) {
  PMD_REQUIRE(c_out >= 3, "c_out must be >= 3");
  const long long total = static_cast<long long>(N) * H * W;
  long long blocks = ceil_div_ll(total, 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  augment_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_u8, d_aug, N, H, W, pad, c_out, static_cast<__nv_bfloat16*>(d_out));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}
