// Input pipeline on the GPU: zero-pad random crop + horizontal flip + ToTensor + Normalize(0.5, 0.5),
// uint8 NHWC RGB -> NHWC bf16 with the channel dim padded to `c_out`.  Replaces the PIL/torchvision
// transforms that run in 8 DataLoader worker processes per rank in the reference
// (ref:data.py:11-19, 41-53).  Two sources of a batch: (a) uint8 images + crop/flip parameters staged
// from the host (pmdist/data/loader.py), (b) a dataset that lives in device memory, gathered by an index vector
// with crop/flip drawn in the kernel from Philox(seed, epoch, dataset index).  The host can reproduce (b)'s draws
// (pmdist/data/philox.py), so both forms are bit-comparable with the CPU path.
#include "aug.cuh"
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

// Augmentation fused with the stem convolution's im2col: every output pixel carries the 3x3x3
// neighbourhood of the AUGMENTED image (tap-major, channel-minor: column = (r*3+s)*3 + c, the KRSC order
// of the stem weight) padded with zeros to 64 columns.  The stem conv (ref:model/resnet.py:79-80) then is
// a 64-wide 1x1 convolution the tcgen05 path handles, and so is its weight gradient.  Positions outside
// the image are the convolution's zero padding (0), positions inside the image but outside the crop are
// torchvision's pre-normalisation zero pad (-1 after Normalize).  Only columns 0..31 of a pixel are written (27 values
// + 5 zeros); columns 32..63 must be zero in the caller's buffer and stay untouched.
template <bool GATHER>
__global__ void augment_im2col_kernel(const unsigned char* __restrict__ u8, const int* __restrict__ aug, GatherSrc g, int N,
                                      int H, int W, int pad, __nv_bfloat16* __restrict__ out) {
  const long long total = static_cast<long long>(N) * H * W;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x = static_cast<int>(i % W);
    const int y = static_cast<int>((i / W) % H);
    const int n = static_cast<int>(i / (static_cast<long long>(W) * H));
    int ox, oy, fl, src = n;
    if (GATHER) {
      gather_params(g, n, pad, &src, &ox, &oy, &fl);
      if (x == 0 && y == 0) {            // one thread per sample also hands label and validity to the loss kernel
        g.labels_out[n] = reinterpret_cast<const long long*>(g.ds[1])[src];
        g.valid_out[n] = g.idx[n] < 0 ? 0 : 1;
      }
    } else {
      ox = aug[n * 3];
      oy = aug[n * 3 + 1];
      fl = aug[n * 3 + 2];
    }
    const unsigned char* img =
        (GATHER ? reinterpret_cast<const unsigned char*>(g.ds[0]) : u8) + static_cast<long long>(src) * H * W * 3;
    float v[28];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int ay = y + r - 1, ax = x + q - 1;            // position in the augmented image
        float t0 = 0.f, t1 = 0.f, t2 = 0.f;
        if (ay >= 0 && ay < H && ax >= 0 && ax < W) {
          const int xc = fl ? (W - 1 - ax) : ax;
          const int sx = xc + ox - pad, sy = ay + oy - pad;   // position in the stored image
          t0 = t1 = t2 = -1.f;
          if (sx >= 0 && sx < W && sy >= 0 && sy < H) {
            const unsigned char* p = img + (static_cast<long long>(sy) * W + sx) * 3;
            t0 = (p[0] / 255.0f - 0.5f) / 0.5f;
            t1 = (p[1] / 255.0f - 0.5f) / 0.5f;
            t2 = (p[2] / 255.0f - 0.5f) / 0.5f;
          }
        }
        v[(r * 3 + q) * 3] = t0;
        v[(r * 3 + q) * 3 + 1] = t1;
        v[(r * 3 + q) * 3 + 2] = t2;
      }
    }
    v[27] = 0.f;
    uint4* o = reinterpret_cast<uint4*>(out + i * 64);
    o[0] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    o[1] = make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]),
                      pack_bf16x2(v[14], v[15]));
    o[2] = make_uint4(pack_bf16x2(v[16], v[17]), pack_bf16x2(v[18], v[19]), pack_bf16x2(v[20], v[21]),
                      pack_bf16x2(v[22], v[23]));
    o[3] = make_uint4(pack_bf16x2(v[24], v[25]), pack_bf16x2(v[26], v[27]), 0u, 0u);
    // columns 32..63 are never written: they are zero from the buffer's allocation on (half the store traffic)
  }
}

// [K][27] bf16 (KRSC stem weight) -> [K][64] bf16, zero padded.
__global__ void pad_stem_weight_kernel(const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ wp, int K,
                                       int rsc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * 64) return;
  const int k = i / 64, c = i % 64;
  wp[i] = c < rsc ? w[k * rsc + c] : __float2bfloat16_rn(0.f);
}

}  // namespace
}  // namespace pmd

using namespace pmd;

extern "C" PMD_API int pmd_augment_im2col(const unsigned char* d_u8, const int* d_aug, int N, int H, int W, int pad,
                                          void* d_out, void* stream) {
  const long long total = static_cast<long long>(N) * H * W;
  long long blocks = ceil_div_ll(total, 128);
  if (blocks > 148 * 32) blocks = 148 * 32;
  augment_im2col_kernel<false><<<static_cast<int>(blocks), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      d_u8, d_aug, GatherSrc{}, N, H, W, pad, static_cast<__nv_bfloat16*>(d_out));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_augment_im2col_gather(const unsigned long long* d_dataset_desc, const int* d_hdr,
                                                 const int* d_idx, int N, int H, int W, int pad, void* d_out,
                                                 long long* d_labels_out, unsigned char* d_valid_out, void* stream) {
  PMD_REQUIRE(d_dataset_desc != nullptr && d_hdr != nullptr && d_idx != nullptr,
              "dataset descriptor, header and index vector are required");
  const long long total = static_cast<long long>(N) * H * W;
  long long blocks = ceil_div_ll(total, 128);
  if (blocks > 148 * 32) blocks = 148 * 32;
  const GatherSrc g{d_hdr, d_idx, d_dataset_desc, d_labels_out, d_valid_out};
  augment_im2col_kernel<true><<<static_cast<int>(blocks), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      nullptr, nullptr, g, N, H, W, pad, static_cast<__nv_bfloat16*>(d_out));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

extern "C" PMD_API int pmd_pad_stem_weight(const void* d_w, void* d_w_padded, int K, int rsc, void* stream) {
  PMD_REQUIRE(rsc <= 64, "stem filter has more than 64 taps*channels");
  pad_stem_weight_kernel<<<ceil_div(K * 64, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(d_w), static_cast<__nv_bfloat16*>(d_w_padded), K, rsc);
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

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
