// Pieces of the on-GPU input pipeline shared by csrc/data_aug.cu (augmentation + im2col into a buffer) and
// csrc/stem_fused.cu (the same patches built straight into the stem convolution's shared-memory operand).
#pragma once
#include "common.cuh"

namespace pmd {

// Philox4x32-10 (Salmon et al., SC'11): counter-based, so the crop/flip of a sample is a pure function of
// (seed, epoch, dataset index) -- no generator state, identical on the device and in the host oracle
// (pmdist/data/philox.py), and independent of the rank that happens to own the sample this epoch.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// Device-resident dataset: which stored image, and which crop / flip, sample n of the batch uses.
// hdr = {seed, stream, epoch, train}; idx[n] = dataset index, or ~index for a sampler wrap-around duplicate that
// must not count in the evaluation metrics.
struct GatherSrc {
  const int* hdr;
  const int* idx;
  const unsigned long long* ds;   // device-resident descriptor {images address, labels address}: read at run time, so
                                  // switching between resident datasets (train / test) needs no new CUDA graph
  long long* labels_out;       // [N] labels of this batch (read by the loss kernel)
  unsigned char* valid_out;    // [N]
};
__device__ __forceinline__ void gather_params(const GatherSrc& g, int n, int pad, int* src, int* ox, int* oy, int* fl) {
  const int raw = g.idx[n];
  const int id = raw < 0 ? ~raw : raw;
  *src = id;
  if (g.hdr[3] != 0) {
    const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(id), static_cast<uint32_t>(g.hdr[2]), 0u, 0u),
                                  make_uint2(static_cast<uint32_t>(g.hdr[0]), static_cast<uint32_t>(g.hdr[1])));
    *ox = static_cast<int>(r.x % static_cast<uint32_t>(2 * pad + 1));
    *oy = static_cast<int>(r.y % static_cast<uint32_t>(2 * pad + 1));
    *fl = static_cast<int>(r.z & 1u);
  } else {
    *ox = pad;
    *oy = pad;
    *fl = 0;
  }
}

}  // namespace pmd
