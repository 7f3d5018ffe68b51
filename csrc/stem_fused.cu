// The network's first layer in one persistent kernel: batch gather + zero-pad random crop + horizontal flip +
// ToTensor/Normalize + the 3x3, 3->64 stem convolution + its BatchNorm statistics (ref:data.py:11-19, 41-53 and
// ref:model/resnet.py:79-80, 96).
//
// The two-kernel form (augment_im2col_kernel writing a [pixels][64] buffer, then a 64-wide 1x1 GEMM reading it back)
// costs 33 + 47 us at 512 images: the first is instruction-bound (81 fp32 divisions per pixel), the second spends its
// time on per-CTA set-up for two MMAs of work.  Here
//   * builder warps (three groups of 128 threads, one output pixel per thread) compute the 27-value patch of their
//     pixel -- uint8 -> bf16 through a 256-entry table built with the reference's exact arithmetic -- and store it
//     straight into the 128B-swizzled shared-memory A tile the tensor core reads (and, when a weight gradient will
//     follow, into columns 0..31 of x0: the stem wgrad's operand).  Crop/flip come from Philox (resident dataset) or
//     from the staged per-sample parameters, once per warp: a 128-pixel tile lies inside one image.  The stored rows
//     a tile reads are first copied into shared memory with coalesced loads (one memory round trip per tile, not one
//     per tap), and a CTA walks a contiguous range of tiles so an image is fetched from DRAM once;
//   * one thread issues two K = 16 MMAs per tile (27 live columns of 32) against the weight tile that stays resident;
//   * two epilogue warpgroups alternate over two TMEM accumulators: tcgen05.ld -> bf16 -> swizzled staging -> TMA
//     store, statistics accumulated in registers over all tiles of the CTA (128 red.global per CTA).
// Per step the layer then moves y (67 MB) + x0 (34 MB) instead of 235 MB.
#include <cstdlib>
#include <cstring>

#include "aug.cuh"
#include "conv_tc.cuh"
#include "ptx.cuh"
#include "../include/pmdist.h"

namespace pmd {

using namespace ptx;

namespace {

// GROUPS = builder groups of 128 threads; warp 0: MMA + TMEM + weights; then the builders; then 8 epilogue warps.
// Builder group b owns A slot b.
__host__ __device__ constexpr int sf_slots(int groups) { return groups; }
__host__ __device__ constexpr int sf_epi_warp0(int groups) { return 1 + 4 * groups; }
__host__ __device__ constexpr int sf_threads(int groups) { return 32 * (sf_epi_warp0(groups) + 8); }
constexpr int kSfTileBytes = 16384;                     // 128 pixels x 128 B
constexpr int kSfWBytes = 8192;                         // 64 output channels x 128 B
constexpr int kSfStageBytes = 8192 + 64;                // a tile's augmented region as bf16 (one per builder group) + read slack

template <int GROUPS>
struct SfBars {
  uint64_t a_full[GROUPS], a_empty[GROUPS];
  uint64_t w_full;
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
  float stat_part[2][128];
  alignas(16) float piv[64];    // statistics pivot per channel (0 when none)
  unsigned short lut[256];      // bf16 bits of Normalize(ToTensor(v)) for every byte value
};
template <int GROUPS>
constexpr int sf_smem() {
  return kSfWBytes + sf_slots(GROUPS) * kSfTileBytes + 2 * kSfTileBytes + GROUPS * kSfStageBytes +
         static_cast<int>(sizeof(SfBars<GROUPS>)) + 1024;
}

__device__ __forceinline__ uint8_t* sf_align_1024(uint8_t* raw) {
  const uint32_t s = smem_u32(raw);
  return raw + ((1024u - (s & 1023u)) & 1023u);
}

template <bool GATHER, int GROUPS>
__global__ void __launch_bounds__(sf_threads(GROUPS), 1) stem_fused_kernel(const __grid_constant__ StemFusedParams p,
                                                                          int total_tiles) {
  constexpr int kSfBuilderGroups = GROUPS, kSfSlots = sf_slots(GROUPS), kSfEpiWarp0 = sf_epi_warp0(GROUPS);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = sf_align_1024(smem_raw);
  uint8_t* smem_w = smem;
  uint8_t* smem_a = smem + kSfWBytes;
  uint8_t* smem_out = smem_a + kSfSlots * kSfTileBytes;
  uint8_t* smem_stage = smem_out + 2 * kSfTileBytes;
  SfBars<GROUPS>* bars = reinterpret_cast<SfBars<GROUPS>*>(smem_stage + kSfBuilderGroups * kSfStageBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kSfSlots; ++s) {
      mbar_init(&bars->a_full[s], 128);
      mbar_init(&bars->a_empty[s], 1);
    }
    mbar_init(&bars->w_full, p.w_raw != nullptr ? 256 : 1);
    mbar_init(&bars->tmem_full[0], 1);
    mbar_init(&bars->tmem_full[1], 1);
    mbar_init(&bars->tmem_empty[0], 128);
    mbar_init(&bars->tmem_empty[1], 128);
    fence_mbar_init();
    prefetch_tmap(&p.w);
    prefetch_tmap(&p.out);
    if (p.x0_tma) prefetch_tmap(&p.x0m);
  }
  if (threadIdx.x >= 32 && threadIdx.x < 288) {
    const int v = threadIdx.x - 32;
    const __nv_bfloat16 h = __float2bfloat16_rn((v / 255.0f - 0.5f) / 0.5f);   // the arithmetic of augment_im2col_kernel
    bars->lut[v] = *reinterpret_cast<const unsigned short*>(&h);
  }
  // columns 32..63 of the A slots are never written by the builders and never read by the MMAs; zero them once so the
  // slot can also be stored to x0 as full 128-byte rows
  for (int i = threadIdx.x; i < kSfSlots * kSfTileBytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem_a)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_proxy_async_smem();
  if (warp == 0) {
    __syncwarp();
    tmem_alloc<128>(&bars->tmem_base);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  // a CTA walks a CONTIGUOUS range of tiles: consecutive tiles share an image, so after its first touch the image's
  // bytes come from L1/L2 instead of DRAM
  const int t_base = total_tiles / static_cast<int>(gridDim.x), t_extra = total_tiles % static_cast<int>(gridDim.x);
  const int b_id = static_cast<int>(blockIdx.x);
  const int t_begin = b_id * t_base + (b_id < t_extra ? b_id : t_extra);
  const uint32_t t_count = static_cast<uint32_t>(t_base + (b_id < t_extra ? 1 : 0));
  pdl_wait();

  if (warp == 0) {
    // ------------------------------------------------------------------ weights + MMA issuer
    if (lane == 0) {
      if (p.w_raw == nullptr) {
        mbar_expect_tx(&bars->w_full, kSfWBytes);
        tma_load_2d(&p.w, &bars->w_full, smem_w, 0, 0);
      }
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, false, false);
      const uint32_t w_addr = smem_u32(smem_w);
      mbar_wait(&bars->w_full, 0);
      tc_fence_after();
      uint32_t j = 0;
      for (; j < t_count; ++j) {
        const uint32_t acc = j & 1, use = j >> 1;
        mbar_wait(&bars->tmem_empty[acc], (use & 1) ^ 1u);
        const uint32_t s = j % kSfSlots;
        mbar_wait(&bars->a_full[s], (j / kSfSlots) & 1);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem_a + s * kSfTileBytes);
#pragma unroll
        for (int k = 0; k < 2; ++k)      // columns 0..31 of the patch rows: 27 live values
          umma_bf16(tmem_base + acc * 64, make_smem_desc_sw128(a_addr + k * 32, 16, 1024),
                    make_smem_desc_sw128(w_addr + k * 32, 16, 1024), idesc, k != 0 ? 1u : 0u);
        umma_commit(&bars->a_empty[s]);
        umma_commit(&bars->tmem_full[acc]);
      }
    }
  } else if (warp < kSfEpiWarp0) {
    // ------------------------------------------------------------------ patch builders
    const int bg = (warp - 1) >> 2;
    const int bt = static_cast<int>(threadIdx.x) - 32 - bg * 128;      // pixel of the tile
    const int HW = p.H * p.W;
    const GatherSrc g{p.hdr, p.idx, p.ds, p.labels_out, p.valid_out};
    unsigned short* region = reinterpret_cast<unsigned short*>(smem_stage + bg * kSfStageBytes);
    const uint32_t region_s = smem_u32(region);
    uint8_t* slot = smem_a + bg * kSfTileBytes;
    const int bar_a = 4 + 2 * bg, bar_b = 5 + 2 * bg;
    uint32_t it = 0;
    bool x0_pending = false;
    for (uint32_t j = bg; j < t_count; j += kSfBuilderGroups, ++it) {
      const int t = t_begin + static_cast<int>(j);
      const int n = (t * 128) / HW;                  // a tile never straddles two images (HW % 128 == 0)
      const int rem0 = t * 128 - n * HW;
      const int y_first = rem0 / p.W, x_first = rem0 - y_first * p.W, y_last = (rem0 + 127) / p.W;
      int src = n, ox, oy, fl;
      if (GATHER) {
        if (lane == 0) gather_params(g, n, p.pad, &src, &ox, &oy, &fl);
        src = __shfl_sync(0xffffffffu, src, 0);
        ox = __shfl_sync(0xffffffffu, ox, 0);
        oy = __shfl_sync(0xffffffffu, oy, 0);
        fl = __shfl_sync(0xffffffffu, fl, 0);
        if (rem0 == 0 && bt == 0) {                  // one thread per sample hands label and validity to the loss kernel
          p.labels_out[n] = reinterpret_cast<const long long*>(p.ds[1])[src];
          p.valid_out[n] = p.idx[n] < 0 ? 0 : 1;
        }
      } else {
        ox = p.aug[n * 3];
        oy = p.aug[n * 3 + 1];
        fl = p.aug[n * 3 + 2];
      }
      const unsigned char* img =
          (GATHER ? reinterpret_cast<const unsigned char*>(p.ds[0]) : p.u8) + static_cast<long long>(src) * HW * 3;
      // Phase 1: the augmented image region the tile's patches cover -- rows y_first-1 .. y_last+1, columns -1 .. W (or
      // x_first-1 .. x_first+128 when the tile lies inside one row) -- is resolved ONCE per pixel (conv padding 0, crop
      // padding -1, flip, uint8 -> bf16 through the table) into shared memory: 3 bf16 per pixel, pixel-major.
      const bool one_row = y_last == y_first;
      const int cx_lo = one_row ? x_first - 1 : -1;
      const int PW = one_row ? 130 : p.W + 2;
      const int npix = (y_last - y_first + 3) * PW;
      {
        const int d_r = 128 / PW, d_x = 128 - d_r * PW;
        int ry = bt / PW, rx = bt - ry * PW;
        for (int i = bt; i < npix; i += 128) {
          const int ay = y_first - 1 + ry, ax = cx_lo + rx;      // position in the augmented image
          unsigned short t0 = 0, t1 = 0, t2 = 0;                 // outside the image: the convolution's zero padding
          if (ay >= 0 && ay < p.H && ax >= 0 && ax < p.W) {
            const int xc = fl ? (p.W - 1 - ax) : ax;
            const int sx = xc + ox - p.pad, sy = ay + oy - p.pad;   // position in the stored image
            t0 = t1 = t2 = 0xBF80;                   // inside the image, outside the crop: torchvision's zero pad, -1 after Normalize
            if (sy >= 0 && sy < p.H && sx >= 0 && sx < p.W) {
              const unsigned char* px = img + (sy * p.W + sx) * 3;
              t0 = bars->lut[__ldg(px)];
              t1 = bars->lut[__ldg(px + 1)];
              t2 = bars->lut[__ldg(px + 2)];
            }
          }
          region[i * 3] = t0;
          region[i * 3 + 1] = t1;
          region[i * 3 + 2] = t2;
          rx += d_x;
          ry += d_r;
          if (rx >= PW) {
            rx -= PW;
            ++ry;
          }
        }
      }
      if (bt == 0 && x0_pending) tma_store_wait_read0();          // the slot's previous x0 store has read it
      asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
      // Phase 2: a pixel's patch = three runs of 9 consecutive bf16 (taps s = 0..2 x 3 channels) of the region; read as
      // aligned words, realigned with funnel shifts, concatenated into the 27 (+5 zero) columns of the A row.
      const int rem = rem0 + bt;
      const int y = rem / p.W, x = rem - y * p.W;
      uint32_t q[3][5];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int e0 = 3 * ((y - y_first + r) * PW + (x - 1 - cx_lo));
        const uint32_t sh = (e0 & 1) * 16;
        const uint32_t addr = region_s + static_cast<uint32_t>(e0 >> 1) * 4;
        uint32_t a[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a[k]) : "r"(addr + k * 4));
#pragma unroll
        for (int k = 0; k < 5; ++k) q[r][k] = __funnelshift_r(a[k], a[k + 1], sh);
      }
      uint32_t w[16];
#pragma unroll
      for (int k = 0; k < 4; ++k) w[k] = q[0][k];
      w[4] = (q[0][4] & 0xffffu) | (q[1][0] << 16);
#pragma unroll
      for (int k = 0; k < 4; ++k) w[5 + k] = __funnelshift_r(q[1][k], q[1][k + 1], 16);
#pragma unroll
      for (int k = 0; k < 4; ++k) w[9 + k] = q[2][k];
      w[13] = q[2][4] & 0xffffu;
      w[14] = w[15] = 0u;
      if (p.x0 != nullptr && !p.x0_tma) {
        uint4* o = reinterpret_cast<uint4*>(p.x0 + static_cast<long long>(t * 128 + bt) * 64);
#pragma unroll
        for (int c = 0; c < 4; ++c) o[c] = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
      }
      mbar_wait(&bars->a_empty[bg], (it & 1) ^ 1u);               // the MMAs of this group's previous tile have read the slot
      uint8_t* rowp = slot + bt * 128;
#pragma unroll
      for (int c = 0; c < 4; ++c)
        *reinterpret_cast<uint4*>(rowp + ((c ^ (bt & 7)) << 4)) = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
      fence_proxy_async_smem();                      // generic-proxy stores -> visible to the tensor core and to TMA
      mbar_arrive(&bars->a_full[bg]);
      asm volatile("bar.sync %0, 128;" ::"r"(bar_b) : "memory");   // slot complete; everybody is done reading the region
      if (p.x0 != nullptr && p.x0_tma && bt == 0) {
        // the im2col rows for the weight gradient: the same shared-memory tile, stored by TMA (the map is 32 columns wide,
        // or 64 with the always-zero upper half of the slot)
        tma_store_2d(&p.x0m, slot, 0, t * 128);
        tma_store_commit();
        x0_pending = true;
      }
    }
    if (bt == 0 && x0_pending) tma_store_wait_read0();
  } else {
    // ------------------------------------------------------------------ two epilogue warpgroups
    const int g = (warp - kSfEpiWarp0) >> 2;
    const int q4 = warp & 3;                         // TMEM lane quarter this warp may read
    const int row = q4 * 32 + lane;                  // accumulator row = pixel of the tile
    const int et = static_cast<int>(threadIdx.x) - 32 * kSfEpiWarp0 - g * 128;
    uint8_t* stage_out = smem_out + g * kSfTileBytes;
    const int bar_id = 1 + g;
    if (p.w_raw != nullptr) {
      // the [64][rsc] weight as the optimizer left it -> K-major SW128 tile, columns rsc..31 zero (no re-pad kernel at the
      // end of every step): 256 threads x 8 elements, one 16-byte chunk each
      const int wt = et + g * 128;                   // 0..255
      const int k = wt >> 2, c0 = (wt & 3) * 8;
      uint32_t w4[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = c0 + 2 * e;
        const unsigned short lo = c < p.w_rsc ? *reinterpret_cast<const unsigned short*>(p.w_raw + k * p.w_rsc + c) : 0;
        const unsigned short hi = c + 1 < p.w_rsc ? *reinterpret_cast<const unsigned short*>(p.w_raw + k * p.w_rsc + c + 1) : 0;
        w4[e] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
      }
      *reinterpret_cast<uint4*>(smem_w + k * 128 + (((c0 >> 3) ^ (k & 7)) << 4)) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
      fence_proxy_async_smem();
      mbar_arrive(&bars->w_full);
    }
    float s1[8], s2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s1[i] = s2[i] = 0.f;
    if (g == 0 && et < 64) bars->piv[et] = (p.stats != nullptr && p.stat_pivot != nullptr) ? p.stat_pivot[et] : 0.f;
    asm volatile("bar.sync 3, 256;" ::: "memory");      // both epilogue groups: the pivots are in shared memory
    uint32_t u = 0;
    bool store_pending = false;
    for (uint32_t j = g; j < t_count; j += 2, ++u) {
      const int t = t_begin + static_cast<int>(j);
      mbar_wait(&bars->tmem_full[g], u & 1);
      tc_fence_after();
      if (et == 0 && store_pending) tma_store_wait_read0();     // the previous store has read the staging buffer
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q4 * 32) << 16) + g * 64 + cc * 32, v);
        tmem_ld_wait();
        uint32_t packed[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) packed[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
        uint8_t* rowp = stage_out + row * 128;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int chunk = (cc * 4 + jj) ^ (row & 7);
          *reinterpret_cast<uint4*>(rowp + chunk * 16) =
              make_uint4(packed[jj * 4], packed[jj * 4 + 1], packed[jj * 4 + 2], packed[jj * 4 + 3]);
        }
      }
      tc_fence_before();
      mbar_arrive(&bars->tmem_empty[g]);             // this thread's TMEM reads are complete
      fence_proxy_async_smem();
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (et == 0) {
        tma_store_2d(&p.out, stage_out, 0, t * 128);
        tma_store_commit();
      }
      store_pending = true;
      if (p.stats != nullptr) {
        // thread -> 16-byte channel group (et & 7) of rows (et >> 3) + 16 i: one LDS.128 covers 8 channels
        const int cg = et & 7;
        const float4 pa = *reinterpret_cast<const float4*>(&bars->piv[cg * 8]);
        const float4 pb = *reinterpret_cast<const float4*>(&bars->piv[cg * 8 + 4]);
        const float piv[8] = {pa.x, pa.y, pa.z, pa.w, pb.x, pb.y, pb.z, pb.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = (et >> 3) + 16 * i;
          const uint4 h4 = *reinterpret_cast<const uint4*>(stage_out + rr * 128 + ((cg ^ (rr & 7)) << 4));
          const uint32_t hw[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 f = unpack_bf16x2(hw[e]);
            f.x -= piv[2 * e];
            f.y -= piv[2 * e + 1];
            s1[2 * e] += f.x;
            s2[2 * e] += f.x * f.x;
            s1[2 * e + 1] += f.y;
            s2[2 * e + 1] += f.y * f.y;
          }
        }
      }
    }
    if (et == 0 && store_pending) tma_store_wait_read0();
    if (p.stats != nullptr) {
      // lanes with equal (lane & 7) hold the same channel group: fold 4 lanes, then the 4 warps via shared memory
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 8);
        s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 16);
        s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 8);
        s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 16);
      }
      float* part = bars->stat_part[g];
      part[et] = 0.f;
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (lane < 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          atomicAdd(&part[lane * 8 + i], s1[i]);
          atomicAdd(&part[64 + lane * 8 + i], s2[i]);
        }
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      red_add_f32(p.stats + et, part[et]);           // [sum 64 | sum of squares 64]
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<128>(tmem_base);
}

}  // namespace

bool stem_fused_enabled() {
  static const bool on = getenv("PMD_STEM_FUSED") == nullptr || atoi(getenv("PMD_STEM_FUSED")) != 0;
  return on;
}

int build_stem_fused_params(int N, int H, int W, int pad, const void* w_pad, void* x0, void* y, StemFusedParams* out) {
  PMD_REQUIRE(stem_fused_supported(N, H, W), "the fused stem kernel needs H*W % 128 == 0, W % 4 == 0 and W <= 336");
  StemFusedParams& p = *out;
  memset(&p, 0, sizeof(p));
  p.N = N;
  p.H = H;
  p.W = W;
  p.pad = pad;
  {
    const uint64_t dims[2] = {64ull, 64ull};
    const uint64_t strides[1] = {128ull};
    const uint32_t box[2] = {64u, 64u};
    if (w_pad != nullptr && encode_tiled_bf16(&p.w, w_pad, 2, dims, strides, box)) return -1;
  }
  const uint64_t dims[2] = {64ull, static_cast<uint64_t>(N) * H * W};
  const uint64_t strides[1] = {128ull};
  const uint32_t box[2] = {64u, 128u};
  if (x0 != nullptr && (getenv("PMD_STEM_X0_TMA") == nullptr || atoi(getenv("PMD_STEM_X0_TMA")) != 0)) {
    // 32 columns wide: the box's upper half falls outside the tensor and is not written.  Should the driver refuse a box
    // wider than the tensor, the full-width map stores the slot's zero upper half as well (x0's is zero anyway).
    const uint64_t dims32[2] = {32ull, dims[1]};
    if (encode_tiled_bf16(&p.x0m, x0, 2, dims32, strides, box) != 0 && encode_tiled_bf16(&p.x0m, x0, 2, dims, strides, box) != 0)
      return -1;
    p.x0_tma = 1;
  }
  return encode_tiled_bf16(&p.out, y, 2, dims, strides, box);
}

template <int GROUPS>
static int launch_stem_fused_t(const StemFusedParams& p, cudaStream_t stream) {
  static PerDevice<int> sms_of;      // also marks "shared-memory attribute set on this device"
  int* sms = sms_of.get();
  if (sms == nullptr) return -1;
  if (*sms == 0) {
    PMD_CHECK_CUDA(cudaFuncSetAttribute(stem_fused_kernel<true, GROUPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, sf_smem<GROUPS>()));
    PMD_CHECK_CUDA(cudaFuncSetAttribute(stem_fused_kernel<false, GROUPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, sf_smem<GROUPS>()));
    int dev = 0;
    PMD_CHECK_CUDA(cudaGetDevice(&dev));
    PMD_CHECK_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const int total = static_cast<int>(static_cast<long long>(p.N) * p.H * p.W / 128);
  const int grid = total < *sms ? total : *sms;
  if (p.ds != nullptr)
    PMD_CHECK_CUDA(launch_pdl(stem_fused_kernel<true, GROUPS>, dim3(grid), dim3(sf_threads(GROUPS)), sf_smem<GROUPS>(), stream, p, total));
  else
    PMD_CHECK_CUDA(launch_pdl(stem_fused_kernel<false, GROUPS>, dim3(grid), dim3(sf_threads(GROUPS)), sf_smem<GROUPS>(), stream, p, total));
  return 0;
}

int launch_stem_fused(const StemFusedParams& p, cudaStream_t stream) {
  PMD_REQUIRE((p.ds != nullptr && p.hdr != nullptr && p.idx != nullptr && p.labels_out != nullptr && p.valid_out != nullptr) ||
                  (p.u8 != nullptr && p.aug != nullptr),
              "a batch source is required: resident dataset (descriptor, header, indices, label/valid outputs) or staged images");
  static const int groups = getenv("PMD_STEM_GROUPS") != nullptr ? atoi(getenv("PMD_STEM_GROUPS")) : 4;
  int rc;
  if (groups == 2) rc = launch_stem_fused_t<2>(p, stream);
  else if (groups == 3) rc = launch_stem_fused_t<3>(p, stream);
  else rc = launch_stem_fused_t<4>(p, stream);
  if (rc != 0) return -1;
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

}  // namespace pmd

using namespace pmd;

extern "C" PMD_API int pmd_stem_fused(const PmdStemFused* a, void* stream) {
  PMD_REQUIRE(a != nullptr && a->w_pad != nullptr && a->y != nullptr, "weights and output are required");
  PMD_REQUIRE(a->w_rsc >= 0 && a->w_rsc <= 32, "w_rsc must be 0 (padded [64][64] weight) or the row length of the raw weight (<= 32)");
  StemFusedParams p;
  if (build_stem_fused_params(a->N, a->H, a->W, a->pad, a->w_pad, a->x0, a->y, &p)) return -1;
  p.ds = a->d_dataset_desc;
  p.hdr = a->d_hdr;
  p.idx = a->d_idx;
  p.u8 = a->d_u8;
  p.aug = a->d_aug;
  p.labels_out = a->d_labels_out;
  p.valid_out = a->d_valid_out;
  p.x0 = static_cast<__nv_bfloat16*>(a->x0);
  p.stats = a->stats;
  p.stat_pivot = a->stat_pivot;
  if (a->w_rsc > 0) {
    p.w_raw = static_cast<const __nv_bfloat16*>(a->w_pad);
    p.w_rsc = a->w_rsc;
  }
  return launch_stem_fused(p, static_cast<cudaStream_t>(stream));
}
