// Implicit-GEMM convolution for sm_100a: tcgen05.mma with TMEM accumulators, operands staged into
// 128B-swizzled shared memory by TMA, warp-specialised (1 TMA thread, 1 MMA thread, 4 epilogue warps).
//
// Replaces cuDNN's fprop / dgrad / wgrad under nn.Conv2d (ref:model/resnet.py:20-31, 48-61, 79;
// ref:main.py:45 cudnn.benchmark).
//
// fprop / dgrad  -- "gather GEMM":  Out[pixel, n] = sum over k-blocks  A_kb[pixel, 64] * B_kb[n, 64]
//   * a CTA owns a tile of 128 output pixels (a w_box x h_box x n_box brick of the NHWC grid) and
//     BN in {64,128} output channels; the accumulator is a 128-lane x BN-column fp32 TMEM tile;
//   * a k-block is (filter tap, 64-channel chunk).  Its A operand is the same brick shifted by the
//     tap offset, fetched with ONE 4-D TMA box -- zero padding is the TMA out-of-bounds fill, stride-2
//     convolutions read one of four parity views of the input (separate tensor maps), and stride-2
//     dgrad runs one output-parity class per blockIdx.z.  No im2col buffer ever exists;
//   * fprop epilogue: bf16 round, per-channel sum / sum-of-squares for BatchNorm and a
//     TMA store; dgrad epilogue can fold in a second convolution's k-blocks (projection shortcut) and an
//     addend tensor (identity shortcut), so the branch sum of a residual block needs no extra pass.
// wgrad -- dW[(tap,c), k] = sum over pixels X_tap[pixel, c] * dY[pixel, k]
//   * both operands are "MN-major" (the reduction dim, pixels, is the slow dim of NHWC), which
//     tcgen05 consumes directly from the TMA-written tiles: no transposes;
//   * M = 128 = two (tap, 64-channel) boxes of X, N = up to 256 output channels, reduction split over
//     gridDim.z CTAs; partial tiles go straight into the fp32 gradient slab with red.global.add.f32.
#include <cstdlib>

#include "conv_tc.cuh"
#include "ptx.cuh"

namespace pmd {

using namespace ptx;

namespace {

constexpr int kThreads = 192;      // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2..5: epilogue
constexpr int kABytes = 16384;     // 128 rows x 128 B (fprop/dgrad A; wgrad: two 64-pixel boxes)

__device__ __forceinline__ void tma_load_3d(const CUtensorMap* m, uint64_t* bar, void* dst, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

template <int STAGES>
struct PipeBars {
  uint64_t full[STAGES];
  uint64_t empty[STAGES];
  uint64_t tmem_full;
  uint32_t tmem_base;
  uint32_t invalid_rows[4];   // per epilogue warp (TMEM lane quarter): rows of the tile outside the tensor (zero-filled)
};

__device__ __forceinline__ uint4 ld_stream16(const void* p) {     // 16-byte global load, L2 evict-first
  uint4 u;
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
               : "l"(p), "l"(pol)
               : "memory");
  return u;
}

__device__ __forceinline__ uint8_t* align_smem_1024(uint8_t* raw) {
  const uint32_t s = smem_u32(raw);
  return raw + ((1024u - (s & 1023u)) & 1023u);
}

// ============================================================================ fprop / dgrad
// B_MN: the B operand is consumed MN-major -- dgrad reading the SAME KRSC weights as fprop ([k][tap][c]: for
// one tap a [k rows][c contiguous] matrix, i.e. the output-channel dim of dgrad is the contiguous one), so no
// transposed RSCK copy of the weights has to be maintained.  The tile is BN/64 boxes of [64 k][64 c].
__device__ __forceinline__ void trace_mark(const ConvGemmParams& p, int slot) {
  if (p.trace == nullptr) return;
  const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  if (cta < p.trace_cap) p.trace[static_cast<long long>(cta) * kTraceSlots + slot] = clock64();
}
__device__ __forceinline__ void trace_begin(const ConvGemmParams& p, int tiles) {
  if (p.trace == nullptr) return;
  const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  if (cta >= p.trace_cap) return;
  long long* t = p.trace + static_cast<long long>(cta) * kTraceSlots;
  unsigned long long ns;
  unsigned sm;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
  asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
  t[kTrEntry] = clock64();
  t[kTrGlobalNs] = static_cast<long long>(ns);
  t[kTrSm] = sm;
  t[kTrTiles] = tiles;
}

template <int BN>
struct RedExtra {          // extra shared state of the RED (dgrad + BatchNorm-backward reduction) instantiations
  uint64_t y_full;
  uint64_t pad_;           // keeps sc/sh 16-byte aligned (read as float4)
  float sc[BN], sh[BN];
};

// TWO: CTA-pair form (tcgen05 cta_group::2).  Two CTAs of a cluster -- neighbouring pixel bricks, same channel tile --
// run ONE 256 x BN MMA per k-step: each stages its own brick (A) and only HALF of the weight slice (BN/2 rows of B),
// the leader (cluster rank 0) issues the MMAs for both, each CTA's tensor core reads the missing half of B from its
// partner's shared memory.  Per-CTA operand traffic out of shared memory drops from 4 KB + BN*32 B to 4 KB + BN*16 B
// per k-step (BN = 128: 128 -> 96 B/clk, under the SM's 128 B/clk) and the weight slice crosses L2 -> SM once per
// pair instead of once per CTA.  Epilogues stay per CTA (each holds its 128 accumulator rows in its own TMEM).
// RED: 0 plain; 1 = the dgrad also does the BatchNorm-backward reduction of the layer below, ReLU mask recomputed from
// that layer's y (inner BatchNorms); 2 = the same for a residual block's OUTPUT BatchNorm(s): the mask comes from the
// ReLU bit array the forward apply recorded, one or two branches (main path + projection shortcut) share it, and the
// dgrad may be strided (one parity view of y per class).
template <int BN, int STAGES, bool B_MN = false, int RED = 0, bool TWO = false>
__global__ void __launch_bounds__(kThreads) conv_gemm_kernel(const __grid_constant__ ConvGemmParams p) {
  constexpr int kBBytes = (TWO ? BN / 2 : BN) * 128;      // this CTA's share of the weight slice
  constexpr int kStageBytes = kABytes + kBBytes;
  static_assert(!TWO || BN >= 128, "CTA pairs need at least 64 weight rows per CTA");
  const bool leader = !TWO || cluster_ctarank() == 0;
  const int half = TWO ? static_cast<int>(cluster_ctarank()) : 0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  PipeBars<STAGES>* bars = reinterpret_cast<PipeBars<STAGES>*>(smem + STAGES * kStageBytes);
  RedExtra<BN>* rx = reinterpret_cast<RedExtra<BN>*>(reinterpret_cast<uint8_t*>(bars) + ((sizeof(PipeBars<STAGES>) + 15) & ~15));
  static_assert(RED == 0 || STAGES * kStageBytes >= (RED + 1) * (BN / 64) * kABytes,
                "the y tile(s) do not fit beside the output staging");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cls = blockIdx.z, ntile = blockIdx.y;
  int t = blockIdx.x;
  const int tw = t % p.tiles_w;
  t /= p.tiles_w;
  const int th = t % p.tiles_h;
  const int tn = t / p.tiles_h;
  const int w0 = tw * p.w_box, h0 = th * p.h_box, n0 = tn * p.n_box;
  const int kb0 = p.kb_begin[cls], nkb = p.kb_count[cls];

  if (threadIdx.x == 0) {
    trace_begin(p, 1);
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
    }
    mbar_init(&bars->tmem_full, 1);
    if (RED) mbar_init(&rx->y_full, 1);
    fence_mbar_init();
    prefetch_tmap(&p.a[0]);
    prefetch_tmap(&p.b[0]);
    prefetch_tmap(&p.out[cls]);
  }
  if (warp == 1) {
    if (TWO) tmem_alloc_pair<BN>(&bars->tmem_base); else tmem_alloc<BN>(&bars->tmem_base);
  }
  tc_fence_before();
  if (TWO) cluster_sync_all(); else __syncthreads();     // pair: the partner's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  if (threadIdx.x == 0) trace_mark(p, kTrSetup);
  pdl_wait();        // everything above is private to this CTA; operands below come from the previous kernel
  if (RED == 1) {
    for (int c = threadIdx.x; c < BN; c += kThreads) {
      rx->sc[c] = p.red_scale[ntile * BN + c];
      rx->sh[c] = p.red_shift[ntile * BN + c];
    }
    __syncthreads();
  }

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&bars->empty[s], ph ^ 1u);
        const KBlock kb = p.kb[kb0 + i];
        uint8_t* sa = smem + s * kStageBytes;
        uint8_t* sb = sa + kABytes;
        if (TWO) {
          // the leader's barrier counts the bytes of BOTH CTAs' loads; each CTA fills its own shared memory
          if (leader) mbar_expect_tx(&bars->full[s], 2 * kStageBytes);
          tma_load_4d_pair(&p.a[kb.a_map], &bars->full[s], sa, kb.a_ch, w0 + kb.dw, h0 + kb.dh, n0);
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 128; ++j)
              tma_load_3d_pair(&p.b[kb.b_map], &bars->full[s], sb + j * 8192, ntile * BN + (half * (BN / 128) + j) * 64,
                               kb.tap, kb.b_c0);
          } else if (p.b_tap_last) {
            tma_load_3d_pair(&p.b[kb.b_map], &bars->full[s], sb, kb.b_c0, ntile * BN + half * (BN / 2), kb.tap);
          } else {
            tma_load_3d_pair(&p.b[kb.b_map], &bars->full[s], sb, kb.b_c0, kb.tap, ntile * BN + half * (BN / 2));
          }
          if (i == 0) trace_mark(p, kTrFirstLoad);
          continue;
        }
        mbar_expect_tx(&bars->full[s], kStageBytes);
        if ((p.l2_hint & (B_MN ? 4 : 2)) != 0)
          tma_load_4d_stream(&p.a[kb.a_map], &bars->full[s], sa, kb.a_ch, w0 + kb.dw, h0 + kb.dh, n0);
        else
          tma_load_4d(&p.a[kb.a_map], &bars->full[s], sa, kb.a_ch, w0 + kb.dw, h0 + kb.dh, n0);
        if (B_MN) {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)   // KRSC map (c, tap, k): 64 output channels x 64 reduction rows per box
            tma_load_3d(&p.b[kb.b_map], &bars->full[s], sb + j * 8192, ntile * BN + j * 64, kb.tap, kb.b_c0);
        } else if (p.b_tap_last) {
          tma_load_3d(&p.b[kb.b_map], &bars->full[s], sb, kb.b_c0, ntile * BN, kb.tap);
        } else {
          tma_load_3d(&p.b[kb.b_map], &bars->full[s], sb, kb.b_c0, kb.tap, ntile * BN);
        }
        if (i == 0) trace_mark(p, kTrFirstLoad);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = make_idesc_bf16(TWO ? 256 : 128, BN, false, B_MN);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&bars->full[s], ph);
        tc_fence_after();
        if (i == 0) trace_mark(p, kTrFirstFull);
        const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
        const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
        for (int k = 0; k < 4; ++k) {   // 64-deep reduction chunk = 4 x (K = 16)
          const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_addr + k * 2048, 8192, 1024)
                                      : make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
          const uint64_t adesc = make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
          if (TWO) umma_bf16_pair(tmem_base, adesc, bdesc, idesc, (i | k) != 0 ? 1u : 0u);
          else umma_bf16(tmem_base, adesc, bdesc, idesc, (i | k) != 0 ? 1u : 0u);
        }
        // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
        if (TWO) umma_commit_pair(&bars->empty[s]); else umma_commit(&bars->empty[s]);
      }
      if (nkb > 0) {
        if (TWO) umma_commit_pair(&bars->tmem_full); else umma_commit(&bars->tmem_full);
      }
      trace_mark(p, kTrMmaIssued);
    }
  } else {
    // ------------------------------------------------------------------ epilogue (128 threads)
    const int q = warp & 3;              // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;       // accumulator row == pixel inside the brick
    const int wi = row % p.w_box;
    const int hi = (row / p.w_box) % p.h_box;
    const int ni = row / (p.w_box * p.h_box);
    const bool valid = (w0 + wi < p.GW[cls]) && (h0 + hi < p.GH[cls]) && (n0 + ni < p.N);
    const __nv_bfloat16* add_row = nullptr;
    const unsigned char* add_bits = nullptr;   // optional ReLU bits of the addend (see ConvGemmParams::addend_bits)
    if (p.addend != nullptr && valid) {
      const long long pix = (static_cast<long long>(n0 + ni) * p.GH[cls] + (h0 + hi)) * p.GW[cls] + (w0 + wi);
      add_row = p.addend + pix * p.out_channels + ntile * BN;
      if (p.addend_bits != nullptr) add_bits = p.addend_bits + (pix * p.out_channels + ntile * BN) / 8;
    }
    if (add_row != nullptr) {
      // the epilogue warps idle while the MMAs run: pull this row's addend (BN * 2 bytes) towards the SM now, so the
      // loads in the epilogue hit cache instead of paying a DRAM round trip each
#pragma unroll
      for (int b = 0; b < BN * 2; b += 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(add_row) + b));
    }
    if (nkb > 0) {
      mbar_wait(&bars->tmem_full, 0);
      tc_fence_after();
    }
    if (threadIdx.x == 64) trace_mark(p, kTrAccReady);
    uint8_t* stage_out = smem;           // pipeline buffers are idle now: reuse them as the store staging
    uint8_t* stage_y = smem + (BN / 64) * kABytes;
    uint8_t* stage_y2 = stage_y + (BN / 64) * kABytes;   // RED == 2 with two branches: the shortcut conv's y tile
    const int red_nb = RED == 2 ? p.red_nb : 1;
    if (RED) {
      if (threadIdx.x == 64) {           // every MMA has completed (tmem_full): the ring can take the y tile(s)
        mbar_expect_tx(&rx->y_full, red_nb * (BN / 64) * kABytes);
#pragma unroll
        for (int hf = 0; hf < BN / 64; ++hf)
          tma_load_4d(&p.ymap[RED == 2 ? cls : 0], &rx->y_full, stage_y + hf * kABytes, ntile * BN + hf * 64, w0, h0, n0);
        if (RED == 2 && red_nb == 2) {
#pragma unroll
          for (int hf = 0; hf < BN / 64; ++hf)
            tma_load_4d(&p.ymap2[cls], &rx->y_full, stage_y2 + hf * kABytes, ntile * BN + hf * 64, w0, h0, n0);
        }
      }
      mbar_wait(&rx->y_full, 0);
    }
    const unsigned char* bits_row = nullptr;   // RED == 2: this row's bytes of the ReLU bit array (full-resolution pixel)
    if (RED == 2 && valid) {
      const int st = p.red_stride, ph = cls / st, pw = cls % st;
      const long long pix = (static_cast<long long>(n0 + ni) * p.red_H + (st * (h0 + hi) + ph)) * p.red_W + (st * (w0 + wi) + pw);
      bits_row = p.red_bits + (pix * p.out_channels + ntile * BN) / 8;
    }
#pragma unroll
    for (int hf = 0; hf < BN / 64; ++hf) {
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t v[32];
        if (nkb > 0) {
          tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + hf * 64 + cc * 32, v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        uint32_t packed[16];
        if (add_row != nullptr) {
          const uint4* ap = reinterpret_cast<const uint4*>(add_row + hf * 64 + cc * 32);
          const uint32_t am = add_bits != nullptr ? *reinterpret_cast<const uint32_t*>(add_bits + (hf * 64 + cc * 32) / 8)
                                                  : 0xffffffffu;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint4 a4 = ap[jj];   // plain load: the addend may alias the output (in-place accumulate)
            const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float2 af = unpack_bf16x2(aw[e]);
              if (!((am >> (jj * 8 + e * 2)) & 1u)) af.x = 0.f;
              if (!((am >> (jj * 8 + e * 2 + 1)) & 1u)) af.y = 0.f;
              packed[jj * 4 + e] = pack_bf16x2(__uint_as_float(v[jj * 8 + e * 2]) + af.x,
                                               __uint_as_float(v[jj * 8 + e * 2 + 1]) + af.y);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            packed[j] = pack_bf16x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
        }
        if (RED == 2) {
          // ReLU mask of the block output this gradient belongs to: 32 channels = 4 bytes of the bit array
          const uint32_t m = bits_row != nullptr ? *reinterpret_cast<const uint32_t*>(bits_row + (hf * 64 + cc * 32) / 8) : 0u;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float2 gf = unpack_bf16x2(packed[j]);
            if (!((m >> (2 * j)) & 1u)) gf.x = 0.f;
            if (!((m >> (2 * j + 1)) & 1u)) gf.y = 0.f;
            packed[j] = pack_bf16x2(gf.x, gf.y);
          }
        }
        if (RED == 1) {
          // ReLU mask of the layer below, recomputed from its pre-activation y: g = (scale*y + shift > 0) ? dz : 0
          const uint8_t* yrow = stage_y + hf * kABytes + row * 128;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint4 y4 = *reinterpret_cast<const uint4*>(yrow + (((cc * 4 + jj) ^ (row & 7)) << 4));
            const uint32_t yw[4] = {y4.x, y4.y, y4.z, y4.w};
            const float4 sa = *reinterpret_cast<const float4*>(&rx->sc[hf * 64 + cc * 32 + jj * 8]);
            const float4 sb = *reinterpret_cast<const float4*>(&rx->sc[hf * 64 + cc * 32 + jj * 8 + 4]);
            const float4 ha = *reinterpret_cast<const float4*>(&rx->sh[hf * 64 + cc * 32 + jj * 8]);
            const float4 hb = *reinterpret_cast<const float4*>(&rx->sh[hf * 64 + cc * 32 + jj * 8 + 4]);
            const float scv[8] = {sa.x, sa.y, sa.z, sa.w, sb.x, sb.y, sb.z, sb.w};
            const float shv[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 yf = unpack_bf16x2(yw[e]);
              float2 gf = unpack_bf16x2(packed[jj * 4 + e]);
              if (!(fmaf(yf.x, scv[2 * e], shv[2 * e]) > 0.f)) gf.x = 0.f;
              if (!(fmaf(yf.y, scv[2 * e + 1], shv[2 * e + 1]) > 0.f)) gf.y = 0.f;
              packed[jj * 4 + e] = pack_bf16x2(gf.x, gf.y);
            }
          }
        }
        if (!valid) {
#pragma unroll
          for (int j = 0; j < 16; ++j) packed[j] = 0u;
        }
        uint8_t* rowp = stage_out + hf * kABytes + row * 128;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int chunk = (cc * 4 + jj) ^ (row & 7);     // SWIZZLE_128B, as the output tensor map expects
          *reinterpret_cast<uint4*>(rowp + chunk * 16) =
              make_uint4(packed[jj * 4], packed[jj * 4 + 1], packed[jj * 4 + 2], packed[jj * 4 + 3]);
        }
      }
    }
    if (p.stats != nullptr) {
      const uint32_t inv = __popc(__ballot_sync(0xffffffffu, !valid));
      if (lane == 0) bars->invalid_rows[q] = inv;
    }
    fence_proxy_async_smem();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int et = threadIdx.x - 64;
    if (et == 0) trace_mark(p, kTrStaged);
    if (et == 0) {
#pragma unroll
      for (int hf = 0; hf < BN / 64; ++hf)
        tma_store_4d(&p.out[cls], stage_out + hf * kABytes, ntile * BN + hf * 64, w0, h0, n0);
      tma_store_commit();
    }
    if (p.stats != nullptr) {
      // per-channel sum and sum of squares of the bf16-rounded tile, taken around the channel's pivot (see
      // ConvGemmParams::stat_pivot): thread -> (channel, half of the rows)
      const int c = et & 63, rh = et >> 6;
      const float ninv = static_cast<float>(bars->invalid_rows[2 * rh] + bars->invalid_rows[2 * rh + 1]);
#pragma unroll
      for (int hf = 0; hf < BN / 64; ++hf) {
        const float piv = p.stat_pivot != nullptr ? p.stat_pivot[ntile * BN + hf * 64 + c] : 0.f;
        float s1 = 0.f, s2 = 0.f;
        const uint8_t* base = stage_out + hf * kABytes;
#pragma unroll 8
        for (int r = rh * 64; r < rh * 64 + 64; ++r) {
          const __nv_bfloat16 hv =
              *reinterpret_cast<const __nv_bfloat16*>(base + r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2);
          const float f = __bfloat162float(hv) - piv;
          s1 += f;
          s2 += f * f;
        }
        s1 += ninv * piv;            // zero-filled rows outside the tensor were counted as (0 - pivot)
        s2 -= ninv * piv * piv;
        red_add_f32(p.stats + ntile * BN + hf * 64 + c, s1);
        red_add_f32(p.stats + p.out_channels + ntile * BN + hf * 64 + c, s2);
      }
    }
    if (RED) {
      // per channel: S_g = sum(g), S_gy = sum(g*y) over this tile; sum(g*xhat) = (S_gy - mean*S_g)*invstd is linear in them
      const int c = et & 63, rh = et >> 6;
#pragma unroll
      for (int hf = 0; hf < BN / 64; ++hf) {
        float s1 = 0.f, s2 = 0.f, s3 = 0.f;
        const uint8_t* gb = stage_out + hf * kABytes;
        const uint8_t* yb = stage_y + hf * kABytes;
        const uint8_t* yb2 = stage_y2 + hf * kABytes;
#pragma unroll 8
        for (int r = rh * 64; r < rh * 64 + 64; ++r) {
          const int off = r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2;
          const float g = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(gb + off));
          const float yv = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(yb + off));
          s1 += g;
          s2 = fmaf(g, yv, s2);
          if (RED == 2 && red_nb == 2)
            s3 = fmaf(g, __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(yb2 + off)), s3);
        }
        const int ch = ntile * BN + hf * 64 + c;
        red_add_f32(p.red_sums + ch, s1);
        red_add_f32(p.red_sums + p.out_channels + ch, (s2 - p.red_mean[ch] * s1) * p.red_invstd[ch]);
        if (RED == 2 && red_nb == 2) {
          red_add_f32(p.red_sums2 + ch, s1);
          red_add_f32(p.red_sums2 + p.out_channels + ch, (s3 - p.red_mean2[ch] * s1) * p.red_invstd2[ch]);
        }
      }
    }
    if (et == 0) tma_store_wait_read0();
    if (et == 0) trace_mark(p, kTrEpiDone);
  }
  tc_fence_before();
  if (TWO) cluster_sync_all(); else __syncthreads();     // pair: neither CTA leaves (or frees TMEM) while its partner works
  if (threadIdx.x == 0) trace_mark(p, kTrExit);
  if (warp == 1) {
    if (TWO) tmem_dealloc_pair<BN>(tmem_base); else tmem_dealloc<BN>(tmem_base);
  }
}

constexpr int kColPatchBytes = 8 * 18 * 128;      // 18432: one 8 x 18-pixel column patch (64 channels)

// ============================================================================ 3x3, C = K = 64: weights resident
// The 64-channel 3x3 layers (32x32 maps) are the worst case for the per-tap kernel: a 128x64 tile needs only 36
// MMAs (1152 tensor cycles) but pulls 216 KB of operands through L2 (the activation brick nine times plus the
// 72 KB of weights once per tile) and pays the whole CTA set-up for it; tools/conv_trace.py shows a CTA lifetime
// of ~13.7k cycles of which the tensor pipe works ~1.2k.  This kernel removes both costs:
//   * persistent: one CTA per SM walks tiles t = blockIdx.x, + gridDim.x, ...; barriers, TMEM and tensor maps are
//     set up once;
//   * the nine 64x64 weight slices (72 KB) are loaded ONCE per CTA and stay in shared memory;
//   * activations arrive as column patches: the CTA's output tile is 8 (w) x 16 (h) pixels and THREE TMA boxes of
//     8 x 18 pixels are loaded, one per horizontal tap offset s (columns w0+s-1 .. w0+s+6, rows h0-1 .. h0+16).  In
//     such a patch a pixel row is exactly one 8-row swizzle atom (1024 B), so the A operand of tap (r, s) is the
//     ordinary K-major SW128 tile that starts r atoms into patch s (every start stays 1024-byte aligned).  3 x 18 KB
//     per tile instead of 9 x 16 KB, through a ring of six patch slots that runs ahead across tile boundaries;
//   * two accumulators in TMEM and TWO epilogue warpgroups (tiles alternate between them), each with its own
//     staging buffer, so tcgen05.ld -> bf16 -> smem -> TMA store of tile j overlaps the MMAs of tiles j+1, j+2;
//   * BatchNorm statistics are accumulated in registers over all tiles of the CTA and leave as 128 red.global
//     per CTA instead of 256 per tile.
// L2->SM traffic per tile drops from 216 KB to 54 KB.  B_MN = true is dgrad straight from the KRSC weights (the
// very same shared-memory image, read MN-major, taps flipped).
constexpr int kResSlots = 6;
constexpr int kResThreads = 64 + 256;
constexpr int kResWBytes = 9 * 8192;
struct ResBars {
  uint64_t a_full[kResSlots], a_empty[kResSlots];
  uint64_t w_full;
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
  float stat_part[2][128];      // per epilogue group: [sum 64 | sum of squares 64]
};
constexpr int kResSmem = kResWBytes + kResSlots * kColPatchBytes + 2 * kABytes + sizeof(ResBars) + 1024;

// RED (dgrad only): the BatchNorm-backward reduction of the layer below in the epilogue.  The tile is staged unmasked;
// a second pass in the statistics mapping (thread = 8 channels x 8 rows) reads the matching cells of y straight from
// global memory (coalesced 16-byte loads issued before the accumulator wait), masks, accumulates sum(g) and sum(g*y) in
// registers and writes the masked cells back before the TMA store: no y tile in shared memory.
template <bool B_MN, bool RED = false>
__global__ void __launch_bounds__(kResThreads) conv3x3_resident_kernel(const __grid_constant__ ConvResidentParams p,
                                                                       int total_tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* smem_w = smem;
  uint8_t* smem_a = smem + kResWBytes;
  uint8_t* smem_out = smem_a + kResSlots * kColPatchBytes;
  ResBars* bars = reinterpret_cast<ResBars*>(smem_out + 2 * kABytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_per_img = p.tiles_w * p.tiles_h;

  const bool tracing = p.trace != nullptr && static_cast<int>(blockIdx.x) < p.trace_cap;
  long long* trec = tracing ? p.trace + static_cast<long long>(blockIdx.x) * kTraceSlots : nullptr;
  if (threadIdx.x == 0) {
    if (tracing) {
      unsigned long long ns;
      unsigned sm;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
      asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
      trec[kRtEntry] = clock64();
      trec[kTrGlobalNs] = static_cast<long long>(ns);
      trec[kTrSm] = sm;
      trec[kTrTiles] = (total_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    }
#pragma unroll
    for (int s = 0; s < kResSlots; ++s) {
      mbar_init(&bars->a_full[s], 1);
      mbar_init(&bars->a_empty[s], 1);
    }
    mbar_init(&bars->w_full, 1);
    mbar_init(&bars->tmem_full[0], 1);
    mbar_init(&bars->tmem_full[1], 1);
    mbar_init(&bars->tmem_empty[0], 128);
    mbar_init(&bars->tmem_empty[1], 128);
    fence_mbar_init();
    prefetch_tmap(&p.a);
    prefetch_tmap(&p.b);
    prefetch_tmap(&p.out);
  }
  if (warp == 1) tmem_alloc<128>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  if (threadIdx.x == 0 && tracing) trec[kRtSetup] = clock64();
  pdl_wait();
  long long w0c = 0, w1c = 0, w2c = 0, tt = 0;      // trace accumulators (dead code unless tracing)
#define RT_WAIT(acc, stmt)          \
  do {                              \
    if (tracing) tt = clock64();    \
    stmt;                           \
    if (tracing) acc += clock64() - tt; \
  } while (0)

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_expect_tx(&bars->w_full, kResWBytes);
#pragma unroll
      for (int tap = 0; tap < 9; ++tap)      // KRSC map (c, tap, k): box = 64 k-rows x 64 c
        tma_load_3d(&p.b, &bars->w_full, smem_w + tap * 8192, 0, tap, 0);
      uint32_t it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int n0 = t / tiles_per_img, r = t % tiles_per_img;
        const int h0 = (r / p.tiles_w) * 16, w0 = (r % p.tiles_w) * 8;
#pragma unroll
        for (int q = 0; q < 3; ++q, ++it) {
          const int s = it % kResSlots;
          RT_WAIT(w0c, mbar_wait(&bars->a_empty[s], ((it / kResSlots) & 1) ^ 1u));
          mbar_expect_tx(&bars->a_full[s], kColPatchBytes);
          if ((p.l2_hint & (B_MN ? 4 : 1)) != 0)
            tma_load_4d_stream(&p.a, &bars->a_full[s], smem_a + s * kColPatchBytes, 0, w0 + q - 1, h0 - 1, n0);
          else
            tma_load_4d(&p.a, &bars->a_full[s], smem_a + s * kColPatchBytes, 0, w0 + q - 1, h0 - 1, n0);
        }
      }
      if (tracing) trec[kRtProdWaitEmpty] = w0c;
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, false, B_MN);
      const uint32_t w_addr = smem_u32(smem_w);
      mbar_wait(&bars->w_full, 0);
      tc_fence_after();
      uint32_t it = 0, j = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
        const uint32_t acc = j & 1, use = j >> 1;
        RT_WAIT(w1c, mbar_wait(&bars->tmem_empty[acc], (use & 1) ^ 1u));
        tc_fence_after();
        const uint32_t d_addr = tmem_base + acc * 64;
#pragma unroll
        for (int q = 0; q < 3; ++q, ++it) {
          const int s = it % kResSlots;
          RT_WAIT(w0c, mbar_wait(&bars->a_full[s], (it / kResSlots) & 1));
          tc_fence_after();
          const uint32_t patch = smem_u32(smem_a + s * kColPatchBytes);
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const int tap = r * 3 + q;
            const uint32_t b_addr = w_addr + (p.flip_taps ? 8 - tap : tap) * 8192;
            const uint32_t a_addr = patch + r * 1024;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_addr + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
              umma_bf16(d_addr, make_smem_desc_sw128(a_addr + k * 32, 16, 1024), bdesc, idesc,
                        (q | r | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(&bars->a_empty[s]);
        }
        umma_commit(&bars->tmem_full[acc]);
      }
      if (tracing) {
        trec[kRtMmaWaitFull] = w0c;
        trec[kRtMmaWaitAcc] = w1c;
      }
    }
  } else {
    // ------------------------------------------------------------------ two epilogue warpgroups
    const int g = (warp - 2) >> 2;                 // group 0: warps 2-5, group 1: warps 6-9
    const int q4 = warp & 3;                       // TMEM lane quarter this warp may read
    const int row = q4 * 32 + lane;                // accumulator row = h_i * 8 + w_i
    const int wi = row & 7, hi = row >> 3;
    const int et = threadIdx.x - 64 - g * 128;     // 0..127 inside the group
    uint8_t* stage_out = smem_out + g * kABytes;
    const int bar_id = 1 + g;
    float s1[8], s2[8], piv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      s1[i] = s2[i] = 0.f;
      piv[i] = (p.stats != nullptr && p.stat_pivot != nullptr) ? p.stat_pivot[(et & 7) * 8 + i] : 0.f;
    }
    float rsc[8], rsh[8];       // RED: scale / shift of this thread's 8 channels (et & 7)
    if (RED) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        rsc[i] = p.red_scale[(et & 7) * 8 + i];
        rsh[i] = p.red_shift[(et & 7) * 8 + i];
      }
    }
    uint32_t u = 0;
    bool store_pending = false;
    for (int t = blockIdx.x + g * gridDim.x; t < total_tiles; t += 2 * gridDim.x, ++u) {
      const int n0 = t / tiles_per_img, r = t % tiles_per_img;
      const int h0 = (r / p.tiles_w) * 16, w0 = (r % p.tiles_w) * 8;
      const __nv_bfloat16* add_row = nullptr;
      const unsigned char* add_bits = nullptr;
      if (p.addend != nullptr) {
        const long long pix = (static_cast<long long>(n0) * p.GH + (h0 + hi)) * p.GW + (w0 + wi);
        add_row = p.addend + pix * 64;
        if (p.addend_bits != nullptr) add_bits = p.addend_bits + pix * 8;
        // this row's addend (128 bytes) is read after the accumulator wait: pull its line towards the SM now
        asm volatile("prefetch.global.L2 [%0];" ::"l"(add_row));
      }
      uint4 yv[8];
      if (RED) {
        // this thread's cells of the y tile: 8 channels (et & 7) of rows (et >> 3) + 16 i; a warp reads whole 128-byte lines
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = (et >> 3) + 16 * i;
          const long long pix = (static_cast<long long>(n0) * p.GH + (h0 + (rr >> 3))) * p.GW + (w0 + (rr & 7));
          const uint4* yp = reinterpret_cast<const uint4*>(p.red_y + pix * 64 + (et & 7) * 8);
          yv[i] = (p.l2_hint & 32) != 0 ? ld_stream16(yp) : __ldg(yp);
        }
      }
      RT_WAIT(w0c, mbar_wait(&bars->tmem_full[g], u & 1));
      tc_fence_after();
      // the previous store of this group must have finished reading the staging buffer
      if (et == 0 && store_pending) RT_WAIT(w1c, tma_store_wait_read0());
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (tracing) tt = clock64();
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q4 * 32) << 16) + g * 64 + cc * 32, v);
        tmem_ld_wait();
        uint32_t packed[16];
        if (add_row != nullptr) {
          const uint4* ap = reinterpret_cast<const uint4*>(add_row + cc * 32);
          const uint32_t am = add_bits != nullptr ? *reinterpret_cast<const uint32_t*>(add_bits + cc * 4) : 0xffffffffu;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint4 a4 = ap[jj];
            const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float2 af = unpack_bf16x2(aw[e]);
              if (!((am >> (jj * 8 + e * 2)) & 1u)) af.x = 0.f;
              if (!((am >> (jj * 8 + e * 2 + 1)) & 1u)) af.y = 0.f;
              packed[jj * 4 + e] = pack_bf16x2(__uint_as_float(v[jj * 8 + e * 2]) + af.x,
                                               __uint_as_float(v[jj * 8 + e * 2 + 1]) + af.y);
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) packed[i] = pack_bf16x2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
        }
        uint8_t* rowp = stage_out + row * 128;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int chunk = (cc * 4 + jj) ^ (row & 7);
          *reinterpret_cast<uint4*>(rowp + chunk * 16) =
              make_uint4(packed[jj * 4], packed[jj * 4 + 1], packed[jj * 4 + 2], packed[jj * 4 + 3]);
        }
      }
      tc_fence_before();
      mbar_arrive(&bars->tmem_empty[g]);          // this thread's TMEM reads are complete: accumulator g is free
      if (RED) {
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");     // the whole tile is staged
        const int cg = et & 7;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = (et >> 3) + 16 * i;
          uint4* cell = reinterpret_cast<uint4*>(stage_out + rr * 128 + ((cg ^ (rr & 7)) << 4));
          const uint4 g4 = *cell;
          uint32_t gw[4] = {g4.x, g4.y, g4.z, g4.w};
          const uint32_t yw[4] = {yv[i].x, yv[i].y, yv[i].z, yv[i].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 gf = unpack_bf16x2(gw[e]);
            const float2 yf = unpack_bf16x2(yw[e]);
            if (!(fmaf(yf.x, rsc[2 * e], rsh[2 * e]) > 0.f)) gf.x = 0.f;
            if (!(fmaf(yf.y, rsc[2 * e + 1], rsh[2 * e + 1]) > 0.f)) gf.y = 0.f;
            s1[2 * e] += gf.x;
            s2[2 * e] = fmaf(gf.x, yf.x, s2[2 * e]);
            s1[2 * e + 1] += gf.y;
            s2[2 * e + 1] = fmaf(gf.y, yf.y, s2[2 * e + 1]);
            gw[e] = pack_bf16x2(gf.x, gf.y);
          }
          *cell = make_uint4(gw[0], gw[1], gw[2], gw[3]);
        }
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (et == 0) {
        tma_store_4d(&p.out, stage_out, 0, w0, h0, n0);
        tma_store_commit();
      }
      store_pending = true;
      if (p.stats != nullptr) {
        // thread -> 16-byte channel group (et & 7) of rows (et >> 3) + 16 i: one LDS.128 covers 8 channels
        const int cg = et & 7;
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
      if (tracing) w2c += clock64() - tt;
    }
    if (tracing && g == 0 && et == 0) {
      trec[kRtEpiWaitAcc] = w0c;
      trec[kRtEpiWaitStore] = w1c;
      trec[kRtEpiBusy] = w2c;
    }
    if (et == 0 && store_pending) tma_store_wait_read0();
    if (p.stats != nullptr || RED) {
      // lanes with equal (lane & 7) hold the same channel group: fold 4 lanes, then the 4 warps via shared memory
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 8);
        s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 16);
        s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 8);
        s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 16);
      }
      float* part = bars->stat_part[g];
      if (et < 128) part[et] = 0.f;
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (lane < 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          atomicAdd(&part[lane * 8 + i], s1[i]);
          atomicAdd(&part[64 + lane * 8 + i], s2[i]);
        }
      }
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
      if (RED) {
        // S_g = sum(g), S_gy = sum(g*y): sum(g*xhat) = (S_gy - mean*S_g)*invstd is linear in them
        if (et < 64) {
          red_add_f32(p.red_sums + et, part[et]);
          red_add_f32(p.red_sums + 64 + et, (part[64 + et] - p.red_mean[et] * part[et]) * p.red_invstd[et]);
        }
      } else {
        red_add_f32(p.stats + (et < 64 ? et : p.out_channels + et - 64), part[et]);
      }
    }
  }
#undef RT_WAIT
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0 && tracing) trec[kRtExit] = clock64();
  if (warp == 1) tmem_dealloc<128>(tmem_base);
}

// ============================================================================ wgrad
// MB = number of 128-row accumulators (pairs of x boxes) per CTA.  With MB = 2 the dy tile of a stage feeds twice as
// many MMAs: (32 + 32) KB per 1024 tensor cycles for BNW = 256 instead of (16 + 32) KB per 512, i.e. a third less
// operand traffic out of L2, which is what bounds this kernel.
template <int BNW, int STAGES, int MB = 1>
__global__ void __launch_bounds__(kThreads) conv_wgrad_kernel(const __grid_constant__ WgradParams p) {
  constexpr int kBoxBytes = 8192;                       // 64 pixels x 128 B
  constexpr int kABox = 2 * MB;                         // x boxes per stage
  constexpr int kStageBytes = kABox * kBoxBytes + (BNW / 64) * kBoxBytes;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  PipeBars<STAGES>* bars = reinterpret_cast<PipeBars<STAGES>*>(smem + STAGES * kStageBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int box0 = kABox * blockIdx.x;
  const int k0 = blockIdx.y * BNW;
  const int tiles_total = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (tiles_total + gridDim.z - 1) / gridDim.z;
  const int pt_begin = blockIdx.z * per;
  const int pt_end = min(pt_begin + per, tiles_total);
  const int niter = max(pt_end - pt_begin, 0);

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
    }
    mbar_init(&bars->tmem_full, 1);
    fence_mbar_init();
    prefetch_tmap(&p.dy);
    prefetch_tmap(&p.x[0]);
  }
  if (warp == 1) tmem_alloc<MB * BNW>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      WBox bx[kABox];
#pragma unroll
      for (int j = 0; j < kABox; ++j) bx[j] = p.box[min(box0 + j, p.n_boxes - 1)];
      for (int i = 0; i < niter; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        int t = pt_begin + i;
        const int w0 = (t % p.tiles_w) * p.w_box;
        t /= p.tiles_w;
        const int h0 = (t % p.tiles_h) * p.h_box;
        const int n0 = (t / p.tiles_h) * p.n_box;
        mbar_wait(&bars->empty[s], ph ^ 1u);
        uint8_t* sa = smem + s * kStageBytes;
        uint8_t* sb = sa + kABox * kBoxBytes;
        mbar_expect_tx(&bars->full[s], kStageBytes);
#pragma unroll
        for (int j = 0; j < kABox; ++j)
          if ((p.l2_hint & 8) != 0)
            tma_load_4d_stream(&p.x[bx[j].a_map], &bars->full[s], sa + j * kBoxBytes, bx[j].a_ch, w0 + bx[j].dw, h0 + bx[j].dh, n0);
          else
            tma_load_4d(&p.x[bx[j].a_map], &bars->full[s], sa + j * kBoxBytes, bx[j].a_ch, w0 + bx[j].dw, h0 + bx[j].dh, n0);
#pragma unroll
        for (int j = 0; j < BNW / 64; ++j)
          if ((p.l2_hint & 16) != 0) tma_load_4d_stream(&p.dy, &bars->full[s], sb + j * kBoxBytes, k0 + j * 64, w0, h0, n0);
          else tma_load_4d(&p.dy, &bars->full[s], sb + j * kBoxBytes, k0 + j * 64, w0, h0, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BNW, true, true);
      for (int i = 0; i < niter; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&bars->full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
        const uint32_t b_addr = a_addr + kABox * kBoxBytes;
#pragma unroll
        for (int m = 0; m < MB; ++m) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {   // 64 pixels = 4 x (K = 16 rows of 128 B)
            umma_bf16(tmem_base + m * BNW, make_smem_desc_sw128(a_addr + m * 2 * kBoxBytes + k * 2048, kBoxBytes, 1024),
                      make_smem_desc_sw128(b_addr + k * 2048, kBoxBytes, 1024), idesc, (i | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(&bars->empty[s]);
      }
      if (niter > 0) umma_commit(&bars->tmem_full);
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int c = row & 63;
    if (niter > 0) {
      mbar_wait(&bars->tmem_full, 0);
      tc_fence_after();
    }
#pragma unroll 1
    for (int m = 0; m < MB; ++m) {
      const int bx = 2 * m + (row >> 6);
      const bool live = (niter > 0) && (box0 + bx < p.n_boxes) && (c < p.c_valid);
      const int out_col = p.box[min(box0 + bx, p.n_boxes - 1)].out_col + c;
      if (niter <= 0) break;
#pragma unroll 1
      for (int cc = 0; cc < BNW / 32; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + m * BNW + cc * 32, v);
        tmem_ld_wait();
        if (live) {
          float* dst = p.dw + static_cast<long long>(k0 + cc * 32) * p.RSC + out_col;
#pragma unroll
          for (int j = 0; j < 32; ++j) red_add_f32(dst + static_cast<long long>(j) * p.RSC, __uint_as_float(v[j]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<MB * BNW>(tmem_base);
}

// ============================================================================ wgrad, all nine taps per CTA
// For 3x3 stride-1 layers with few output channels the kernel above is bound by x traffic: every 64-pixel tile of x is
// fetched once per tap (9 x 8 KB) and the small dy tile (8 KB) cannot amortise it (layer1: 786 MB through L2 = the
// whole 78 us).  Here a CTA owns one 64-channel chunk of C, one 64-channel chunk of K and ALL nine taps: per 8x8-pixel
// tile it loads three column patches of x (8 w x 10 h pixels, one per horizontal tap offset, as in
// conv3x3_resident_kernel) and one dy tile -- 38 KB instead of 96 KB -- and issues five M=128 MMAs per 16 pixels, each
// covering two taps: the A descriptor is MN-major, so its two 64-row halves may sit anywhere in shared memory (LBO =
// distance between the two tap views; a vertical tap offset is a whole 1024-byte swizzle atom, so every view starts
// atom-aligned).  Five accumulators of 64 columns live in TMEM; tap order inside the accumulators is (s, r).
constexpr int kWPatchBytes = 8 * 10 * 128;                 // 10240
constexpr int kWPatchStage = 3 * kWPatchBytes + 8192;      // 38912 = 38 atoms
__device__ __forceinline__ uint32_t wpatch_view(int u) {   // byte offset of tap view u = s*3 + r inside a stage
  return static_cast<uint32_t>((u / 3) * kWPatchBytes + (u % 3) * 1024);
}

template <int STAGES>
__global__ void __launch_bounds__(kThreads) conv_wgrad_patch_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  PipeBars<STAGES>* bars = reinterpret_cast<PipeBars<STAGES>*>(smem + STAGES * kWPatchStage);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.x * 64, k0 = blockIdx.y * 64;
  const int tiles_total = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (tiles_total + gridDim.z - 1) / gridDim.z;
  const int pt_begin = blockIdx.z * per;
  const int pt_end = min(pt_begin + per, tiles_total);
  const int niter = max(pt_end - pt_begin, 0);

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
    }
    mbar_init(&bars->tmem_full, 1);
    fence_mbar_init();
    prefetch_tmap(&p.dy);
    prefetch_tmap(&p.x[0]);
  }
  if (warp == 1) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < niter; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        int t = pt_begin + i;
        const int w0 = (t % p.tiles_w) * 8;
        t /= p.tiles_w;
        const int h0 = (t % p.tiles_h) * 8;
        const int n0 = t / p.tiles_h;
        mbar_wait(&bars->empty[s], ph ^ 1u);
        uint8_t* sa = smem + s * kWPatchStage;
        mbar_expect_tx(&bars->full[s], kWPatchStage);
#pragma unroll
        for (int q = 0; q < 3; ++q)
          if ((p.l2_hint & 8) != 0) tma_load_4d_stream(&p.x[0], &bars->full[s], sa + q * kWPatchBytes, c0, w0 + q - 1, h0 - 1, n0);
          else tma_load_4d(&p.x[0], &bars->full[s], sa + q * kWPatchBytes, c0, w0 + q - 1, h0 - 1, n0);
        if ((p.l2_hint & 16) != 0) tma_load_4d_stream(&p.dy, &bars->full[s], sa + 3 * kWPatchBytes, k0, w0, h0, n0);
        else tma_load_4d(&p.dy, &bars->full[s], sa + 3 * kWPatchBytes, k0, w0, h0, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, true, true);
      for (int i = 0; i < niter; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(&bars->full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * kWPatchStage);
        const uint32_t b_addr = a_addr + 3 * kWPatchBytes;
#pragma unroll
        for (int pr = 0; pr < 5; ++pr) {
          const uint32_t v0 = wpatch_view(2 * pr);
          const uint32_t lbo = pr < 4 ? wpatch_view(2 * pr + 1) - v0 : 1024u;   // pair 4: second half is padding
#pragma unroll
          for (int k = 0; k < 4; ++k) {   // 64 pixels = 4 x (K = 16 pixel rows of 128 B)
            umma_bf16(tmem_base + pr * 64, make_smem_desc_sw128(a_addr + v0 + k * 2048, lbo, 1024),
                      make_smem_desc_sw128(b_addr + k * 2048, 8192, 1024), idesc, (i | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(&bars->empty[s]);
      }
      if (niter > 0) umma_commit(&bars->tmem_full);
    }
  } else if (niter > 0) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int c = row & 63;
    mbar_wait(&bars->tmem_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int pr = 0; pr < 5; ++pr) {
      const int u = 2 * pr + (row >> 6);              // tap view (s, r) of this accumulator row
      const bool live = u < 9;
      const int tap = (u % 3) * 3 + (u / 3);          // weight tap index r*3 + s
      const int out_col = tap * p.C + c0 + c;
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + pr * 64 + cc * 32, v);
        tmem_ld_wait();
        if (live) {
          float* dst = p.dw + static_cast<long long>(k0 + cc * 32) * p.RSC + out_col;
#pragma unroll
          for (int j = 0; j < 32; ++j) red_add_f32(dst + static_cast<long long>(j) * p.RSC, __uint_as_float(v[j]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ============================================================================ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

int encode_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
               const uint32_t* box) {
  EncodeTiledFn fn = get_encode_fn();
  PMD_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  cuuint64_t gd[5], gs[5];
  cuuint32_t bx[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
  }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gd, gs, bx, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(r)));
    return -1;
  }
  return 0;
}

// NHWC bf16 activation [N,H,W,C]; view = every `step`-th row/column starting at (ph, pw).
int act_map(CUtensorMap* m, const void* base, int N, int H, int W, int C, int step, int ph, int pw, int w_box,
            int h_box, int n_box) {
  const int Hv = (H - ph + step - 1) / step, Wv = (W - pw + step - 1) / step;
  const char* b = static_cast<const char*>(base) + (static_cast<long long>(ph) * W + pw) * C * 2;
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)Wv, (uint64_t)Hv, (uint64_t)N};
  const uint64_t strides[3] = {(uint64_t)step * C * 2, (uint64_t)step * W * C * 2, (uint64_t)H * W * C * 2};
  const uint32_t box[4] = {64u, (uint32_t)w_box, (uint32_t)h_box, (uint32_t)n_box};
  return encode_map(m, b, 4, dims, strides, box);
}

int largest_pow2_le(int x) {
  int p = 1;
  while (p * 2 <= x) p *= 2;
  return p;
}

// Split P (= 64 or 128) pixels into a w x h x n brick; prefer box sides that divide the grid.
void choose_brick(int P, int GW, int GH, int* w_box, int* h_box, int* n_box) {
  auto pick = [](int limit, int extent) {
    int best = 1;
    for (int c = 1; c <= limit; c *= 2)
      if (extent % c == 0) best = c;
    if (best < 4 && extent >= 4) best = largest_pow2_le(extent < limit ? extent : limit);   // awkward extent: accept OOB waste
    return best;
  };
  *w_box = pick(P < 32 ? P : 32, GW);
  *h_box = pick(P / *w_box, GH);
  *n_box = P / (*w_box * *h_box);
}

}  // namespace

int conv_l2_hint_mask() {
  static const int mask = getenv("PMD_L2_HINT") == nullptr ? 27 : atoi(getenv("PMD_L2_HINT"));
  return mask;
}

int encode_tiled_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box) {
  return encode_map(m, base, rank, dims, strides_bytes, box);
}

// Output-channel tile of the fprop/dgrad GEMM.  256-wide tiles halve the A traffic and take an N=256 MMA under the
// shared-memory operand limit (96 B/clk instead of 128), but leave half as many CTAs; PMD_GEMM_BN256: 0 never,
// 1 whenever the channel count allows, 2 (default) only when at least `148` CTAs remain.
static int pick_bn_tile(int out_channels, int pixel_tiles, int classes) {
  if (const char* f = getenv("PMD_GEMM_BN")) {      // experiments: force the tile width where the channel count allows it
    const int bn = atoi(f);
    if ((bn == 64 || bn == 128 || bn == 256) && out_channels % bn == 0) return bn;
  }
  static const int mode = getenv("PMD_GEMM_BN256") == nullptr ? 2 : atoi(getenv("PMD_GEMM_BN256"));
  if (out_channels % 256 == 0 && mode != 0) {
    const int ctas = pixel_tiles * (out_channels / 256) * classes;
    if (mode == 1 || ctas >= 148) return 256;
  }
  return out_channels % 128 == 0 ? 128 : 64;
}

// CTA pairs (cta_group::2) for 128- and 256-wide channel tiles when the pixel tiles pair up: PMD_GEMM_2CTA=1.
// Off by default: correct, but measured no faster than one CTA per tile (profiles/r2_conv_cta_pairs.md).
static int pick_two_cta(int bn_tile, int pixel_tiles) {
  const char* v = getenv("PMD_GEMM_2CTA");
  const bool on = v != nullptr && atoi(v) != 0;
  return (on && bn_tile >= 128 && pixel_tiles % 2 == 0) ? 1 : 0;
}

int build_fprop_params(const PmdConvShape& s, const void* x, const void* w_krsc, void* y, float* stats,
                       ConvGemmParams* out) {
  PMD_REQUIRE(conv_tc_supported(s), "conv shape not supported by the tensor-core path");
  ConvGemmParams& p = *out;
  memset(&p, 0, sizeof(p));
  p.l2_hint = conv_l2_hint_mask();
  const int pad = s.R / 2, RS = s.R * s.R;
  const int Ho = conv_out_dim(s.H, s.R, s.stride), Wo = conv_out_dim(s.W, s.R, s.stride);
  choose_brick(128, Wo, Ho, &p.w_box, &p.h_box, &p.n_box);
  p.tiles_w = ceil_div(Wo, p.w_box);
  p.tiles_h = ceil_div(Ho, p.h_box);
  p.tiles_n = ceil_div(s.N, p.n_box);
  p.N = s.N;
  p.n_classes = 1;
  p.GW[0] = Wo;
  p.GH[0] = Ho;
  p.b_tap_last = 0;
  p.out_channels = s.K;
  p.bn_tile = pick_bn_tile(s.K, p.tiles_w * p.tiles_h * p.tiles_n, 1);
  p.two_cta = pick_two_cta(p.bn_tile, p.tiles_w * p.tiles_h * p.tiles_n);
  p.stats = stats;
  p.addend = nullptr;
  if (s.stride == 1) {
    if (act_map(&p.a[0], x, s.N, s.H, s.W, s.C, 1, 0, 0, p.w_box, p.h_box, p.n_box)) return -1;
  } else {
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw)
        if (act_map(&p.a[ph * 2 + pw], x, s.N, s.H, s.W, s.C, 2, ph, pw, p.w_box, p.h_box, p.n_box)) return -1;
  }
  {
    const uint64_t dims[3] = {(uint64_t)s.C, (uint64_t)RS, (uint64_t)s.K};
    const uint64_t strides[2] = {(uint64_t)s.C * 2, (uint64_t)RS * s.C * 2};
    const uint32_t box[3] = {64u, 1u, (uint32_t)(p.two_cta ? p.bn_tile / 2 : p.bn_tile)};
    if (encode_map(&p.b[0], w_krsc, 3, dims, strides, box)) return -1;
  }
  if (act_map(&p.out[0], y, s.N, Ho, Wo, s.K, 1, 0, 0, p.w_box, p.h_box, p.n_box)) return -1;
  int n = 0;
  for (int r = 0; r < s.R; ++r)
    for (int q = 0; q < s.R; ++q) {
      int a_map = 0, dh, dw;
      if (s.stride == 1) {
        dh = r - pad;
        dw = q - pad;
      } else {   // input row = 2*ho - pad + r = 2*(ho + dh) + ph
        const int oh = r - pad, ow = q - pad;
        const int ph = ((oh % 2) + 2) % 2, pw = ((ow % 2) + 2) % 2;
        dh = (oh - ph) / 2;
        dw = (ow - pw) / 2;
        a_map = ph * 2 + pw;
      }
      for (int c0 = 0; c0 < s.C; c0 += 64) {
        PMD_REQUIRE(n < kMaxKBlocks, "too many k-blocks for one convolution");
        KBlock& kb = p.kb[n++];
        kb.dh = (int8_t)dh;
        kb.dw = (int8_t)dw;
        kb.a_map = (uint8_t)a_map;
        kb.b_map = 0;
        kb.a_ch = (uint16_t)c0;
        kb.b_c0 = (uint16_t)c0;
        kb.tap = (uint8_t)(r * s.R + q);
      }
    }
  p.kb_begin[0] = 0;
  p.kb_count[0] = n;
  return 0;
}

int build_dgrad_params(const PmdConvShape& s, const void* dy, const void* w_rsck, const PmdConvShape* s2,
                       const void* dy2, const void* w2_rsck, const void* addend, void* dx, ConvGemmParams* out,
                       bool w_is_krsc) {
  PMD_REQUIRE(conv_tc_supported(s), "conv shape not supported by the tensor-core path");
  ConvGemmParams& p = *out;
  memset(&p, 0, sizeof(p));
  p.l2_hint = conv_l2_hint_mask();
  const int Ho = conv_out_dim(s.H, s.R, s.stride), Wo = conv_out_dim(s.W, s.R, s.stride);
  if (s2 != nullptr) {
    PMD_REQUIRE(conv_tc_supported(*s2) && s2->N == s.N && s2->H == s.H && s2->W == s.W && s2->C == s.C &&
                    s2->stride == s.stride && conv_out_dim(s2->H, s2->R, s2->stride) == Ho,
                "fused second dgrad must consume the same input tensor with the same stride");
  }
  PMD_REQUIRE(addend == nullptr || s.stride == 1, "addend is only supported for stride-1 dgrad");
  const int st = s.stride;
  p.n_classes = st * st;
  int maxGW = 0, maxGH = 0;
  for (int ph = 0; ph < st; ++ph)
    for (int pw = 0; pw < st; ++pw) {
      const int cls = ph * st + pw;
      p.GH[cls] = (s.H - ph + st - 1) / st;
      p.GW[cls] = (s.W - pw + st - 1) / st;
      maxGW = p.GW[cls] > maxGW ? p.GW[cls] : maxGW;
      maxGH = p.GH[cls] > maxGH ? p.GH[cls] : maxGH;
    }
  choose_brick(128, maxGW, maxGH, &p.w_box, &p.h_box, &p.n_box);
  p.tiles_w = ceil_div(maxGW, p.w_box);
  p.tiles_h = ceil_div(maxGH, p.h_box);
  p.tiles_n = ceil_div(s.N, p.n_box);
  p.N = s.N;
  p.b_tap_last = 1;
  p.out_channels = s.C;
  p.bn_tile = pick_bn_tile(s.C, p.tiles_w * p.tiles_h * p.tiles_n, st * st);
  p.two_cta = pick_two_cta(p.bn_tile, p.tiles_w * p.tiles_h * p.tiles_n);
  p.stats = nullptr;
  p.addend = static_cast<const __nv_bfloat16*>(addend);
  // A operands: the output-gradient tensors on the (Ho, Wo) grid
  if (act_map(&p.a[0], dy, s.N, Ho, Wo, s.K, 1, 0, 0, p.w_box, p.h_box, p.n_box)) return -1;
  if (s2 != nullptr && act_map(&p.a[1], dy2, s.N, Ho, Wo, s2->K, 1, 0, 0, p.w_box, p.h_box, p.n_box)) return -1;
  // B operands: RSCK weights as (k inner, c rows, tap)
  p.b_mn_major = w_is_krsc ? 1 : 0;
  auto wmap = [&](CUtensorMap* m, const void* w, const PmdConvShape& cs) {
    if (w_is_krsc) {   // (c inner, tap, k): boxes of 64 c x 64 k, consumed MN-major
      const uint64_t dims[3] = {(uint64_t)cs.C, (uint64_t)(cs.R * cs.R), (uint64_t)cs.K};
      const uint64_t strides[2] = {(uint64_t)cs.C * 2, (uint64_t)(cs.R * cs.R) * cs.C * 2};
      const uint32_t box[3] = {64u, 1u, 64u};
      return encode_map(m, w, 3, dims, strides, box);
    }
    const uint64_t dims[3] = {(uint64_t)cs.K, (uint64_t)cs.C, (uint64_t)(cs.R * cs.R)};
    const uint64_t strides[2] = {(uint64_t)cs.K * 2, (uint64_t)cs.C * cs.K * 2};
    const uint32_t box[3] = {64u, (uint32_t)(p.two_cta ? p.bn_tile / 2 : p.bn_tile), 1u};
    return encode_map(m, w, 3, dims, strides, box);
  };
  if (wmap(&p.b[0], w_rsck, s)) return -1;
  if (s2 != nullptr && wmap(&p.b[1], w2_rsck, *s2)) return -1;
  // outputs: one parity view of dx per class
  for (int ph = 0; ph < st; ++ph)
    for (int pw = 0; pw < st; ++pw)
      if (act_map(&p.out[ph * st + pw], dx, s.N, s.H, s.W, s.C, st, ph, pw, p.w_box, p.h_box, p.n_box)) return -1;
  int n = 0;
  for (int ph = 0; ph < st; ++ph)
    for (int pw = 0; pw < st; ++pw) {
      const int cls = ph * st + pw;
      p.kb_begin[cls] = n;
      for (int which = 0; which < (s2 != nullptr ? 2 : 1); ++which) {
        const PmdConvShape& cs = which == 0 ? s : *s2;
        const int pad = cs.R / 2;
        for (int r = 0; r < cs.R; ++r)
          for (int q = 0; q < cs.R; ++q) {
            // input row h = st*ho - pad + r  ->  ho = (h + pad - r)/st with h = st*a + ph
            const int nh = ph + pad - r, nw = pw + pad - q;
            if (((nh % st) + st) % st != 0 || ((nw % st) + st) % st != 0) continue;
            const int dh = nh / st, dw = nw / st;    // exact (checked above), may be negative
            for (int k0 = 0; k0 < cs.K; k0 += 64) {
              PMD_REQUIRE(n < kMaxKBlocks, "too many k-blocks for one dgrad");
              KBlock& kb = p.kb[n++];
              kb.dh = (int8_t)dh;
              kb.dw = (int8_t)dw;
              kb.a_map = (uint8_t)which;
              kb.b_map = (uint8_t)which;
              kb.a_ch = (uint16_t)k0;
              kb.b_c0 = (uint16_t)k0;
              kb.tap = (uint8_t)(r * cs.R + q);
            }
          }
      }
      p.kb_count[cls] = n - p.kb_begin[cls];
    }
  return 0;
}

int build_wgrad_params(const PmdConvShape& s, const void* x, const void* dy, float* dw, int split, int sm_count,
                       WgradParams* out, int rsc_override, int c_valid) {
  PMD_REQUIRE(conv_tc_supported(s), "conv shape not supported by the tensor-core path");
  WgradParams& p = *out;
  memset(&p, 0, sizeof(p));
  p.l2_hint = conv_l2_hint_mask();
  const int pad = s.R / 2, RS = s.R * s.R;
  const int Ho = conv_out_dim(s.H, s.R, s.stride), Wo = conv_out_dim(s.W, s.R, s.stride);
  {
    // PMD_WGRAD_PATCH: 1 (default) all-taps-per-CTA kernel for 3x3 stride-1 layers with K == 64, 2 for every K, 0 never
    static const int patch_mode = getenv("PMD_WGRAD_PATCH") == nullptr ? 1 : atoi(getenv("PMD_WGRAD_PATCH"));
    const bool shape_ok = s.R == 3 && s.stride == 1 && s.W % 8 == 0 && s.H % 8 == 0 && rsc_override == 0 && c_valid == 64;
    if (shape_ok && (patch_mode == 2 || (patch_mode == 1 && s.K == 64))) {
      p.patch = 1;
      p.C = s.C;
      p.K = s.K;
      p.RSC = RS * s.C;
      p.c_valid = 64;
      p.dw = dw;
      p.w_box = 8;
      p.h_box = 8;
      p.n_box = 1;
      p.tiles_w = s.W / 8;
      p.tiles_h = s.H / 8;
      p.tiles_n = s.N;
      if (act_map(&p.x[0], x, s.N, s.H, s.W, s.C, 1, 0, 0, 8, 10, 1)) return -1;
      if (act_map(&p.dy, dy, s.N, s.H, s.W, s.K, 1, 0, 0, 8, 8, 1)) return -1;
      const int tiles_total = p.tiles_w * p.tiles_h * p.tiles_n;
      const int ctas = (s.C / 64) * (s.K / 64);
      if (split <= 0) {
        split = (sm_count > 0 ? sm_count : 148) / ctas;     // one CTA per SM (152 KB of stages, all of TMEM): one wave
        const int max_split = ceil_div(tiles_total, 8);
        if (split > max_split) split = max_split;
      }
      if (split > tiles_total) split = tiles_total;
      if (split < 1) split = 1;
      p.split = split;
      return 0;
    }
  }
  choose_brick(64, Wo, Ho, &p.w_box, &p.h_box, &p.n_box);
  p.tiles_w = ceil_div(Wo, p.w_box);
  p.tiles_h = ceil_div(Ho, p.h_box);
  p.tiles_n = ceil_div(s.N, p.n_box);
  p.K = s.K;
  p.RSC = rsc_override > 0 ? rsc_override : RS * s.C;
  p.c_valid = c_valid;
  p.dw = dw;
  p.bn_tile = (s.K % 256 == 0) ? 256 : (s.K % 128 == 0 ? 128 : 64);
  if (s.stride == 1) {
    if (act_map(&p.x[0], x, s.N, s.H, s.W, s.C, 1, 0, 0, p.w_box, p.h_box, p.n_box)) return -1;
  } else {
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw)
        if (act_map(&p.x[ph * 2 + pw], x, s.N, s.H, s.W, s.C, 2, ph, pw, p.w_box, p.h_box, p.n_box)) return -1;
  }
  if (act_map(&p.dy, dy, s.N, Ho, Wo, s.K, 1, 0, 0, p.w_box, p.h_box, p.n_box)) return -1;
  int n = 0;
  for (int r = 0; r < s.R; ++r)
    for (int q = 0; q < s.R; ++q) {
      int a_map = 0, dh, dw_;
      if (s.stride == 1) {
        dh = r - pad;
        dw_ = q - pad;
      } else {
        const int oh = r - pad, ow = q - pad;
        const int ph = ((oh % 2) + 2) % 2, pw = ((ow % 2) + 2) % 2;
        dh = (oh - ph) / 2;
        dw_ = (ow - pw) / 2;
        a_map = ph * 2 + pw;
      }
      for (int c0 = 0; c0 < s.C; c0 += 64) {
        PMD_REQUIRE(n < kMaxWBoxes, "too many operand boxes for one wgrad");
        WBox& b = p.box[n++];
        b.dh = (int8_t)dh;
        b.dw = (int8_t)dw_;
        b.a_map = (uint8_t)a_map;
        b.a_ch = (uint16_t)c0;
        b.out_col = (r * s.R + q) * s.C + c0;
      }
    }
  p.n_boxes = n;
  const int tiles_total = p.tiles_w * p.tiles_h * p.tiles_n;
  // PMD_WGRAD_MB: 0 (default) two accumulators per CTA when the boxes fill groups of four exactly or there are at
  // least 8 of them (<= 25 % padding; measured better than MB = 1 even for the 9-box 64-channel layers), 1, 2
  static const int mb_mode = getenv("PMD_WGRAD_MB") == nullptr ? 0 : atoi(getenv("PMD_WGRAD_MB"));
  p.mb = mb_mode == 0 ? ((n % 4 == 0 || n >= 8) ? 2 : 1) : (mb_mode == 2 ? 2 : 1);
  if (split <= 0) {
    const int sms = sm_count > 0 ? sm_count : 148;
    const int ctas = ceil_div(n, 2 * p.mb) * (s.K / p.bn_tile);
    if (p.mb == 2) {
      split = sms / ctas;                                // one CTA per SM fits (>= 144 KB of stages): exactly one wave
      if (split < 1) split = 1;
    } else {
      static const int target = getenv("PMD_WGRAD_TARGET") == nullptr ? 0 : atoi(getenv("PMD_WGRAD_TARGET"));
      split = target > 0 ? ceil_div(target * sms, ctas) : (2 * sms) / ctas;   // 0: exactly one wave of 2 CTAs per SM
      if (split < 1) split = 1;
    }
    const int max_split = ceil_div(tiles_total, 8);     // keep >= 8 pipeline iterations per CTA
    if (split > max_split) split = max_split;
  }
  if (split > tiles_total) split = tiles_total;
  if (split < 1) split = 1;
  p.split = split;
  return 0;
}

static long long* g_trace_buf = nullptr;   // pmd_conv_set_trace
static int g_trace_cap = 0;

template <typename Kernel>
static int set_smem(Kernel k, int bytes) {
  PMD_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return 0;
}

// ---------------------------------------------------------------------------- weight-resident 3x3 kernel, host side
static int build_resident_common(const PmdConvShape& s, const void* in, const void* w_krsc, void* outp,
                                 ConvResidentParams* o) {
  PMD_REQUIRE(conv_resident_supported(s), "shape not eligible for the weight-resident 3x3 kernel");
  ConvResidentParams& p = *o;
  memset(&p, 0, sizeof(p));
  p.l2_hint = conv_l2_hint_mask();
  p.tiles_w = s.W / 8;
  p.tiles_h = s.H / 16;
  p.N = s.N;
  p.GW = s.W;
  p.GH = s.H;
  p.out_channels = 64;
  if (act_map(&p.a, in, s.N, s.H, s.W, 64, 1, 0, 0, 8, 18, 1)) return -1;
  if (act_map(&p.out, outp, s.N, s.H, s.W, 64, 1, 0, 0, 8, 16, 1)) return -1;
  const uint64_t dims[3] = {64ull, 9ull, 64ull};   // KRSC: (c inner, tap, k)
  const uint64_t strides[2] = {64ull * 2, 9ull * 64 * 2};
  const uint32_t box[3] = {64u, 1u, 64u};
  return encode_map(&p.b, w_krsc, 3, dims, strides, box);
}

int build_fprop_resident_params(const PmdConvShape& s, const void* x, const void* w_krsc, void* y, float* stats,
                                ConvResidentParams* out) {
  if (build_resident_common(s, x, w_krsc, y, out)) return -1;
  out->stats = stats;
  return 0;
}

int build_dgrad_resident_params(const PmdConvShape& s, const void* dy, const void* w_krsc, const void* addend, void* dx,
                                ConvResidentParams* out) {
  // same-size 3x3 stride-1: dx = conv(dy, flipped w); the KRSC weights are consumed MN-major
  if (build_resident_common(s, dy, w_krsc, dx, out)) return -1;
  out->b_mn = 1;
  out->flip_taps = 1;
  out->addend = static_cast<const __nv_bfloat16*>(addend);
  return 0;
}

void attach_bn_reduce_resident(ConvResidentParams* p, const void* y, const float* scale, const float* shift,
                               const float* mean, const float* invstd, float* sums) {
  p->red_y = static_cast<const __nv_bfloat16*>(y);
  p->red_scale = scale;
  p->red_shift = shift;
  p->red_mean = mean;
  p->red_invstd = invstd;
  p->red_sums = sums;
}

template <bool B_MN, bool RED = false>
static int launch_conv_resident_t(const ConvResidentParams& p, cudaStream_t stream) {
  static PerDevice<int> sms_of;      // also marks "shared-memory attribute set on this device"
  int* sms = sms_of.get();
  if (sms == nullptr) return -1;
  if (*sms == 0) {
    if (set_smem(conv3x3_resident_kernel<B_MN, RED>, kResSmem)) return -1;
    int dev = 0;
    PMD_CHECK_CUDA(cudaGetDevice(&dev));
    PMD_CHECK_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
  }
  const int total = p.tiles_w * p.tiles_h * p.N;
  const int grid = total < *sms ? total : *sms;
  PMD_CHECK_CUDA(launch_pdl(conv3x3_resident_kernel<B_MN, RED>, dim3(grid), dim3(kResThreads), kResSmem, stream, p, total));
  return 0;
}

int launch_conv_resident(const ConvResidentParams& p_in, cudaStream_t stream) {
  ConvResidentParams p_traced;
  if (g_trace_buf != nullptr) {
    p_traced = p_in;
    p_traced.trace = g_trace_buf;
    p_traced.trace_cap = g_trace_cap;
  }
  const ConvResidentParams& p = g_trace_buf != nullptr ? p_traced : p_in;
  PMD_REQUIRE(p.red_y == nullptr || p.b_mn, "the fused BatchNorm-backward reduction belongs to the dgrad form");
  const int rc = p.red_y != nullptr ? launch_conv_resident_t<true, true>(p, stream)
                 : p.b_mn         ? launch_conv_resident_t<true>(p, stream)
                                  : launch_conv_resident_t<false>(p, stream);
  if (rc != 0) return -1;
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

// PMD_CONV_RESIDENT: 5 (default) fprop and every 64-channel dgrad, the dgrads with their fused BatchNorm-backward
// reduction / ReLU-bit addend; 2 fprop only; 1 fprop and dgrad without the fused reduction; 3 dgrad only; 4 fprop + a
// block's first-convolution dgrad; 0 off.  Without the fused reduction the resident dgrads lose (an extra reduce pass
// per layer: 1.479 / 1.471 ms/step for modes 1 / 4 against 1.461 for mode 2); with it they win (1.433).  Leaving SMs to
// the weight-gradient stream (fewer CTAs) does not help (1.438 / 1.443 at 140 / 128 CTAs).
static int conv_resident_mode() {
  return getenv("PMD_CONV_RESIDENT") == nullptr ? 5 : atoi(getenv("PMD_CONV_RESIDENT"));   // read per call (plan builders, tests)
}
bool conv_resident_enabled() { return conv_resident_mode() == 1 || conv_resident_mode() == 2 || conv_resident_mode() >= 4; }
bool conv_resident_dgrad_enabled() { return conv_resident_mode() == 1 || conv_resident_mode() == 3 || conv_resident_mode() == 5; }
bool conv_resident_red_enabled() { return conv_resident_mode() == 5; }
// mode 4: fprop + only the dgrads that carry no BatchNorm-backward reduction (a block's first convolution)
bool conv_resident_block_dgrad_enabled() { return conv_resident_dgrad_enabled() || conv_resident_mode() == 4; }

template <int BN, int ST, bool B_MN = false, int RED = 0, bool TWO = false>
static int launch_conv_gemm_t(const ConvGemmParams& p, const dim3& grid, cudaStream_t stream) {
  constexpr int smem =
      ST * (kABytes + (TWO ? BN / 2 : BN) * 128) + sizeof(PipeBars<ST>) + 16 + (RED ? sizeof(RedExtra<BN>) : 0) + 1024;
  static PerDevice<bool> ready;
  bool* once = ready.get();
  if (once == nullptr) return -1;
  if (!*once) {
    if (set_smem(conv_gemm_kernel<BN, ST, B_MN, RED, TWO>, smem)) return -1;
    *once = true;
  }
  PMD_CHECK_CUDA(launch_pdl_cluster(conv_gemm_kernel<BN, ST, B_MN, RED, TWO>, grid, dim3(kThreads), smem, stream, TWO ? 2 : 1, p));
  return 0;
}

// Pipeline depth trades smem per CTA against resident CTAs per SM (the epilogue of one CTA overlaps the
// main loop of its neighbours).  Tunable for experiments: PMD_GEMM64_STAGES / PMD_GEMM128_STAGES in {2,3,4}.
static int stages_from_env(const char* name, int dflt) {
  const char* v = getenv(name);
  if (v == nullptr) return dflt;
  const int s = atoi(v);
  return (s >= 2 && s <= 4) ? s : dflt;
}

// PMD_DGRAD_BN_REDUCE=0 keeps the stand-alone reduction kernels.
int attach_bn_reduce(ConvGemmParams* p, const PmdConvShape& s, const void* y, const float* scale, const float* shift,
                     const float* mean, const float* invstd, float* sums) {
  static const bool on = getenv("PMD_DGRAD_BN_REDUCE") == nullptr || atoi(getenv("PMD_DGRAD_BN_REDUCE")) != 0;
  if (!on || !p->b_mn_major || p->n_classes != 1 || s.stride != 1) return 0;   // (an addend is summed in before the mask)
  if (p->out_channels % 256 == 0 && p->bn_tile == 256) {
    // the y tile does not fit beside a 256-wide output staging area: fall back to 128-wide tiles for this dgrad
    p->bn_tile = 128;
  }
  if (p->bn_tile != 64 && p->bn_tile != 128) return 0;
  for (int i = 0; i < p->kb_count[0]; ++i)
    if (p->kb[p->kb_begin[0] + i].a_map != 0) return 0;    // fused shortcut branch: two A tensors, not handled
  if (act_map(&p->ymap[0], y, s.N, s.H, s.W, s.C, 1, 0, 0, p->w_box, p->h_box, p->n_box)) return -1;
  p->red_scale = scale;
  p->red_shift = shift;
  p->red_mean = mean;
  p->red_invstd = invstd;
  p->red_sums = sums;
  p->bn_reduce = 1;
  return 1;
}

int attach_block_bn_reduce(ConvGemmParams* p, const PmdConvShape& s, const unsigned char* relu_bits, int nb, const void* y0,
                           const float* mean0, const float* invstd0, float* sums0, const void* y1, const float* mean1,
                           const float* invstd1, float* sums1) {
  static const bool on = getenv("PMD_DGRAD_BN_REDUCE") == nullptr || atoi(getenv("PMD_DGRAD_BN_REDUCE")) != 0;
  if (!on || !p->b_mn_major || relu_bits == nullptr || (nb != 1 && nb != 2) || s.C % 64 != 0) return 0;
  if (p->out_channels % 256 == 0 && p->bn_tile == 256) p->bn_tile = 128;   // staging + y tile(s) must fit the ring
  if (p->bn_tile != 64 && p->bn_tile != 128) return 0;
  p->two_cta = 0;
  const int st = s.stride;
  for (int ph = 0; ph < st; ++ph)
    for (int pw = 0; pw < st; ++pw) {
      if (act_map(&p->ymap[ph * st + pw], y0, s.N, s.H, s.W, s.C, st, ph, pw, p->w_box, p->h_box, p->n_box)) return -1;
      if (nb == 2 && act_map(&p->ymap2[ph * st + pw], y1, s.N, s.H, s.W, s.C, st, ph, pw, p->w_box, p->h_box, p->n_box)) return -1;
    }
  p->red_nb = nb;
  p->red_stride = st;
  p->red_H = s.H;
  p->red_W = s.W;
  p->red_bits = relu_bits;
  p->red_mean = mean0;
  p->red_invstd = invstd0;
  p->red_sums = sums0;
  p->red_mean2 = mean1;
  p->red_invstd2 = invstd1;
  p->red_sums2 = sums1;
  p->red_scale = p->red_shift = nullptr;
  p->bn_reduce = 2;
  return 1;
}

int launch_conv_gemm(const ConvGemmParams& p_in, cudaStream_t stream) {
  ConvGemmParams p_traced;
  if (g_trace_buf != nullptr) {                 // tracing is a tool mode: the copy costs nothing when it is off
    p_traced = p_in;
    p_traced.trace = g_trace_buf;
    p_traced.trace_cap = g_trace_cap;
  }
  const ConvGemmParams& p = g_trace_buf != nullptr ? p_traced : p_in;
  const dim3 grid(p.tiles_w * p.tiles_h * p.tiles_n, p.out_channels / p.bn_tile, p.n_classes);
  // One tile per CTA with 2-4 resident CTAs per SM (a persistent double-buffered variant measured slower on these
  // shapes, profiles/r1_conv_variants.md, and was removed).  PMD_GEMM64_STAGES / PMD_GEMM128_STAGES in {2,3,4}.
  static const int st64 = stages_from_env("PMD_GEMM64_STAGES", 2), st128 = stages_from_env("PMD_GEMM128_STAGES", 3);
  int rc;
  if (p.two_cta) {           // CTA pairs: 128- / 256-wide tiles, pixel tiles in pairs along grid.x
    PMD_REQUIRE(grid.x % 2 == 0 && p.bn_tile >= 128, "internal: CTA-pair kernel needs an even pixel-tile count");
    // a pair stage is 8 KB (BN = 128) / 16 KB (BN = 256) smaller than a single-CTA stage: room for a deeper ring
    const int s128 = stages_from_env("PMD_GEMM2_STAGES128", 3), s256 = stages_from_env("PMD_GEMM2_STAGES256", 2);
    if (p.b_mn_major && p.bn_reduce)
      rc = s128 == 4 ? launch_conv_gemm_t<128, 4, true, 1, true>(p, grid, stream)
                     : launch_conv_gemm_t<128, 3, true, 1, true>(p, grid, stream);
    else if (p.b_mn_major)
      rc = p.bn_tile == 128 ? (s128 == 4 ? launch_conv_gemm_t<128, 4, true, 0, true>(p, grid, stream)
                                         : launch_conv_gemm_t<128, 3, true, 0, true>(p, grid, stream))
                            : (s256 == 3 ? launch_conv_gemm_t<256, 3, true, 0, true>(p, grid, stream)
                                         : launch_conv_gemm_t<256, 2, true, 0, true>(p, grid, stream));
    else
      rc = p.bn_tile == 128 ? (s128 == 4 ? launch_conv_gemm_t<128, 4, false, 0, true>(p, grid, stream)
                                         : launch_conv_gemm_t<128, 3, false, 0, true>(p, grid, stream))
                            : (s256 == 3 ? launch_conv_gemm_t<256, 3, false, 0, true>(p, grid, stream)
                                         : launch_conv_gemm_t<256, 2, false, 0, true>(p, grid, stream));
  } else if (p.b_mn_major && p.bn_reduce == 2) {
    rc = p.bn_tile == 64 ? launch_conv_gemm_t<64, 2, true, 2>(p, grid, stream)
                         : launch_conv_gemm_t<128, 3, true, 2>(p, grid, stream);
  } else if (p.b_mn_major && p.bn_reduce) {
    rc = p.bn_tile == 64 ? launch_conv_gemm_t<64, 2, true, 1>(p, grid, stream)
                         : launch_conv_gemm_t<128, 3, true, 1>(p, grid, stream);
  } else if (p.b_mn_major) {
    rc = p.bn_tile == 64 ? launch_conv_gemm_t<64, 2, true>(p, grid, stream)
                         : (p.bn_tile == 128 ? launch_conv_gemm_t<128, 3, true>(p, grid, stream)
                                             : launch_conv_gemm_t<256, 2, true>(p, grid, stream));
  } else if (p.bn_tile == 256) {
    rc = launch_conv_gemm_t<256, 2>(p, grid, stream);
  } else if (p.bn_tile == 64) {
    rc = st64 == 2 ? launch_conv_gemm_t<64, 2>(p, grid, stream)
                   : (st64 == 3 ? launch_conv_gemm_t<64, 3>(p, grid, stream) : launch_conv_gemm_t<64, 4>(p, grid, stream));
  } else {
    rc = st128 == 2 ? launch_conv_gemm_t<128, 2>(p, grid, stream)
                    : (st128 == 4 ? launch_conv_gemm_t<128, 4>(p, grid, stream) : launch_conv_gemm_t<128, 3>(p, grid, stream));
  }
  if (rc != 0) return -1;
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

template <int BNW, int ST, int MB = 1>
static int launch_wgrad_t(const WgradParams& p, cudaStream_t stream) {
  constexpr int smem = ST * (2 * MB * 8192 + (BNW / 64) * 8192) + sizeof(PipeBars<ST>) + 1024;
  static_assert(smem <= 227 * 1024, "wgrad pipeline does not fit in shared memory");
  static PerDevice<bool> ready;
  bool* once = ready.get();
  if (once == nullptr) return -1;
  if (!*once) {
    if (set_smem(conv_wgrad_kernel<BNW, ST, MB>, smem)) return -1;
    *once = true;
  }
  const dim3 grid(ceil_div(p.n_boxes, 2 * MB), p.K / BNW, p.split);
  PMD_CHECK_CUDA(launch_pdl(conv_wgrad_kernel<BNW, ST, MB>, grid, dim3(kThreads), smem, stream, p));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

static int launch_wgrad_patch(const WgradParams& p, cudaStream_t stream) {
  constexpr int ST = 4;
  constexpr int smem = ST * kWPatchStage + sizeof(PipeBars<ST>) + 1024;
  static PerDevice<bool> ready;
  bool* once = ready.get();
  if (once == nullptr) return -1;
  if (!*once) {
    if (set_smem(conv_wgrad_patch_kernel<ST>, smem)) return -1;
    *once = true;
  }
  const dim3 grid(p.C / 64, p.K / 64, p.split);
  PMD_CHECK_CUDA(launch_pdl(conv_wgrad_patch_kernel<ST>, grid, dim3(kThreads), smem, stream, p));
  count_launch();
  PMD_LAUNCH_CHECK();
  return 0;
}

int launch_wgrad(const WgradParams& p, cudaStream_t stream) {
  if (p.patch) return launch_wgrad_patch(p, stream);
  if (p.mb == 2) {
    if (p.bn_tile == 256) return launch_wgrad_t<256, 3, 2>(p, stream);
    if (p.bn_tile == 128) return launch_wgrad_t<128, 4, 2>(p, stream);
    return launch_wgrad_t<64, 4, 2>(p, stream);
  }
  if (p.bn_tile == 256) return launch_wgrad_t<256, 4>(p, stream);
  if (p.bn_tile == 128) return launch_wgrad_t<128, 4>(p, stream);
  return launch_wgrad_t<64, 4>(p, stream);
}

}  // namespace pmd

// ============================================================================ C ABI
using namespace pmd;

extern "C" PMD_API int pmd_conv_fprop(const PmdConvShape* s, const void* d_x, const void* d_w_krsc, void* d_y, float* d_stats,
                              void* stream) {
  if (conv_resident_enabled() && conv_resident_supported(*s)) {
    ConvResidentParams hp;
    if (build_fprop_resident_params(*s, d_x, d_w_krsc, d_y, d_stats, &hp)) return -1;
    return launch_conv_resident(hp, static_cast<cudaStream_t>(stream));
  }
  ConvGemmParams p;
  if (build_fprop_params(*s, d_x, d_w_krsc, d_y, d_stats, &p)) return -1;
  return launch_conv_gemm(p, static_cast<cudaStream_t>(stream));
}

extern "C" PMD_API int pmd_conv_set_trace(void* d_buf, int cap_ctas) {
  g_trace_buf = static_cast<long long*>(d_buf);
  g_trace_cap = d_buf != nullptr ? cap_ctas : 0;
  return 0;
}

extern "C" PMD_API int pmd_conv_dgrad(const PmdConvShape* s, const void* d_dy, const void* d_w_rsck, const PmdConvShape* s2,
                              const void* d_dy2, const void* d_w2_rsck, const void* d_addend, void* d_dx,
                              void* stream) {
  ConvGemmParams p;
  if (build_dgrad_params(*s, d_dy, d_w_rsck, s2, d_dy2, d_w2_rsck, d_addend, d_dx, &p)) return -1;
  return launch_conv_gemm(p, static_cast<cudaStream_t>(stream));
}

extern "C" PMD_API int pmd_conv_dgrad_krsc(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc,
                                           const PmdConvShape* s2, const void* d_dy2, const void* d_w2_krsc,
                                           const void* d_addend, void* d_dx, void* stream) {
  if (s2 == nullptr && conv_resident_dgrad_enabled() && conv_resident_supported(*s)) {
    ConvResidentParams hp;
    if (build_dgrad_resident_params(*s, d_dy, d_w_krsc, d_addend, d_dx, &hp)) return -1;
    return launch_conv_resident(hp, static_cast<cudaStream_t>(stream));
  }
  ConvGemmParams p;
  if (build_dgrad_params(*s, d_dy, d_w_krsc, s2, d_dy2, d_w2_krsc, d_addend, d_dx, &p, true)) return -1;
  return launch_conv_gemm(p, static_cast<cudaStream_t>(stream));
}

extern "C" PMD_API int pmd_conv_dgrad_bn_reduce_addend(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc,
                                                       const void* d_addend, const void* d_y_prev, const float* d_scale,
                                                       const float* d_shift, const float* d_mean, const float* d_invstd,
                                                       float* d_sums, void* d_dz, void* stream) {
  if (conv_resident_red_enabled() && conv_resident_supported(*s)) {
    ConvResidentParams hp;
    if (build_dgrad_resident_params(*s, d_dy, d_w_krsc, d_addend, d_dz, &hp)) return -1;
    attach_bn_reduce_resident(&hp, d_y_prev, d_scale, d_shift, d_mean, d_invstd, d_sums);
    return launch_conv_resident(hp, static_cast<cudaStream_t>(stream));
  }
  ConvGemmParams p;
  if (build_dgrad_params(*s, d_dy, d_w_krsc, nullptr, nullptr, nullptr, d_addend, d_dz, &p, true)) return -1;
  const int rc = attach_bn_reduce(&p, *s, d_y_prev, d_scale, d_shift, d_mean, d_invstd, d_sums);
  if (rc < 0) return -1;
  PMD_REQUIRE(rc == 1, "this dgrad cannot carry the BatchNorm reduction (needs stride 1, C % 64 == 0, and the fusion enabled)");
  return launch_conv_gemm(p, static_cast<cudaStream_t>(stream));
}

extern "C" PMD_API int pmd_conv_dgrad_bn_reduce(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, const void* d_y_prev,
                                                const float* d_scale, const float* d_shift, const float* d_mean,
                                                const float* d_invstd, float* d_sums, void* d_dz, void* stream) {
  return pmd_conv_dgrad_bn_reduce_addend(s, d_dy, d_w_krsc, nullptr, d_y_prev, d_scale, d_shift, d_mean, d_invstd, d_sums, d_dz,
                                         stream);
}

extern "C" PMD_API int pmd_conv_dgrad_block_bn_reduce(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc,
                                                      const PmdConvShape* s2, const void* d_dy2, const void* d_w2_krsc,
                                                      const void* d_addend, const unsigned char* d_relu_bits, int nb,
                                                      const void* d_y0, const float* d_mean0, const float* d_invstd0,
                                                      float* d_sums0, const void* d_y1, const float* d_mean1,
                                                      const float* d_invstd1, float* d_sums1, void* d_dx, void* stream) {
  ConvGemmParams p;
  if (build_dgrad_params(*s, d_dy, d_w_krsc, s2, d_dy2, d_w2_krsc, d_addend, d_dx, &p, true)) return -1;
  const int rc = attach_block_bn_reduce(&p, *s, d_relu_bits, nb, d_y0, d_mean0, d_invstd0, d_sums0, d_y1, d_mean1, d_invstd1,
                                        d_sums1);
  if (rc < 0) return -1;
  PMD_REQUIRE(rc == 1, "this dgrad cannot carry the block-output BatchNorm reduction");
  return launch_conv_gemm(p, static_cast<cudaStream_t>(stream));
}

extern "C" PMD_API int pmd_conv_wgrad(const PmdConvShape* s, const void* d_x, const void* d_dy, float* d_dw_krsc, int split_k,
                              void* stream) {
  WgradParams p;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (build_wgrad_params(*s, d_x, d_dy, d_dw_krsc, split_k, sms, &p)) return -1;
  return launch_wgrad(p, static_cast<cudaStream_t>(stream));
}
