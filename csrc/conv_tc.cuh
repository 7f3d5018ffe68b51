// Internal interface of the tensor-core convolution kernels (conv_tc.cu): parameter blocks that
// are built once per layer (tensor maps + k-block tables) and launched every step by the engine.
#pragma once

#include "common.cuh"
#include "../include/pmdist.h"

namespace pmd {

constexpr int kMaxKBlocks = 96;
constexpr int kMaxWBoxes = 96;

// One pipeline stage of the fprop/dgrad implicit GEMM: which activation window and which weight
// slice to fetch.  (dh, dw) shift the 128-pixel box inside activation map `a_map`; zero padding
// is TMA out-of-bounds fill.
struct KBlock {
  int8_t dh, dw;
  uint8_t a_map, b_map;
  uint16_t a_ch;   // first channel of the 64-wide chunk in the activation tensor
  uint16_t b_c0;   // first reduction-dim element of the 64-wide chunk in the weight tensor
  uint8_t tap;     // r*S + s
  uint8_t pad_[3];
};

struct ConvGemmParams {
  int l2_hint;               // L2 eviction hints of the operand loads (set at launch from PMD_L2_HINT, see conv_l2_hint_mask)
  CUtensorMap a[4];
  CUtensorMap b[2];
  CUtensorMap out[4];
  KBlock kb[kMaxKBlocks];
  int kb_begin[4], kb_count[4];
  int GW[4], GH[4];          // valid pixel grid per class (blockIdx.z)
  int N;
  int w_box, h_box, n_box, tiles_w, tiles_h, tiles_n;
  int n_classes;
  int b_tap_last;            // weight-map coordinate order: 0 = (c, tap, n) [KRSC], 1 = (c, n, tap) [RSCK]
  int b_mn_major;            // dgrad straight from KRSC weights: B tile = [64 k][64 c] boxes, MN-major descriptors
  int out_channels;          // total channels of the output tensor
  int bn_tile;               // 64, 128 or 256
  int two_cta;               // CTA-pair kernel (tcgen05 cta_group::2): the weight map's box holds bn_tile / 2 rows
  float* stats;              // optional [2 * out_channels]: per channel sum(y - pivot), sum((y - pivot)^2)
  // Optional [out_channels] pivot of the statistics (null = 0).  E[y^2] - E[y]^2 in fp32 cancels when |mean| >> std;
  // sums taken around a pivot near the mean do not.  The engine passes the previous step's batch mean.
  const float* stat_pivot;
  const __nv_bfloat16* addend;  // optional NHWC tensor shaped like the (single-class) output
  // Optional ReLU bit array of the addend (uint8 [pixels * out_channels / 8], pmd_bn_apply_bits): elements whose bit is
  // clear count as zero.  An identity shortcut's gradient is dz_out masked by the block output's ReLU: with the bits the
  // dgrad reads dz_out directly and the BatchNorm backward need not write the masked copy.
  const unsigned char* addend_bits;
  long long* trace;          // optional [trace_cap][kTraceSlots] per-CTA timestamps (pmd_conv_set_trace), else null
  int trace_cap;
  // dgrad + BatchNorm-backward reduction in one kernel (attach_bn_reduce): the output dz = dL/dz of the previous
  // layer's activation z = relu(BN(y)).  The epilogue loads the y tile, applies the ReLU mask to dz before storing
  // it and accumulates per channel  sum(g), sum(g * xhat)  into red_sums -- exactly what the stand-alone
  // bn_bwd_reduce kernel would compute from a second pass over dz and y.
  int bn_reduce;                         // 0 off, 1 inner BatchNorm (mask from y), 2 block-output BatchNorm(s) (mask from bits)
  CUtensorMap ymap[4];                   // y of that BatchNorm, same shape and box as the output (mode 2: one parity view per class)
  const float *red_scale, *red_shift;    // forward affine (mask = scale*y + shift > 0)
  const float *red_mean, *red_invstd;
  float* red_sums;                       // [2 * out_channels], accumulated with red.global.add
  // mode 2 (attach_block_bn_reduce): the gradient this dgrad produces is dL/dz of a residual block's output
  // z = relu(BN(y) + BN_sc(y_sc) | + identity); its mask is the ReLU bit array of z, and up to two BatchNorms reduce it
  int red_nb;                            // 1 or 2 branches
  int red_stride, red_H, red_W;          // output (= z) geometry: class (ph, pw) pixel (h, w) is z pixel (stride*h + ph, stride*w + pw)
  const unsigned char* red_bits;         // uint8 [N * red_H * red_W * out_channels / 8]
  CUtensorMap ymap2[4];                  // the shortcut convolution's y (second branch)
  const float *red_mean2, *red_invstd2;
  float* red_sums2;
};

// Per-CTA trace record written by the fprop/dgrad kernels when tracing is on (tools/conv_trace.py decodes it).
// Slots hold clock64() of the CTA's SM unless noted.
enum {
  kTrEntry = 0,      // kernel entry
  kTrSetup = 1,      // barriers initialised, TMEM allocated, CTA-wide sync passed
  kTrFirstLoad = 2,  // producer: first stage's TMA issued
  kTrFirstFull = 3,  // MMA thread: first stage has landed
  kTrMmaIssued = 4,  // MMA thread: last MMA + commit issued (first tile for the persistent kernel)
  kTrAccReady = 5,   // epilogue: accumulator complete (tmem_full observed)
  kTrStaged = 6,     // epilogue: tile converted and staged in smem, store about to be issued
  kTrEpiDone = 7,    // epilogue: statistics reduced, store has read the staging buffer
  kTrExit = 8,       // after the final CTA-wide sync
  kTrGlobalNs = 9,   // %globaltimer at entry (ns, comparable across SMs)
  kTrSm = 10,        // %smid
  kTrTiles = 11,     // tiles processed by this CTA
  kTraceSlots = 12
};

struct WBox {
  int8_t dh, dw;
  uint8_t a_map, pad0_;
  uint16_t a_ch, pad1_;
  int32_t out_col;           // tap * C + first channel: column of dW's [K][R*S*C] matrix
};

struct WgradParams {
  int l2_hint;
  CUtensorMap x[4];
  CUtensorMap dy;
  WBox box[kMaxWBoxes];
  int n_boxes;
  int w_box, h_box, n_box, tiles_w, tiles_h, tiles_n;
  int K, RSC;
  int bn_tile;               // 64, 128 or 256 output-channel columns per CTA
  int split;
  int c_valid;                 // rows (channels within a box) >= c_valid are not written (stem: 27 of 64)
  int mb;                      // 128-row accumulators per CTA: 1 (two x boxes) or 2 (four)
  int patch;                   // 3x3 stride-1, 64 output channels per CTA: conv_wgrad_patch_kernel (all nine taps per CTA)
  int C;
  float* dw;
};

inline int conv_out_dim(int in, int R, int stride) { return (in + 2 * (R / 2) - R) / stride + 1; }
inline bool conv_tc_supported(const PmdConvShape& s) {
  return s.C % 64 == 0 && s.K % 64 == 0 && (s.R == 1 || s.R == 3) && (s.stride == 1 || s.stride == 2);
}

// 3x3 stride-1 convolutions with C = K = 64 on maps with W % 8 == 0 and H % 16 == 0 (the 32x32 stage):
// conv3x3_resident_kernel -- persistent, weights resident in shared memory, activations as three column patches
// per tile (L2->SM activation traffic 2.7x lower than per-tap boxes).
struct ConvResidentParams {
  int l2_hint;
  CUtensorMap a;       // input, box (64 ch, 8, 18, 1): one column patch
  CUtensorMap b;       // KRSC weights (c, tap, k), box 64 x 1 x 64
  CUtensorMap out;     // output, box (64 ch, 8, 16, 1)
  int tiles_w, tiles_h, N, GW, GH;
  int flip_taps;       // dgrad: weight tap (2-r, 2-s) pairs with patch offset (r, s)
  int b_mn;            // dgrad straight from the KRSC weights (MN-major B operand)
  int out_channels;    // 64
  float* stats;
  const float* stat_pivot;   // as in ConvGemmParams
  const __nv_bfloat16* addend;
  const unsigned char* addend_bits;   // optional ReLU bits of the addend (as ConvGemmParams::addend_bits)
  // Fused BatchNorm-backward reduction of the layer below (dgrad only; as RED == 1 of conv_gemm_kernel): the output is
  // masked with (red_scale * y + red_shift > 0) and red_sums[c] += sum(g), red_sums[64 + c] += sum(g * xhat).
  const __nv_bfloat16* red_y;
  const float *red_scale, *red_shift, *red_mean, *red_invstd;
  float* red_sums;
  long long* trace;    // optional [trace_cap][kTraceSlots] per-CTA record (kRt* slots), else null
  int trace_cap;
};
// Trace record of conv3x3_resident_kernel: total cycles each role spent WAITING over all tiles of the CTA.
enum {
  kRtEntry = 0, kRtSetup = 1,
  kRtProdWaitEmpty = 2,   // producer: waiting for a free patch slot (consumer-bound when large)
  kRtMmaWaitFull = 3,     // MMA thread: waiting for a patch to land (load-bound when large and the above is small)
  kRtMmaWaitAcc = 4,      // MMA thread: waiting for the epilogue to release an accumulator
  kRtEpiWaitAcc = 5,      // epilogue group 0: waiting for an accumulator to complete
  kRtEpiBusy = 6,         // epilogue group 0: tcgen05.ld .. statistics, excluding the waits
  kRtEpiWaitStore = 7,    // epilogue group 0: waiting for the previous TMA store to release the staging buffer
  kRtExit = 8             // then kTrGlobalNs, kTrSm, kTrTiles as in the per-tile kernels
};
inline bool conv_resident_supported(const PmdConvShape& s) {
  return conv_tc_supported(s) && s.R == 3 && s.stride == 1 && s.W % 8 == 0 && s.H % 16 == 0 && s.C == 64 && s.K == 64;
}
// PMD_L2_HINT bit mask: operand loads issued with an L2 evict-first policy (the tensor is not read again soon, so it should
// not push out the one being written for the next kernel): 1 resident fprop activations, 2 per-tap fprop activations,
// 4 dgrad dy, 8 wgrad x, 16 wgrad dy, 32 the y cells of the resident dgrad reduction.  Default 27 (1 + 2 + 8 + 16): measured on
// one box in one run, ms/step: 0 -> 1.4531, 19 -> 1.4437, 19 + BatchNorm streaming loads (csrc/bn.cu) -> 1.4155; on another
// box 19 -> 1.4252, 27 -> 1.4185, 51 -> 1.4271, 59 -> 1.4205; bit 4 (dy is read by the dgrad AND the wgrad) costs ~10 us.
int conv_l2_hint_mask();
bool conv_resident_enabled();        // fprop
bool conv_resident_dgrad_enabled();
bool conv_resident_red_enabled();           // mode 5: every 64-channel dgrad resident, with the fused reductions
bool conv_resident_block_dgrad_enabled();   // a block's first-convolution dgrad (modes 1, 3, 4, 5; see conv_resident_mode in conv_tc.cu)
int build_fprop_resident_params(const PmdConvShape& s, const void* x, const void* w_krsc, void* y, float* stats,
                                ConvResidentParams* out);
int build_dgrad_resident_params(const PmdConvShape& s, const void* dy, const void* w_krsc, const void* addend, void* dx,
                                ConvResidentParams* out);
// The layer below's BatchNorm-backward reduction in a resident dgrad's epilogue (see ConvResidentParams::red_y).
void attach_bn_reduce_resident(ConvResidentParams* p, const void* y, const float* scale, const float* shift,
                               const float* mean, const float* invstd, float* sums);
int launch_conv_resident(const ConvResidentParams& p, cudaStream_t stream);

// Builders return 0 / -1 (error text via pmd_last_error()).
int build_fprop_params(const PmdConvShape& s, const void* x, const void* w_krsc, void* y, float* stats,
                       ConvGemmParams* out);
int build_dgrad_params(const PmdConvShape& s, const void* dy, const void* w_rsck, const PmdConvShape* s2,
                       const void* dy2, const void* w2_rsck, const void* addend, void* dx, ConvGemmParams* out,
                       bool w_is_krsc = false);
int build_wgrad_params(const PmdConvShape& s, const void* x, const void* dy, float* dw, int split, int sm_count,
                       WgradParams* out, int rsc_override = 0, int c_valid = 64);
// Fuses the BatchNorm-backward reduction of the layer below into a stride-1, single-branch dgrad (see ConvGemmParams).
// Returns 1 if attached, 0 if this dgrad cannot take it (the caller then runs the stand-alone reduce), -1 on error.
int attach_bn_reduce(ConvGemmParams* p, const PmdConvShape& s, const void* y, const float* scale, const float* shift,
                     const float* mean, const float* invstd, float* sums);
// The same for the BatchNorm(s) that produced a residual block's OUTPUT, fused into the dgrad of the NEXT block's first
// convolution (any stride, with or without a fused projection-shortcut dgrad / addend).  nb = 1: y1 etc. null.
int attach_block_bn_reduce(ConvGemmParams* p, const PmdConvShape& s, const unsigned char* relu_bits, int nb, const void* y0,
                           const float* mean0, const float* invstd0, float* sums0, const void* y1, const float* mean1,
                           const float* invstd1, float* sums1);
int launch_conv_gemm(const ConvGemmParams& p, cudaStream_t stream);
int launch_wgrad(const WgradParams& p, cudaStream_t stream);

// Tiled bf16 tensor map with 128-byte swizzle (the encoder behind every map in this file), for kernels in other files.
int encode_tiled_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box);

// stem_fused_kernel (csrc/stem_fused.cu): gather + crop/flip + normalise + im2col + the 3->64 stem convolution + its
// BatchNorm statistics in ONE persistent kernel.  Builder warps write the 27-column patches of 128 consecutive pixels
// straight into the swizzled shared-memory A tile (and, for the weight gradient, into x0); no im2col read-back.
struct StemFusedParams {
  CUtensorMap w;       // padded stem weight [64 k][64 c], box 64 x 64 (used when w_raw is null)
  const __nv_bfloat16* w_raw;   // or the KRSC stem weight [64][w_rsc] itself: the kernel pads it into shared memory
  int w_rsc;
  CUtensorMap out;     // y viewed as [M pixels][64], box 64 x 128
  CUtensorMap x0m;     // x0 viewed as [M pixels][32 of 64 columns] (or all 64), box 64 x 128: store of an A slot
  int x0_tma;          // x0m is valid: x0 leaves through TMA; 0 = per-thread stores
  // batch source: resident dataset (desc/hdr/idx, see pmd_augment_im2col_gather) or staged uint8 images + crop/flip
  const unsigned long long* ds;
  const int* hdr;
  const int* idx;
  const unsigned char* u8;
  const int* aug;
  long long* labels_out;
  unsigned char* valid_out;
  int N, H, W, pad;
  __nv_bfloat16* x0;         // optional [M][64]: columns 0..31 written (the stem wgrad's operand); null = not needed
  float* stats;              // optional [128]: per channel sum(y - pivot), sum((y - pivot)^2)
  const float* stat_pivot;   // optional [64]
};
inline bool stem_fused_supported(int N, int H, int W) {
  // W % 4: image rows are copied as 32-bit words; W <= 336: the (at most four, for W >= 128) stored rows one 128-pixel
  // tile reads fit its 4 KB stage
  return (H * W) % 128 == 0 && W % 4 == 0 && W <= 336 && static_cast<long long>(N) * H * W < (1LL << 31);
}
bool stem_fused_enabled();   // PMD_STEM_FUSED (default 1)
// x0 (may be null) = the [M][64] im2col buffer a training launch fills; a launch uses it only if its p.x0 is set.
int build_stem_fused_params(int N, int H, int W, int pad, const void* w_pad, void* x0, void* y, StemFusedParams* out);
int launch_stem_fused(const StemFusedParams& p, cudaStream_t stream);

}  // namespace pmd
