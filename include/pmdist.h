/* libpmdist -- C ABI of the B200-native (sm_100a) data-parallel training runtime.
 *
 * Every entry point returns 0 on success and -1 on failure (pmd_last_error() has the text).
 * Pointers named d_* are device pointers; `stream` is a cudaStream_t passed as void*.
 * Activations are NHWC bf16; convolution weights are KRSC (fprop/wgrad) or RSCK (dgrad) bf16;
 * master parameters, gradients and momentum are fp32.
 *
 * The Python package binds this header with ctypes (pmdist/ops/_lib.py).  "ref:" citations name
 * the place in the reference (MOONJOOYOUNG/pytorch_multiprocessing-distributed) whose behaviour
 * the entry point replaces; the reference itself ships no native code.
 */
#ifndef PMDIST_H_
#define PMDIST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* pmd_last_error(void);
/* Number of kernels this library has launched in this process (bench.py "gpu_launches"). */
unsigned long long pmd_launch_count(void);
int pmd_version(void);

/* ------------------------------------------------------------------ convolution (conv_tc.cu)
 * Implicit-GEMM convolution on tcgen05 tensor cores with TMEM accumulators, operands staged by
 * TMA.  Replaces cuDNN fprop/dgrad/wgrad under nn.Conv2d (ref:model/resnet.py:20-31, 48-61, 79;
 * SURVEY.md K1-K3).  R == S in {1, 3}, pad = R/2, stride in {1, 2}; C and K multiples of 64. */
typedef struct {
  int N, H, W, C; /* input  [N, H, W, C] */
  int K;          /* output [N, Ho, Wo, K], Ho = (H + 2*(R/2) - R)/stride + 1 */
  int R;          /* square filter size: 1 or 3 */
  int stride;     /* 1 or 2 */
} PmdConvShape;

/* y = conv(x, w).  If d_stats != NULL, the epilogue also accumulates per-output-channel
 * sum(y) into d_stats[0..K) and sum(y*y) into d_stats[K..2K) (of the bf16-rounded outputs): the
 * BatchNorm batch statistics at no extra pass.  d_stats must be zeroed by the caller. */
int pmd_conv_fprop(const PmdConvShape* s, const void* d_x, const void* d_w_krsc, void* d_y, float* d_stats,
                   void* stream);
/* dx = conv_transpose(dy, w) [+ conv_transpose(dy2, w2)] [+ addend].  The optional second
 * convolution (s2 != NULL) must consume the same input tensor (a residual block's projection
 * shortcut); it is accumulated in the same TMEM tile, so the branch sum costs no extra pass.
 * d_addend (may be NULL) is an NHWC bf16 tensor shaped like dx (identity-shortcut gradient). */
int pmd_conv_dgrad(const PmdConvShape* s, const void* d_dy, const void* d_w_rsck, const PmdConvShape* s2,
                   const void* d_dy2, const void* d_w2_rsck, const void* d_addend, void* d_dx, void* stream);
/* Same as pmd_conv_dgrad but reads the KRSC weights fprop uses (consumed MN-major by tcgen05): no transposed RSCK
 * copy of the weights is needed at all. */
int pmd_conv_dgrad_krsc(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, const PmdConvShape* s2,
                        const void* d_dy2, const void* d_w2_krsc, const void* d_addend, void* d_dx, void* stream);
/* dgrad (stride 1, KRSC weights) fused with the BatchNorm-backward reduction of the layer below, whose output
 * z = relu(scale*y_prev + shift) was this convolution's input: writes d_dz = mask(z > 0) * conv_transpose(dy, w) and
 * accumulates d_sums[0..C) += sum(dz), d_sums[C..2C) += sum(dz * (y_prev - mean) * invstd) over all pixels (d_sums is
 * an accumulator the caller zeroes).  Replaces pmd_conv_dgrad_krsc + pmd_bn_bwd_reduce_multi(mask_mode 2). */
int pmd_conv_dgrad_bn_reduce(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, const void* d_y_prev,
                             const float* d_scale, const float* d_shift, const float* d_mean, const float* d_invstd,
                             float* d_sums, void* d_dz, void* stream);
/* The same with an addend tensor (an identity shortcut's gradient, shaped like d_dz) summed in before the mask:
 * the first residual block's dgrad, whose output feeds only the stem's BatchNorm backward. */
int pmd_conv_dgrad_bn_reduce_addend(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, const void* d_addend,
                                    const void* d_y_prev, const float* d_scale, const float* d_shift, const float* d_mean,
                                    const float* d_invstd, float* d_sums, void* d_dz, void* stream);
/* dgrad of a residual block's FIRST convolution (optionally with the projection shortcut's dgrad fused as s2 / d_dy2 /
 * d_w2_krsc, or an identity shortcut's gradient as d_addend) that also does the BatchNorm-backward reduction of the
 * PREVIOUS block's output: d_dx = dL/dz_prev is masked with that output's ReLU bit array (pmd_bn_apply_bits) and, per
 * channel, sum(g) and sum(g * xhat) are accumulated for nb = 1 (main path) or 2 (main path + projection shortcut)
 * BatchNorms -- what pmd_bn_bwd_reduce_multi would compute from a second pass over d_dx, y0, y1 and z.  Any stride. */
int pmd_conv_dgrad_block_bn_reduce(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, const PmdConvShape* s2,
                                   const void* d_dy2, const void* d_w2_krsc, const void* d_addend,
                                   const unsigned char* d_relu_bits, int nb, const void* d_y0, const float* d_mean0,
                                   const float* d_invstd0, float* d_sums0, const void* d_y1, const float* d_mean1,
                                   const float* d_invstd1, float* d_sums1, void* d_dx, void* stream);
/* dw (fp32, KRSC) += x^T * dy, reduction over all N*Ho*Wo pixels split over `split_k` CTAs per
 * output tile (0 = choose); partial tiles are added with red.global.add.f32, so d_dw is an
 * accumulator the caller zeroes (it is a slice of the flat gradient bucket slab). */
int pmd_conv_wgrad(const PmdConvShape* s, const void* d_x, const void* d_dy, float* d_dw_krsc, int split_k,
                   void* stream);
/* Tool mode (tools/conv_trace.py): while d_buf is non-NULL every fprop/dgrad tensor-core launch writes, per CTA,
 * PMD_CONV_TRACE_SLOTS long-long timestamps (clock64 per pipeline milestone, %globaltimer, %smid, tile count; slot
 * meaning in csrc/conv_tc.cuh) for its first cap_ctas CTAs.  d_buf = NULL switches tracing off.  Not thread-safe. */
#define PMD_CONV_TRACE_SLOTS 12
int pmd_conv_set_trace(void* d_buf, int cap_ctas);

/* Direct (SIMT) convolution for shapes the tensor-core path does not take (the C=3 stem) and as
 * an in-library cross-check.  Same layouts; C, K arbitrary.  x may carry `x_cstride` >= C channels
 * per pixel (the augmentation kernel pads RGB to 4). */
int pmd_conv_fprop_direct(const PmdConvShape* s, int x_cstride, const void* d_x, const void* d_w_krsc, void* d_y,
                          void* stream);
int pmd_conv_dgrad_direct(const PmdConvShape* s, const void* d_dy, const void* d_w_krsc, void* d_dx, void* stream);
int pmd_conv_wgrad_direct(const PmdConvShape* s, int x_cstride, const void* d_x, const void* d_dy,
                          float* d_dw_krsc, void* stream);

/* ------------------------------------------------------------------ BatchNorm (bn.cu)
 * Replaces torch's SyncBatchNorm kernels (ref:main.py:43). */
/* per-channel sum / sum-of-squares of a bf16 [rows, C] tensor -> d_stats[2C] (accumulates). */
int pmd_bn_stats(const void* d_y, long long rows, int C, float* d_stats, void* stream);
/* The same sums taken around a per-channel pivot P (d_pivot[C], NULL = 0): d_stats[c] += sum(y - P),
 * d_stats[C + c] += sum((y - P)^2).  With P near the channel mean the variance  s2/n - (s1/n)^2  does not suffer the
 * fp32 cancellation of E[y^2] - E[y]^2 when |mean| >> std (torch's SyncBatchNorm avoids it with Welford merges).
 * pmd_bn_finalize_pivot consumes such sums (mean = P + s1/n) and can store the batch mean as the next pivot; the
 * convolution epilogues (pmd_conv_fprop through the engine) produce them the same way.  All ranks must use the same P. */
int pmd_bn_stats_pivot(const void* d_y, long long rows, int C, float* d_stats, const float* d_pivot, void* stream);
int pmd_bn_finalize_pivot(const float* const* d_stats_ranks, int world, float count_per_rank, int C,
                          const float* d_gamma, const float* d_beta, float* d_running_mean, float* d_running_var,
                          long long* d_num_batches, float momentum, float eps, float* d_scale, float* d_shift,
                          float* d_mean, float* d_invstd, const float* d_pivot, float* d_pivot_new, void* stream);
/* Combine statistics and produce the affine the apply kernel uses.
 *   d_stats_ranks: `world` pointers (device array of float*) to each rank's [2C] sums -- with
 *   world == 1 pass a 1-element table.  count_per_rank = rows each rank contributed.
 *   Writes scale[c] = gamma*invstd, shift[c] = beta - mean*scale, save_mean, save_invstd; updates
 *   running_mean/var (momentum, unbiased variance) and num_batches_tracked like torch. */
int pmd_bn_finalize(const float* const* d_stats_ranks, int world, float count_per_rank, int C, const float* d_gamma,
                    const float* d_beta, float* d_running_mean, float* d_running_var, long long* d_num_batches,
                    float momentum, float eps, float* d_scale, float* d_shift, float* d_mean, float* d_invstd,
                    void* stream);
/* eval mode: scale/shift from running statistics. */
int pmd_bn_eval_affine(int C, const float* d_gamma, const float* d_beta, const float* d_running_mean,
                       const float* d_running_var, float eps, float* d_scale, float* d_shift, void* stream);
/* z = act(scale*y + shift [+ (res_scale*res + res_shift | res)]); relu != 0 applies max(.,0). */
int pmd_bn_apply(const void* d_y, const float* d_scale, const float* d_shift, const void* d_res,
                 const float* d_res_scale, const float* d_res_shift, int relu, void* d_z, long long rows, int C,
                 void* stream);
/* The same, also recording WHICH outputs passed the ReLU: d_relu_bits = uint8 [rows * C / 8], bit j of byte i set iff
 * element 8 i + j was > 0 before clamping.  The backward kernels take this array as their mask (mask_mode 3) instead of
 * reading the 16x larger z tensor (threshold_backward fused into the BatchNorm backward). */
int pmd_bn_apply_bits(const void* d_y, const float* d_scale, const float* d_shift, const void* d_res,
                      const float* d_res_scale, const float* d_res_shift, int relu, void* d_z, void* d_relu_bits,
                      long long rows, int C, void* stream);
/* backward reduce: g = dz * (z > 0 if relu); d_sums[0..C) += sum(g), d_sums[C..2C) += sum(g * xhat),
 * xhat = (y - mean) * invstd. */
int pmd_bn_bwd_reduce(const void* d_dz, const void* d_z, const void* d_y, const float* d_mean, const float* d_invstd,
                      int relu, float* d_sums, long long rows, int C, void* stream);
/* backward elementwise: dy = gamma*invstd * (g - sum_g/Ntot - xhat * sum_gx/Ntot); sums come from
 * `world` ranks (summed in rank order); also adds dgamma += d_local_sums[C..2C), dbeta += d_local_sums[0..C).
 * d_g_out (may be NULL) receives g itself: the gradient an identity shortcut carries past the block. */
int pmd_bn_bwd_apply(const void* d_dz, const void* d_z, const void* d_y, const float* d_mean, const float* d_invstd,
                     const float* d_gamma, const float* const* d_sums_ranks, int world, const float* d_local_sums,
                     float count_total, int relu, void* d_dy, float* d_dgamma, float* d_dbeta, void* d_g_out,
                     long long rows, int C, void* stream);

/* Same result as pmd_bn_bwd_apply, for callers that cannot allocate device memory (a host that gave the
 * whole GPU to its own pool, or a loop that must not incur cudaMalloc's device-wide synchronisation).  It
 * has no counterpart in the reference.  The [2C] table-sum scratch is a __device__ array built into the
 * library (8 KB per GPU for the life of the process) instead of a cudaMalloc'ed buffer:
 *   - C may be at most 1024; larger C returns -1 with an explanatory pmd_last_error() and launches nothing;
 *   - the first call on each device resolves the array's address (cudaGetSymbolAddress): make that call at
 *     start-up, before any stream capture begins; no call ever allocates or frees device memory;
 *   - the array is shared by every call and keeps the previous call's contents (nothing needs to be zero:
 *     entries [0, 2C) are rewritten before they are read); calls that can overlap on the GPU race on it. */
int pmd_bn_bwd_apply_static(const void* d_dz, const void* d_z, const void* d_y, const float* d_mean,
                            const float* d_invstd, const float* d_gamma, const float* const* d_sums_ranks, int world,
                            const float* d_local_sums, float count_total, int relu, void* d_dy, float* d_dgamma,
                            float* d_dbeta, void* d_g_out, long long rows, int C, void* stream);

/* One or two BatchNorms that share the incoming gradient dz and its ReLU mask (a residual block's last
 * main-path BN and its projection-shortcut BN) handled in one pass each.  mask_mode: 0 none, 1 z > 0,
 * 2 recomputed as mask_scale*y0 + mask_shift > 0 from branch 0's forward affine (no read of z),
 * 3 d_z is the ReLU bit array written by pmd_bn_apply_bits / pmd_bn_finalize_apply_bits (uint8 [rows * C / 8]). */
typedef struct {
  const void* y;        /* conv output this BN normalised (bf16 NHWC) */
  const float* mean;
  const float* invstd;
  const float* gamma;   /* apply only */
  float* sums;          /* [2C] local sums: += by reduce, read by apply (dgamma/dbeta) */
  const float* totals;  /* [2C] totals over ranks; NULL = sums */
  void* dy;             /* apply only */
  float* dgamma;        /* apply only, may be NULL; += */
  float* dbeta;
  /* pivot rotation (optional, both NULL = off): at the end of a step's backward pass the statistics pivot of the NEXT
   * step becomes this step's batch mean (pivot[c] = pivot_new[c]); see pmd_bn_stats_pivot */
  float* pivot;
  const float* pivot_new;
} PmdBnBranch;
int pmd_bn_bwd_reduce_multi(const void* d_dz, const void* d_z, const float* d_mask_scale, const float* d_mask_shift,
                            int mask_mode, const PmdBnBranch* br, int nb, long long rows, int C, void* stream);
int pmd_bn_bwd_apply_multi(const void* d_dz, const void* d_z, const float* d_mask_scale, const float* d_mask_shift,
                           int mask_mode, const PmdBnBranch* br, int nb, float count_total, void* d_g_out,
                           long long rows, int C, void* stream);

/* Forward finalize + apply in ONE launch (no separate pmd_bn_finalize between the convolution and the apply):
 * z = act(BN(y) [+ res | + BN_res(res)]) with the batch statistics taken from `stats` ([sum, sumsq], as left by
 * pmd_conv_fprop); writes scale/shift/mean/invstd and updates the running statistics exactly like
 * pmd_bn_finalize.  f_res describes the projection shortcut's BatchNorm when the residual needs its own affine
 * (then d_res is that branch's conv output).  Single-rank statistics: under multi-rank SyncBatchNorm
 * pmd_bn_finalize_push (pmdist_comm.h) exchanges and finalises them and pmd_bn_apply follows. */
typedef struct {
  const float* stats;           /* [2C] local sums */
  const float* gamma;
  const float* beta;
  float* running_mean;          /* may be NULL */
  float* running_var;
  long long* num_batches;       /* may be NULL */
  float *scale, *shift, *mean, *invstd; /* [C] outputs */
  const float* pivot;           /* [C] pivot the statistics were taken around (NULL = 0) */
  float* pivot_new;             /* [C] receives the batch mean (NULL: not kept); must not alias pivot */
} PmdBnFinalize;
int pmd_bn_finalize_apply(const PmdBnFinalize* f, const PmdBnFinalize* f_res, float count, float momentum, float eps,
                          const void* d_y, const void* d_res, int relu, void* d_z, long long rows, int C, void* stream);
int pmd_bn_finalize_apply_bits(const PmdBnFinalize* f, const PmdBnFinalize* f_res, float count, float momentum, float eps,
                               const void* d_y, const void* d_res, int relu, void* d_z, void* d_relu_bits, long long rows,
                               int C, void* stream);   /* + ReLU bit array, see pmd_bn_apply_bits */

/* ------------------------------------------------------------------ head + loss (head.cu)
 * Global average pool + Linear, cross-entropy with top-1/top-5 counters (ref:model/resnet.py:102-104,
 * ref:main.py:48,105, ref:utils.py:64-77). */
int pmd_pool_fc_fwd(const void* d_z, int N, int HW, int C, const float* d_w, const float* d_b, int classes,
                    float* d_feat, float* d_logits, void* stream);
/* metrics (double[4]) += {sum loss, top-1 hits, top-5 hits, samples} over samples with valid != 0;
 * d_dlogits = (softmax - onehot) * grad_scale (NULL for evaluation). */
int pmd_ce_fwd_bwd(const float* d_logits, const long long* d_labels, const unsigned char* d_valid, int N,
                   int classes, float grad_scale, float* d_dlogits, double* d_metrics, void* stream);
int pmd_pool_fc_bwd(const float* d_dlogits, const float* d_feat, const float* d_w, int N, int HW, int C,
                    int classes, float* d_dw, float* d_db, void* d_dz, void* stream);

/* ------------------------------------------------------------------ optimiser (optim.cu)
 * SGD with weight decay, momentum and Nesterov on flat fp32 slabs; writes the bf16 working copy
 * in the same pass (ref:main.py:51-55,110; math = torch/optim/sgd.py). */
int pmd_sgd_step(float* d_param, const float* d_grad, float* d_mom, void* d_param_bf16, long long n, float lr,
                 float momentum, float weight_decay, int nesterov, float grad_scale, void* stream);
int pmd_cast_f32_to_bf16(const float* d_src, void* d_dst, long long n, void* stream);

/* ------------------------------------------------------------------ input pipeline (data_aug.cu)
 * uint8 NHWC RGB -> zero-pad crop + horizontal flip + ToTensor + Normalize(0.5, 0.5) -> NHWC bf16
 * with `c_out` >= 3 channels (extra channels zero) (ref:data.py:11-19).
 * aug = int32 [N,3] = (x offset, y offset, flip), offsets in [0, 2*pad]. */
int pmd_augment(const unsigned char* d_u8, const int* d_aug, int N, int H, int W, int pad, int c_out, void* d_out,
                void* stream);
/* Same augmentation fused with the stem convolution's im2col: out is [N, H, W, 64] bf16 whose first 27
 * columns are the 3x3x3 neighbourhood ((r*3+s)*3 + c order, zero outside the image), the rest zero.  The
 * 3->64 stem conv (ref:model/resnet.py:79-80) and its weight gradient then run as 64-wide 1x1
 * convolutions on the tensor-core path; pmd_pad_stem_weight makes the matching [K][64] weight.
 * d_out must be ZERO-INITIALISED once by the caller: the kernel writes columns 0..31 of every pixel only. */
int pmd_augment_im2col(const unsigned char* d_u8, const int* d_aug, int N, int H, int W, int pad, void* d_out,
                       void* stream);
/* The same kernel fed from a dataset that is RESIDENT in device memory (replaces the DataLoader workers, the pinned
 * staging and the per-step image H2D of ref:data.py:41-53 / ref:main.py:101): sample n of the
 * batch is stored image d_idx[n] of the dataset named by d_dataset_desc = device uint64 [2] {address of the uint8
 * [M,H,W,3] images, address of the int64 [M] labels} (read when the kernel runs) (a negative entry ~i means "image i, but do not count it": the evaluation
 * sampler's wrap-around duplicates), its label and validity are written to d_labels_out / d_valid_out for the
 * loss kernel, and crop offsets / flip are drawn inside the kernel from Philox4x32-10 with key (hdr[0], hdr[1]) and
 * counter (dataset index, hdr[2] = epoch, 0, 0): x = r0 % (2 pad + 1), y = r1 % (2 pad + 1), flip = r2 & 1
 * (hdr[3] == 0: evaluation, centre crop, no flip).  pmdist/data/philox.py reproduces the draws on the host. */
int pmd_augment_im2col_gather(const unsigned long long* d_dataset_desc, const int* d_hdr, const int* d_idx, int N, int H,
                              int W, int pad, void* d_out, long long* d_labels_out, unsigned char* d_valid_out,
                              void* stream);
int pmd_pad_stem_weight(const void* d_w_k_rsc, void* d_w_padded_k64, int K, int rsc, void* stream);

/* The whole first layer in one persistent kernel (csrc/stem_fused.cu): batch gather (resident dataset: d_dataset_desc,
 * d_hdr, d_idx, d_labels_out, d_valid_out as in pmd_augment_im2col_gather) or staged images (d_u8 [N,H,W,3], d_aug
 * [N,3]), crop/flip/normalise, the 3x3 3->64 stem convolution against the padded weight of pmd_pad_stem_weight, bf16
 * output y [N,H,W,64] and, when stats != NULL, its BatchNorm statistics around stat_pivot (may be NULL).  x0 != NULL
 * additionally receives the im2col rows (columns 0..31; the rest of the buffer must be zero) for the weight gradient.
 * Needs H*W % 128 == 0, W % 4 == 0, W <= 336.  Bit-identical to pmd_augment_im2col[_gather] followed by pmd_conv_fprop on x0. */
typedef struct PmdStemFused {
  const unsigned long long* d_dataset_desc;
  const int* d_hdr;
  const int* d_idx;
  long long* d_labels_out;
  unsigned char* d_valid_out;
  const unsigned char* d_u8;
  const int* d_aug;
  int N, H, W, pad;
  const void* w_pad;   /* padded [64][64] weight (w_rsc == 0), or the raw KRSC weight [64][w_rsc] (w_rsc = 27) */
  void* x0;
  void* y;
  float* stats;
  const float* stat_pivot;
  int w_rsc;
} PmdStemFused;
int pmd_stem_fused(const PmdStemFused* a, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PMDIST_H_ */
