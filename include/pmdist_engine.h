/* libpmdist -- the step engine (csrc/engine.cu): a static plan of kernel launches for one CIFAR-style
 * ResNet training / evaluation step over preallocated buffers, replayed as a CUDA graph.
 *
 * It replaces, for the GPU path, everything the reference runs between `output = model(input)` and
 * `optimizer.step()` (ref:main.py:101-114, 144-156): the H2D'd uint8 batch is augmented on the GPU,
 * forward / loss / backward run as hand-written sm_100a kernels (include/pmdist.h), gradients are
 * written straight into the flat bucket slab, and each bucket's fused all-reduce + SGD kernel
 * (include/pmdist_comm.h) is launched on a side stream as soon as backward has produced it.
 * There is no autograd graph, no Python in the step and no host synchronisation.
 */
#ifndef PMDIST_ENGINE_H_
#define PMDIST_ENGINE_H_

#include "pmdist_comm.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct PmdEngine PmdEngine;

typedef struct {
  /* topology: ref:model/resnet.py:74-125 */
  int bottleneck;    /* 0 = BasicBlock, 1 = Bottleneck */
  int blocks[4];     /* blocks per stage */
  int num_classes;
  int image_hw;      /* square input side */
  int max_batch;     /* per-rank batch the buffers are sized for */
  /* parameter slab layout (element offsets; conv order = stem, then per block main path + shortcut) */
  int n_convs;
  const long long* conv_w_off;
  const long long* bn_g_off;
  const long long* bn_b_off;
  const int* conv_w_bucket;
  const int* bn_g_bucket;
  const int* bn_b_bucket;
  long long fc_w_off, fc_b_off;
  int fc_w_bucket, fc_b_bucket;
  long long slab_elems;
  int n_buckets;
  const long long* bucket_start; /* element offsets */
  const long long* bucket_numel;
  /* BatchNorm buffers live in torch-owned memory (they are the module's buffers) */
  float* d_bn_float;             /* base of the flat running_mean / running_var tensor */
  const long long* bn_rm_off;
  const long long* bn_rv_off;
  long long* d_bn_nbt;           /* int64 [n_convs] num_batches_tracked */
  /* hyper-parameters */
  float momentum, weight_decay;
  int nesterov;
  float bn_momentum, bn_eps;
  int sync_bn;   /* exchange BatchNorm statistics across ranks */
  int use_graph; /* capture steps into CUDA graphs */
  int algo;      /* PMD_ALGO_* for the gradient buckets */
  int comm_blocks;
  /* optional per-bucket overrides chosen from the measured all-reduce table (NULL = algo / comm_blocks) */
  const int* bucket_algo;
  const int* bucket_blocks;
} PmdEngineConfig;

typedef struct {
  unsigned long long param, grad, mom, bf16; /* device addresses of the flat slabs (in the symmetric heap) */
  unsigned long long in_u8, in_aug, in_labels, in_valid; /* device staging of one step's inputs */
  unsigned long long in_index; /* int32 [2][4 + max_batch]: per input slot {seed, stream, epoch, train} + dataset indices (resident mode) */
  unsigned long long metrics_snap; /* double [2][4]: cumulative metrics as of the end of the last step run with each slot */
  unsigned long long logits;                 /* fp32 [max_batch, classes] of the last step */
  unsigned long long metrics;                /* double[4] running {loss sum, top1, top5, count} */
  long long slab_elems;
  long long activation_bytes, heap_bytes_used;
} PmdEngineBuffers;

/* `comm` may describe a 1-rank world. The engine allocates its slabs from comm's symmetric heap. */
int pmd_engine_create(const PmdEngineConfig* cfg, PmdComm* comm, PmdEngine** out);
int pmd_engine_destroy(PmdEngine* e);
int pmd_engine_buffers(PmdEngine* e, PmdEngineBuffers* out);
/* After the fp32 parameter slab was written from the host side (init, load_state_dict): optionally
 * broadcast rank 0's slab + BN buffers to all ranks (C2), then rebuild the bf16 / dgrad copies. */
int pmd_engine_sync_params(PmdEngine* e, int broadcast_from_rank0, void* stream);
/* Device-resident dataset: images uint8 [count, hw, hw, 3] and labels int64 [count] in device
 * memory owned by the caller (must outlive the engine or a later call with NULL).  While attached, a step takes its
 * batch from `in_index` (header + dataset indices, ~2 KB of H2D per step) instead of in_u8 / in_aug / in_labels:
 * gather, crop/flip (in-kernel Philox), normalise and stem im2col run in one kernel (pmd_augment_im2col_gather).
 * Passing d_images == NULL detaches.  Switching between resident datasets (train / test) is a synchronous 16-byte
 * descriptor update (twice per epoch); only attaching / detaching re-captures the CUDA graphs. */
int pmd_engine_set_dataset(PmdEngine* e, const void* d_images_u8, const long long* d_labels, long long count);
/* One training step on the n <= max_batch samples currently in the staging buffers.  `stream` is the caller's stream:
 * the step is ordered after the work already enqueued there (the input copy) and before what is enqueued next (the
 * metric read-back).  Replays launch the captured CUDA graph directly INTO that stream (no cross-stream hop per step);
 * the first call per batch size runs eagerly on the engine's own streams, bracketed by events. */
int pmd_engine_train_step(PmdEngine* e, int n, float lr, void* stream);
/* The same on input slot 0 or 1 of `in_index` (device-resident dataset only).  Two slots let a caller copy step i+1's
 * index vector (on its own copy stream) while step i runs, and read step i's metric snapshot (metrics_snap[slot], written
 * by the step's last node) while step i+1 runs: consecutive steps then run back to back on the GPU although every step
 * has a host->device input copy and a device->host result read.  pmd_engine_train_step == slot 0. */
int pmd_engine_train_step_slot(PmdEngine* e, int n, float lr, int slot, void* stream);
/* Forward + loss/metrics only, BatchNorm in evaluation mode. */
int pmd_engine_eval_step(PmdEngine* e, int n, void* stream);
/* Copies the (optionally cross-rank reduced) metric accumulators to the host and optionally clears them. */
int pmd_engine_read_metrics(PmdEngine* e, int reduce_over_ranks, int reset, double* out4, void* stream);
/* Measurement switch for "exposed communication" (BASELINE.json metric): with bit 0 set the
 * SyncBatchNorm statistics stay rank-local, with bit 1 set every gradient bucket is reduced over this rank alone --
 * i.e. the step runs the kernels of a 1-GPU run while the other ranks do the same next to it.  step time with
 * mask 0 minus step time with mask 3 = communication (and rank skew) the step could not hide.  Ranks diverge while a
 * stub is active: use it at the end of a benchmark only.  Drops the cached plans / graphs (next step re-captures). */
int pmd_engine_set_comm_stub(PmdEngine* e, int mask);
/* kernels launched by one train step of batch n (graph nodes count the same as direct launches) */
int pmd_engine_launches_per_step(PmdEngine* e, int n);

#ifdef __cplusplus
}
#endif
#endif /* PMDIST_ENGINE_H_ */
