/* libpmdist -- peer-memory transport (csrc/runtime.cpp, csrc/comm.cu).
 *
 * One process per GPU.  Each rank owns a "symmetric heap": the same allocation size on every rank,
 * every rank's heap mapped into every process, so a buffer is named by its byte OFFSET and rank w's
 * copy lives at bases[w] + offset.  Collectives are kernels that load/store peer heaps directly
 * over NVLink 5 / NVSwitch; cross-GPU ordering uses release/acquire flags in a signal pad at the
 * start of each heap, with monotonically increasing epochs kept on the device (so the kernels can
 * be captured in CUDA graphs and replayed).
 *
 * Replaces dist.init_process_group('nccl'), the NCCL communicator, DDP's Reducer collectives and
 * SyncBatchNorm's all_gather / all_reduce (ref:main.py:43-44, 190-193).
 */
#ifndef PMDIST_COMM_H_
#define PMDIST_COMM_H_

#include <stdint.h>

#include "pmdist.h"

#ifdef __cplusplus
extern "C" {
#endif

#define PMD_MAX_RANKS 8
#define PMD_MAX_SLOTS 4096
/* signal pad: flags uint32 [PMD_MAX_SLOTS][PMD_MAX_RANKS], then three local uint32 [PMD_MAX_SLOTS] arrays:
 * epochs, arrival tickets and published-totals markers (the last two serve the grid-wide fused exchanges) */
#define PMD_FLAGS_BYTES (PMD_MAX_SLOTS * PMD_MAX_RANKS * 4)
#define PMD_SIGNAL_BYTES (PMD_FLAGS_BYTES + 3 * PMD_MAX_SLOTS * 4)

enum { PMD_TRANSPORT_SINGLE = 0, PMD_TRANSPORT_VMM = 1, PMD_TRANSPORT_IPC = 2 };
enum { PMD_COMM_FORCE_IPC = 1, PMD_COMM_NO_MULTICAST = 2 };
enum { PMD_ALGO_AUTO = 0, PMD_ALGO_ONESHOT = 1, PMD_ALGO_TWOSHOT = 2, PMD_ALGO_NVLS = 3, PMD_ALGO_LL = 4 };

typedef struct PmdRendezvous PmdRendezvous;
typedef struct PmdComm PmdComm;

typedef struct {
  int rank, world, device;
  int transport;           /* PMD_TRANSPORT_* */
  int multicast_supported; /* CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED */
  int multicast_mapped;    /* a multimem-capable mapping of the heap exists */
  int p2p_ok;              /* cudaDeviceCanAccessPeer to every peer */
  int pad_;
  unsigned long long heap_bytes;
  unsigned long long local_base; /* this rank's heap */
  unsigned long long mc_base;    /* multicast mapping or 0 */
  unsigned long long d_bases;    /* device array [world] of heap base addresses */
  char note[256];
} PmdCommInfo;

/* Device-visible description of the transport, passed by value to kernels. */
typedef struct {
  int rank, world;
  const unsigned long long* d_bases; /* [world] */
  unsigned long long mc_base;        /* 0 if no multicast */
} PmdSym;

/* ---- host-side rendezvous (usable without CUDA: tests run it on CPU) */
int pmd_rdv_create(int rank, int world, const char* run_id, double timeout_s, PmdRendezvous** out);
int pmd_rdv_destroy(PmdRendezvous* r);
int pmd_rdv_allgather(PmdRendezvous* r, const void* in, int bytes, void* out);
int pmd_rdv_barrier(PmdRendezvous* r);
int pmd_rdv_allgather_fds(PmdRendezvous* r, int fd, int* fds_out);

/* ---- topology probe (no communicator needed): peer reachability, link performance rank, cross-link atomics
 * for every ordered device pair (row-major [max_devices x max_devices]) and multicast/NVLS support per device. */
int pmd_topology_probe(int max_devices, int* n_devices, int* can_access, int* perf_rank, int* native_atomics,
                       int* multicast);

/* ---- communicator */
int pmd_comm_create(int rank, int world, const char* run_id, int device, unsigned long long heap_bytes, int flags,
                    PmdComm** out);
int pmd_comm_destroy(PmdComm* c);
int pmd_comm_info(PmdComm* c, PmdCommInfo* info);
unsigned long long pmd_comm_peer_base(PmdComm* c, int rank);
long long pmd_comm_alloc(PmdComm* c, unsigned long long bytes, unsigned long long align);
int pmd_comm_host_barrier(PmdComm* c);
int pmd_comm_allgather_host(PmdComm* c, const void* in, int bytes, void* out);

/* ---- device collectives (comm.cu).  `slot` selects flag words in the signal pad; a slot must not
 * be used by two kernels that can be in flight at once.  Multi-block kernels use
 * [slot, slot + blocks). */

/* All ranks rendezvous on the device (no data). */
int pmd_dev_barrier(const PmdSym* sym, int slot, void* stream);

/* Fused gradient all-reduce + SGD (ref:main.py:109-110 + DDP Reducer):
 *   grad (fp32, heap offset grad_off) is summed over ranks in rank order, scaled by 1/world, and the
 *   SGD-Nesterov update is applied to param / momentum; the bf16 working copy is rewritten.
 *   ONESHOT: every rank reads all peers' range and updates all of it locally.
 *   TWOSHOT: rank r reduces and updates slice r, then stores the new fp32 + bf16 weights into every
 *            peer's heap (owner-computes; momentum for slice r lives on rank r only).
 *   NVLS:    as TWOSHOT with multimem.ld_reduce / multimem.st through the NVSwitch.
 *   LL:      for ranges of at most PMD_LL_MAX_ELEMS elements (the step's LAST, un-overlappable bucket): every rank pushes
 *            its gradients into every peer's receive area as {value, epoch} words (a->recv_off, PMD_LL_RECV_BYTES(world,
 *            n) bytes, zero-initialised, owned by this range), polls its own, and updates the whole range itself -- one
 *            one-way NVLink trip and no cross-rank barrier (the barrier-based kernels floor at ~16 us at 8 GPUs).
 * All offsets are byte offsets into the symmetric heap; `n` elements starting at element `first`. */
typedef struct {
  unsigned long long param_off, grad_off, mom_off, bf16_off;
  long long first, n;
  float lr, momentum, weight_decay;
  int nesterov;
  int algo;   /* PMD_ALGO_* (AUTO resolved by the caller) */
  int blocks; /* CTAs (== flag slots used) */
  const float* d_lr; /* if non-NULL the learning rate is read from device memory (CUDA-graph replays) */
  unsigned long long recv_off; /* PMD_ALGO_LL only: heap offset of the range's receive area */
} PmdAllreduceSgd;
#define PMD_LL_MAX_ELEMS 16384
int pmd_allreduce_sgd(const PmdSym* sym, const PmdAllreduceSgd* a, int slot, void* stream);

/* Plain all-reduce (sum) of n floats: out[i] = sum over ranks of in_w[i], delivered to every rank's
 * `out_off` (in_off != out_off).  The NCCL-comparable primitive for the bucket-size sweep
 * (BASELINE.json config 5). */
int pmd_allreduce_f32(const PmdSym* sym, unsigned long long in_off, unsigned long long out_off, long long n, int algo,
                      int blocks, int slot, void* stream);

/* Every rank copies `bytes` at heap offset `off` from rank `src` (initial weight broadcast, C2). */
int pmd_broadcast(const PmdSym* sym, unsigned long long off, unsigned long long bytes, int src, int blocks, int slot,
                  void* stream);

/* SyncBatchNorm exchanges (replace the all_gather / all_reduce of torch's SyncBatchNorm,
 * torch/nn/modules/_functions.py:65-79, 155-165).  Push ("LL") protocol: every rank stores its vector into every
 * peer's receive area as 8-byte {value, epoch} words and polls only its own memory -- one one-way NVLink trip.
 * recv_off: heap offset (same on every rank) of a zero-initialised area of PMD_LL_RECV_BYTES(world, n) bytes that
 * belongs to this exchange alone; `slot` names the exchange's device-resident epoch counter.  Single-block kernels,
 * launched with programmatic stream serialization directly behind the kernel that accumulated the vector.
 * (Letting that producer's last CTA do the exchange instead was built and measured slower:
 * profiles/r2_syncbn_exchange_placement.md.) */
#define PMD_LL_RECV_BYTES(world, n) (2ull * (world) * (n) * 8ull)
/* forward: exchange [sum, sumsq] (taken around d_pivot, which must be identical on every rank; NULL = 0), then
 * finalize like pmd_bn_finalize over the rank-ordered totals; d_pivot_new (may be NULL) receives the batch mean */
int pmd_bn_finalize_push(const PmdSym* sym, int slot, const float* d_stats, unsigned long long recv_off,
                         float count_per_rank, int C, const float* d_gamma, const float* d_beta, float* d_running_mean,
                         float* d_running_var, long long* d_num_batches, float momentum, float eps, float* d_scale,
                         float* d_shift, float* d_mean, float* d_invstd, const float* d_pivot, float* d_pivot_new,
                         void* stream);
/* backward / generic: total[i] = sum over ranks (in rank order) of local[i]; one or two vectors per launch */
typedef struct {
  const float* local;          /* this rank's vector (any device pointer) */
  float* total;                /* local output */
  unsigned long long recv_off;
  int n, slot;
} PmdSmallExchange;
int pmd_small_allreduce_push(const PmdSym* sym, const PmdSmallExchange* x, int nvec, void* stream);
/* metrics: double[4] at heap offset `off` on every rank -> d_total (local) = sum over ranks. */
int pmd_metric_allreduce(const PmdSym* sym, int slot, unsigned long long off, double* d_total, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PMDIST_COMM_H_ */
