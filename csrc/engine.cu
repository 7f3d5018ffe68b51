// The step engine: builds a static launch plan for one ResNet training / evaluation step over
// preallocated buffers and replays it (as a CUDA graph after the first call).
//
// Replaces torch autograd + DDP Reducer + optimizer for the GPU path (ref:main.py:87-171; SURVEY.md
// 3.3, 3.4, W10, W25).  What runs per training step, all on the device:
//   memset(grad slab, BN partials) -> augment -> [conv fprop (+BN partials) -> BN finalize (peer
//   exchange when SyncBN) -> BN apply (+residual, ReLU)]* -> pool+FC -> CE (+metrics) -> backward
//   (BN reduce / exchange / apply, wgrad straight into the bucket slab, dgrad with the shortcut
//   branch fused) -> per bucket, as soon as its last gradient exists: fused all-reduce + SGD on a
//   side stream.
#include <cstdlib>
#include <map>
#include <memory>
#include <vector>

#include "common.cuh"
#include "conv_tc.cuh"
#include "../include/pmdist.h"
#include "../include/pmdist_engine.h"

using namespace pmd;
typedef __nv_bfloat16 bf16;

#define E_CHECK(expr)          \
  do {                         \
    if ((expr) != 0) return -1; \
  } while (0)

namespace {

struct Conv {
  PmdConvShape s{};      // s.N filled per plan
  int Ho = 0, Wo = 0;
  bool tc = false;
  bool stem_col = false;   // 3->64 stem run as a 64-wide 1x1 conv over the im2col'ed augmented input
  int real_rsc = 0;        // R*S*C of the parameter as stored in the slab (27 for the stem)
  long long w_off = 0, g_off = 0, b_off = 0;
  int w_bkt = 0, g_bkt = 0, b_bkt = 0;
  float *rm = nullptr, *rv = nullptr;
  long long* nb = nullptr;
  int x_cstride = 0;
  const bf16* x = nullptr;
  bf16 *y = nullptr, *dy = nullptr;
  float *scale = nullptr, *shift = nullptr, *mean = nullptr, *invstd = nullptr;
  float *pivot = nullptr, *pivot_new = nullptr;   // statistics pivot (previous step's batch mean) and its successor
  float *stats = nullptr, *sums = nullptr, *sums_total = nullptr, *stats_total = nullptr;
  unsigned long long stats_off = 0, sums_off = 0;
  unsigned long long recv_fwd_off = 0, recv_bwd_off = 0;   // push-exchange receive areas (multi-rank SyncBN)
  const float** tab_stats = nullptr;   // device: 1-entry pointer tables for the local-statistics path
  const float** tab_sums = nullptr;
  const float** tab_total = nullptr;
};

struct Block {
  int main[3] = {-1, -1, -1};
  int n_main = 0;
  int sc = -1;
  bf16 *z_in = nullptr, *dz_in = nullptr;
  bf16* z_mid[2] = {nullptr, nullptr};
  bf16* dz_mid[2] = {nullptr, nullptr};
  bf16 *z_out = nullptr, *dz_out = nullptr, *g_out = nullptr;
  unsigned char* zbits = nullptr;   // ReLU bit array of z_out (one bit per element): backward's mask instead of z_out
  int Hin = 0, Hout = 0, Cin = 0, Cout = 0;
};

struct ConvPlan {
  ConvGemmParams fprop, fprop_eval;
  ConvResidentParams fprop_r, fprop_eval_r;   // used instead when resident == true (3x3 stride-1, C = K = 64)
  bool resident = false;
  WgradParams wgrad;
};

struct Plan {
  int n = 0;
  std::vector<ConvPlan> conv;
  std::vector<ConvGemmParams> block_dgrad;   // one per block (conv0 + fused shortcut / addend)
  std::vector<ConvGemmParams> sc_dgrad;      // per block: shortcut alone, when it cannot share conv0's tile grid
  std::vector<ConvGemmParams> inner_dgrad;   // per conv index (main[j], j >= 1)
  std::vector<ConvResidentParams> block_dgrad_r, inner_dgrad_r;
  std::vector<char> block_resident, inner_resident;
  std::vector<char> inner_reduced;          // inner dgrad also does the BatchNorm-backward reduction of the layer below
  bool stem_reduced = false;                // the first block's dgrad does it for the stem's BatchNorm
  bool stem_fused = false;                  // input pipeline + stem convolution + statistics in one kernel (stem_fused.cu)
  StemFusedParams stem_f;
  std::vector<char> g_out_unused;           // per identity block: the dgrad masks dz_out itself (ReLU bits), g_out is not written
  std::vector<char> out_reduced;            // per block: the NEXT block's dgrad does it for this block's output BatchNorm(s)
  cudaGraphExec_t train_graph = nullptr, eval_graph = nullptr;
  cudaGraphExec_t train_graph_alt = nullptr;   // the same step reading input slot 1 / writing metric snapshot 1
  int train_launches = 0, eval_launches = 0;
  bool train_warm = false, eval_warm = false;
};

}  // namespace

struct PmdEngine {
  PmdEngineConfig cfg{};
  PmdComm* comm = nullptr;
  PmdCommInfo info{};
  PmdSym sym{};
  std::vector<Conv> convs;
  std::vector<Block> blocks;
  std::vector<long long> bucket_start, bucket_numel;
  std::vector<int> bucket_params;     // number of parameter tensors per bucket
  std::vector<int> bucket_algo, bucket_blocks;
  std::vector<unsigned long long> bucket_recv;   // per bucket: receive area of the push (LL) all-reduce, or 0
  int sm_count = 148;
  // slabs (symmetric heap)
  unsigned long long param_off = 0, grad_off = 0, mom_off = 0, bf16_off = 0, metrics_off = 0, bnpart_off = 0;
  long long bnpart_bytes = 0;
  float *param = nullptr, *grad = nullptr, *mom = nullptr;
  bf16* wbf16 = nullptr;
  double* metrics = nullptr;
  // local buffers
  std::vector<void*> owned;
  // PMD_COMM_STUB-style measurement switch (pmd_engine_set_comm_stub): bit 0 = SyncBatchNorm statistics stay local,
  // bit 1 = gradient buckets are "reduced" over this rank alone.  The kernels are the ones a 1-GPU run uses.
  int comm_stub = 0;
  PmdSym sym_local{};                 // this rank seen as a world of one (d_bases -> d_local_base)
  unsigned long long* d_local_base = nullptr;
  bf16* x0 = nullptr;
  bf16* stem_w_pad = nullptr;         // [64][64] zero-padded stem weight
  unsigned char *in_u8 = nullptr, *in_valid = nullptr;
  int* in_aug = nullptr;
  int* in_index = nullptr;            // [2][4 + max_batch]: Philox header + dataset indices (resident-dataset mode);
  int in_slot = 0;                    // two slots, so step i+1's indices can be copied in while step i runs
  double* metrics_snap = nullptr;     // [2][4]: the cumulative metrics as of the end of the last step of each slot
  unsigned long long* d_ds = nullptr; // device {images address, labels address} of the active resident dataset
  bool ds_on = false;
  long long ds_count = 0;
  long long* in_labels = nullptr;
  float *feat = nullptr, *logits = nullptr, *dlogits = nullptr, *d_lr = nullptr;
  double* metrics_total = nullptr;
  bf16 *z_stem = nullptr, *dz_stem = nullptr;
  long long activation_bytes = 0;
  cudaStream_t comm_stream = nullptr, main_stream = nullptr, wg_stream = nullptr;
  std::vector<cudaEvent_t> wg_events;
  int wg_event_next = 0;
  cudaEvent_t ev_in = nullptr, ev_out = nullptr;
  std::vector<cudaEvent_t> events;
  cudaEvent_t ev_join = nullptr;
  std::map<int, std::unique_ptr<Plan>> plans;
  float last_lr = -1.f;

  template <typename T>
  int dalloc(T** p, long long count) {
    void* q = nullptr;
    const size_t bytes = static_cast<size_t>(count) * sizeof(T);
    PMD_CHECK_CUDA(cudaMalloc(&q, bytes > 0 ? bytes : 16));
    PMD_CHECK_CUDA(cudaMemset(q, 0, bytes > 0 ? bytes : 16));
    owned.push_back(q);
    activation_bytes += static_cast<long long>(bytes);
    *p = static_cast<T*>(q);
    return 0;
  }
  int halloc(unsigned long long* off, unsigned long long bytes) {
    const long long o = pmd_comm_alloc(comm, bytes, 1024);
    if (o < 0) return -1;
    *off = static_cast<unsigned long long>(o);
    return 0;
  }
  template <typename T>
  T* hptr(unsigned long long off) { return reinterpret_cast<T*>(info.local_base + off); }
};

namespace {

bool relu_bits_enabled();

int build_topology(PmdEngine* e) {
  const PmdEngineConfig& c = e->cfg;
  const int B = c.max_batch, HW = c.image_hw;
  const int expansion = c.bottleneck ? 4 : 1;
  const int planes[4] = {64, 128, 256, 512}, strides[4] = {1, 2, 2, 2};
  auto add_conv = [&](int H, int Cin, int K, int R, int stride) {
    Conv cv;
    cv.s = PmdConvShape{B, H, H, Cin, K, R, stride};
    cv.Ho = conv_out_dim(H, R, stride);
    cv.Wo = cv.Ho;
    cv.tc = conv_tc_supported(cv.s);
    cv.x_cstride = Cin;
    e->convs.push_back(cv);
    return static_cast<int>(e->convs.size()) - 1;
  };
  const int stem = add_conv(HW, 64, 64, 1, 1);   // see data_aug.cu: augmentation emits the 27-column im2col, padded to 64
  e->convs[stem].stem_col = true;
  e->convs[stem].real_rsc = 27;
  int H = HW, Cin = 64;
  for (int si = 0; si < 4; ++si)
    for (int bi = 0; bi < c.blocks[si]; ++bi) {
      Block b;
      const int stride = bi == 0 ? strides[si] : 1;
      const int P = planes[si], Cout = P * expansion;
      b.Hin = H;
      b.Cin = Cin;
      b.Cout = Cout;
      if (!c.bottleneck) {
        b.n_main = 2;
        b.main[0] = add_conv(H, Cin, P, 3, stride);
        b.main[1] = add_conv(e->convs[b.main[0]].Ho, P, P, 3, 1);
      } else {
        b.n_main = 3;
        b.main[0] = add_conv(H, Cin, P, 1, 1);
        b.main[1] = add_conv(H, P, P, 3, stride);
        b.main[2] = add_conv(e->convs[b.main[1]].Ho, P, Cout, 1, 1);
      }
      if (stride != 1 || Cin != Cout) b.sc = add_conv(H, Cin, Cout, 1, stride);
      b.Hout = e->convs[b.main[b.n_main - 1]].Ho;
      H = b.Hout;
      Cin = Cout;
      e->blocks.push_back(b);
    }
  PMD_REQUIRE(static_cast<int>(e->convs.size()) == c.n_convs, "conv count disagrees with the Python model spec");
  return 0;
}

int allocate(PmdEngine* e) {
  const PmdEngineConfig& c = e->cfg;
  const long long S = c.slab_elems;
  const int B = c.max_batch;
  E_CHECK(e->halloc(&e->param_off, S * 4));
  E_CHECK(e->halloc(&e->grad_off, S * 4));
  E_CHECK(e->halloc(&e->mom_off, S * 4));
  E_CHECK(e->halloc(&e->bf16_off, S * 2));
  E_CHECK(e->halloc(&e->metrics_off, 64));
  long long part = 0;
  for (Conv& cv : e->convs) part += 4LL * cv.s.K * 4;   // stats[2K] + sums[2K]
  e->bnpart_bytes = part;
  E_CHECK(e->halloc(&e->bnpart_off, part));
  if (e->sym.world > 1 && c.sync_bn) {
    for (Conv& cv : e->convs) {
      const unsigned long long bytes = PMD_LL_RECV_BYTES(e->sym.world, 2 * cv.s.K);
      E_CHECK(e->halloc(&cv.recv_fwd_off, bytes));
      E_CHECK(e->halloc(&cv.recv_bwd_off, bytes));
      PMD_CHECK_CUDA(cudaMemset(e->hptr<char>(cv.recv_fwd_off), 0, bytes));
      PMD_CHECK_CUDA(cudaMemset(e->hptr<char>(cv.recv_bwd_off), 0, bytes));
    }
  }
  e->param = e->hptr<float>(e->param_off);
  e->grad = e->hptr<float>(e->grad_off);
  e->mom = e->hptr<float>(e->mom_off);
  e->wbf16 = e->hptr<bf16>(e->bf16_off);
  e->metrics = e->hptr<double>(e->metrics_off);
  PMD_CHECK_CUDA(cudaMemset(e->param, 0, S * 4));
  PMD_CHECK_CUDA(cudaMemset(e->grad, 0, S * 4));
  PMD_CHECK_CUDA(cudaMemset(e->mom, 0, S * 4));
  PMD_CHECK_CUDA(cudaMemset(e->wbf16, 0, S * 2));
  PMD_CHECK_CUDA(cudaMemset(e->metrics, 0, 64));
  PMD_CHECK_CUDA(cudaMemset(e->hptr<char>(e->bnpart_off), 0, part));

  E_CHECK(e->dalloc(&e->d_local_base, 1));
  PMD_CHECK_CUDA(cudaMemcpy(e->d_local_base, &e->info.local_base, sizeof(unsigned long long), cudaMemcpyHostToDevice));
  e->sym_local.rank = 0;
  e->sym_local.world = 1;
  e->sym_local.d_bases = e->d_local_base;
  e->sym_local.mc_base = 0;
  E_CHECK(e->dalloc(&e->x0, 1LL * B * c.image_hw * c.image_hw * 64));
  E_CHECK(e->dalloc(&e->stem_w_pad, 64 * 64));
  E_CHECK(e->dalloc(&e->in_u8, 1LL * B * c.image_hw * c.image_hw * 3));
  E_CHECK(e->dalloc(&e->in_aug, 3LL * B));
  E_CHECK(e->dalloc(&e->in_index, 2 * (4LL + B)));
  E_CHECK(e->dalloc(&e->metrics_snap, 8));
  E_CHECK(e->dalloc(&e->d_ds, 2));
  E_CHECK(e->dalloc(&e->in_labels, B));
  E_CHECK(e->dalloc(&e->in_valid, B));
  E_CHECK(e->dalloc(&e->feat, 1LL * B * 512 * (c.bottleneck ? 4 : 1)));
  E_CHECK(e->dalloc(&e->logits, 1LL * B * c.num_classes));
  E_CHECK(e->dalloc(&e->dlogits, 1LL * B * c.num_classes));
  E_CHECK(e->dalloc(&e->d_lr, 4));
  E_CHECK(e->dalloc(&e->metrics_total, 4));

  unsigned long long poff = e->bnpart_off;
  for (size_t i = 0; i < e->convs.size(); ++i) {
    Conv& cv = e->convs[i];
    const long long out_elems = 1LL * B * cv.Ho * cv.Wo * cv.s.K;
    E_CHECK(e->dalloc(&cv.y, out_elems));
    E_CHECK(e->dalloc(&cv.dy, out_elems));
    float* f = nullptr;
    E_CHECK(e->dalloc(&f, 10LL * cv.s.K));
    cv.pivot = f + 8 * cv.s.K;       // zero-initialised: the first step takes plain sums
    cv.pivot_new = f + 9 * cv.s.K;
    cv.scale = f;
    cv.shift = f + cv.s.K;
    cv.mean = f + 2 * cv.s.K;
    cv.invstd = f + 3 * cv.s.K;
    cv.sums_total = f + 4 * cv.s.K;
    cv.stats_total = f + 6 * cv.s.K;
    cv.stats_off = poff;
    cv.sums_off = poff + 2ULL * cv.s.K * 4;
    poff += 4ULL * cv.s.K * 4;
    cv.stats = e->hptr<float>(cv.stats_off);
    cv.sums = e->hptr<float>(cv.sums_off);
    const float** tabs = nullptr;
    E_CHECK(e->dalloc(&tabs, 4));
    const float* host_tab[3] = {cv.stats, cv.sums, cv.sums_total};
    PMD_CHECK_CUDA(cudaMemcpy(tabs, host_tab, sizeof(host_tab), cudaMemcpyHostToDevice));
    cv.tab_stats = tabs;
    cv.tab_sums = tabs + 1;
    cv.tab_total = tabs + 2;
    cv.w_off = c.conv_w_off[i];
    cv.g_off = c.bn_g_off[i];
    cv.b_off = c.bn_b_off[i];
    cv.w_bkt = c.conv_w_bucket[i];
    cv.g_bkt = c.bn_g_bucket[i];
    cv.b_bkt = c.bn_b_bucket[i];
    cv.rm = c.d_bn_float + c.bn_rm_off[i];
    cv.rv = c.d_bn_float + c.bn_rv_off[i];
    cv.nb = c.d_bn_nbt + i;
  }
  // activations between convs
  Conv& st = e->convs[0];
  E_CHECK(e->dalloc(&e->z_stem, 1LL * B * st.Ho * st.Wo * st.s.K));
  E_CHECK(e->dalloc(&e->dz_stem, 1LL * B * st.Ho * st.Wo * st.s.K));
  st.x = e->x0;
  bf16 *z_prev = e->z_stem, *dz_prev = e->dz_stem;
  for (Block& b : e->blocks) {
    b.z_in = z_prev;
    b.dz_in = dz_prev;
    const bf16* x = b.z_in;
    for (int j = 0; j < b.n_main; ++j) {
      Conv& cv = e->convs[b.main[j]];
      cv.x = x;
      if (j < b.n_main - 1) {
        const long long el = 1LL * B * cv.Ho * cv.Wo * cv.s.K;
        E_CHECK(e->dalloc(&b.z_mid[j], el));
        E_CHECK(e->dalloc(&b.dz_mid[j], el));
        x = b.z_mid[j];
      }
    }
    if (b.sc >= 0) e->convs[b.sc].x = b.z_in;
    const long long oel = 1LL * B * b.Hout * b.Hout * b.Cout;
    E_CHECK(e->dalloc(&b.z_out, oel));
    E_CHECK(e->dalloc(&b.dz_out, oel));
    if (b.sc < 0) E_CHECK(e->dalloc(&b.g_out, oel));
    if (relu_bits_enabled()) E_CHECK(e->dalloc(&b.zbits, oel / 8));
    z_prev = b.z_out;
    dz_prev = b.dz_out;
  }
  return 0;
}

// PMD_BN_RELU_BITS=0: the block-output BatchNorm backward reads z (16 bits per element) for its ReLU mask instead of
// the bit array the forward apply leaves behind (1 bit per element).
bool relu_bits_enabled() {
  static const bool on = getenv("PMD_BN_RELU_BITS") == nullptr || atoi(getenv("PMD_BN_RELU_BITS")) != 0;
  return on;
}

// Multi-rank SyncBatchNorm is in effect (and not stubbed out for a communication-free timing run).
bool sync_bn_active(const PmdEngine* e) { return e->sym.world > 1 && e->cfg.sync_bn && (e->comm_stub & 1) == 0; }

Plan* get_plan(PmdEngine* e, int n) {
  auto it = e->plans.find(n);
  if (it != e->plans.end()) return it->second.get();
  std::unique_ptr<Plan> pl(new Plan());
  pl->n = n;
  pl->conv.resize(e->convs.size());
  pl->inner_dgrad.resize(e->convs.size());
  pl->block_dgrad.resize(e->blocks.size());
  pl->sc_dgrad.resize(e->blocks.size());
  pl->block_dgrad_r.resize(e->blocks.size());
  pl->inner_dgrad_r.resize(e->convs.size());
  pl->block_resident.assign(e->blocks.size(), 0);
  pl->inner_resident.assign(e->convs.size(), 0);
  pl->inner_reduced.assign(e->convs.size(), 0);
  pl->out_reduced.assign(e->blocks.size(), 0);
  pl->g_out_unused.assign(e->blocks.size(), 0);
  // C = K = 64 3x3 layers: weight-resident persistent kernel
  const bool resident_on = conv_resident_enabled(), resident_dgrad = conv_resident_dgrad_enabled();
  for (size_t i = 0; i < e->convs.size(); ++i) {
    Conv& cv = e->convs[i];
    if (!cv.tc) continue;
    PmdConvShape s = cv.s;
    s.N = n;
    const bf16* w = cv.stem_col ? e->stem_w_pad : e->wbf16 + cv.w_off;
    if (build_fprop_params(s, cv.x, w, cv.y, cv.stats, &pl->conv[i].fprop)) return nullptr;
    pl->conv[i].fprop.stat_pivot = cv.pivot;
    pl->conv[i].fprop_eval = pl->conv[i].fprop;
    pl->conv[i].fprop_eval.stats = nullptr;
    if (!cv.stem_col && resident_on && conv_resident_supported(s)) {
      if (build_fprop_resident_params(s, cv.x, w, cv.y, cv.stats, &pl->conv[i].fprop_r)) return nullptr;
      pl->conv[i].fprop_r.stat_pivot = cv.pivot;
      pl->conv[i].fprop_eval_r = pl->conv[i].fprop_r;
      pl->conv[i].fprop_eval_r.stats = nullptr;
      pl->conv[i].resident = true;
    }
    if (build_wgrad_params(s, cv.x, cv.dy, e->grad + cv.w_off, 0, e->sm_count, &pl->conv[i].wgrad,
                           cv.stem_col ? cv.real_rsc : 0, cv.stem_col ? cv.real_rsc : 64))
      return nullptr;
  }
  if (e->convs[0].stem_col && e->convs[0].tc && stem_fused_enabled() && stem_fused_supported(n, e->cfg.image_hw, e->cfg.image_hw)) {
    if (build_stem_fused_params(n, e->cfg.image_hw, e->cfg.image_hw, 8, e->stem_w_pad, e->x0, e->convs[0].y, &pl->stem_f)) return nullptr;
    pl->stem_fused = true;
  }
  for (size_t bi = 0; bi < e->blocks.size(); ++bi) {
    Block& b = e->blocks[bi];
    for (int j = 1; j < b.n_main; ++j) {
      Conv& cv = e->convs[b.main[j]];
      PmdConvShape s = cv.s;
      s.N = n;
      if (build_dgrad_params(s, cv.dy, e->wbf16 + cv.w_off, nullptr, nullptr, nullptr, nullptr, b.dz_mid[j - 1],
                             &pl->inner_dgrad[b.main[j]], true))
        return nullptr;
      if (resident_dgrad && conv_resident_supported(s)) {
        if (build_dgrad_resident_params(s, cv.dy, e->wbf16 + cv.w_off, nullptr, b.dz_mid[j - 1],
                                        &pl->inner_dgrad_r[b.main[j]]))
          return nullptr;
        pl->inner_resident[b.main[j]] = 1;
        if (conv_resident_red_enabled()) {
          Conv& pv = e->convs[b.main[j - 1]];
          attach_bn_reduce_resident(&pl->inner_dgrad_r[b.main[j]], pv.y, pv.scale, pv.shift, pv.mean, pv.invstd, pv.sums);
          pl->inner_reduced[b.main[j]] = 1;
        }
      } else {
        // dz_mid[j-1] is consumed only by the BatchNorm backward of conv j-1: let the dgrad epilogue do its reduction
        const int prev = b.main[j - 1];
        Conv& pv = e->convs[prev];
        const int rc = attach_bn_reduce(&pl->inner_dgrad[b.main[j]], s, pv.y, pv.scale, pv.shift, pv.mean, pv.invstd, pv.sums);
        if (rc < 0) return nullptr;
        pl->inner_reduced[b.main[j]] = static_cast<char>(rc);
      }
    }
    Conv& c0 = e->convs[b.main[0]];
    PmdConvShape s0 = c0.s;
    s0.N = n;
    if (b.sc >= 0 && e->convs[b.sc].s.stride == c0.s.stride) {
      Conv& cs = e->convs[b.sc];
      PmdConvShape ss = cs.s;
      ss.N = n;
      if (build_dgrad_params(s0, c0.dy, e->wbf16 + c0.w_off, &ss, cs.dy, e->wbf16 + cs.w_off, nullptr, b.dz_in,
                             &pl->block_dgrad[bi], true))
        return nullptr;
    } else if (b.sc >= 0) {
      // Bottleneck stage entry: conv0 is 1x1 stride 1 but the projection is 1x1 stride 2.  The shortcut's
      // dgrad fills dz_in first (zeros at the positions it skips), conv0's dgrad then adds onto it in place.
      Conv& cs = e->convs[b.sc];
      PmdConvShape ss = cs.s;
      ss.N = n;
      if (build_dgrad_params(ss, cs.dy, e->wbf16 + cs.w_off, nullptr, nullptr, nullptr, nullptr, b.dz_in,
                             &pl->sc_dgrad[bi], true))
        return nullptr;
      if (build_dgrad_params(s0, c0.dy, e->wbf16 + c0.w_off, nullptr, nullptr, nullptr, b.dz_in, b.dz_in,
                             &pl->block_dgrad[bi], true))
        return nullptr;
    } else {
      // identity shortcut: its gradient is dz_out under the block output's ReLU mask -- read straight from dz_out with
      // the ReLU bits when they exist (the BatchNorm backward then does not write the masked copy g_out)
      const bool res_block = conv_resident_block_dgrad_enabled() && conv_resident_supported(s0);
      const bool masked_addend = b.zbits != nullptr && (!res_block || conv_resident_red_enabled());
      if (build_dgrad_params(s0, c0.dy, e->wbf16 + c0.w_off, nullptr, nullptr, nullptr, masked_addend ? b.dz_out : b.g_out,
                             b.dz_in, &pl->block_dgrad[bi], true))
        return nullptr;
      if (masked_addend) pl->block_dgrad[bi].addend_bits = b.zbits;
      pl->g_out_unused[bi] = masked_addend ? 1 : 0;
      if (res_block) {
        if (build_dgrad_resident_params(s0, c0.dy, e->wbf16 + c0.w_off, masked_addend ? b.dz_out : b.g_out, b.dz_in,
                                        &pl->block_dgrad_r[bi]))
          return nullptr;
        if (masked_addend) pl->block_dgrad_r[bi].addend_bits = b.zbits;
        pl->block_resident[bi] = 1;
        if (bi == 0 && conv_resident_red_enabled()) {
          Conv& stem = e->convs[0];
          attach_bn_reduce_resident(&pl->block_dgrad_r[bi], stem.y, stem.scale, stem.shift, stem.mean, stem.invstd, stem.sums);
          pl->stem_reduced = true;
        }
      } else if (bi == 0) {
        // dz_stem is consumed only by the stem's BatchNorm backward (mask recomputed from y_stem): this dgrad's
        // epilogue (accumulator + identity-shortcut addend) masks it and accumulates the two channel sums
        Conv& stem = e->convs[0];
        const int rc = attach_bn_reduce(&pl->block_dgrad[bi], s0, stem.y, stem.scale, stem.shift, stem.mean, stem.invstd,
                                        stem.sums);
        if (rc < 0) return nullptr;
        pl->stem_reduced = rc == 1;
      }
    }
  }
  // The dgrad of block bi's first convolution produces dL/dz of block bi-1's output: let its epilogue mask that gradient
  // with the output's ReLU bits and reduce it for the output BatchNorm(s) of block bi-1 (no stand-alone reduce pass).
  // Not when two kernels write dz_in (Bottleneck stage entry with a strided projection) or the resident kernel runs.
  for (size_t bi = 1; bi < e->blocks.size(); ++bi) {
    Block& b = e->blocks[bi];
    Block& pb = e->blocks[bi - 1];
    const bool two_writers = b.sc >= 0 && e->convs[b.sc].s.stride != e->convs[b.main[0]].s.stride;
    // PMD_BLOCK_REDUCE: bit (bi - 1) enables the fusion for the output of block bi - 1 (default: see below)
    static const int sel = getenv("PMD_BLOCK_REDUCE") == nullptr ? 0 : atoi(getenv("PMD_BLOCK_REDUCE"));
    if (pb.zbits == nullptr || two_writers || pl->block_resident[bi] || ((sel >> (bi - 1)) & 1) == 0) continue;
    Conv& c0 = e->convs[b.main[0]];
    PmdConvShape s0 = c0.s;
    s0.N = n;
    Conv& l0 = e->convs[pb.main[pb.n_main - 1]];
    Conv* l1 = pb.sc >= 0 ? &e->convs[pb.sc] : nullptr;
    const int rc = attach_block_bn_reduce(&pl->block_dgrad[bi], s0, pb.zbits, l1 != nullptr ? 2 : 1, l0.y, l0.mean, l0.invstd,
                                          l0.sums, l1 != nullptr ? l1->y : nullptr, l1 != nullptr ? l1->mean : nullptr,
                                          l1 != nullptr ? l1->invstd : nullptr, l1 != nullptr ? l1->sums : nullptr);
    if (rc < 0) return nullptr;
    pl->out_reduced[bi - 1] = static_cast<char>(rc);
  }
  Plan* raw = pl.get();
  e->plans[n] = std::move(pl);
  return raw;
}

// ---------------------------------------------------------------------------------- step bodies
struct BucketTracker {
  PmdEngine* e;
  cudaStream_t main;
  std::vector<int> pending;
  int launched = 0;
  float lr;
  int done(int bkt) {
    if (--pending[bkt] != 0) return 0;
    // bucket complete: hand it to the comm stream
    // gradients of this bucket were produced on the main stream (BN, FC) and on the wgrad stream
    cudaEvent_t ev = e->events[launched % e->events.size()];
    PMD_CHECK_CUDA(cudaEventRecord(ev, main));
    PMD_CHECK_CUDA(cudaStreamWaitEvent(e->comm_stream, ev, 0));
    cudaEvent_t ev2 = e->wg_events[(e->wg_event_next++) % e->wg_events.size()];
    PMD_CHECK_CUDA(cudaEventRecord(ev2, e->wg_stream));
    PMD_CHECK_CUDA(cudaStreamWaitEvent(e->comm_stream, ev2, 0));
    PmdAllreduceSgd a{};
    a.param_off = e->param_off;
    a.grad_off = e->grad_off;
    a.mom_off = e->mom_off;
    a.bf16_off = e->bf16_off;
    a.first = e->bucket_start[bkt];
    a.n = e->bucket_numel[bkt];
    a.lr = lr;
    a.momentum = e->cfg.momentum;
    a.weight_decay = e->cfg.weight_decay;
    a.nesterov = e->cfg.nesterov;
    a.algo = e->bucket_algo[bkt];
    const bool local = e->sym.world == 1 || (e->comm_stub & 2) != 0;
    a.blocks = local ? 4 * e->sm_count : e->bucket_blocks[bkt];
    a.d_lr = e->d_lr;
    a.recv_off = e->bucket_recv[bkt];
    const int slot = 2 * static_cast<int>(e->convs.size()) + 8 + bkt * 128;
    E_CHECK(pmd_allreduce_sgd(local ? &e->sym_local : &e->sym, &a, slot, e->comm_stream));
    ++launched;
    return 0;
  }
};

// Describes conv `ci`'s BatchNorm for the fused finalize+apply kernel.
PmdBnFinalize bn_desc(PmdEngine* e, int ci) {
  Conv& cv = e->convs[ci];
  PmdBnFinalize f{};
  f.stats = cv.stats;
  f.gamma = e->param + cv.g_off;
  f.beta = e->param + cv.b_off;
  f.running_mean = cv.rm;
  f.running_var = cv.rv;
  f.num_batches = cv.nb;
  f.scale = cv.scale;
  f.shift = cv.shift;
  f.mean = cv.mean;
  f.invstd = cv.invstd;
  f.pivot = cv.pivot;
  f.pivot_new = cv.pivot_new;
  return f;
}

// Multi-rank SyncBatchNorm, forward: exchange conv `ci`'s [sum, sumsq] with the peers and finalise the statistics
// (one single-block push-exchange kernel, launched right behind the convolution ON ITS STREAM -- a projection
// shortcut's exchange therefore runs on the side stream, next to the main path's second convolution).
int exchange_forward(PmdEngine* e, int ci, long long rows, cudaStream_t st) {
  Conv& b = e->convs[ci];
  const PmdEngineConfig& c = e->cfg;
  return pmd_bn_finalize_push(&e->sym, ci, b.stats, b.recv_fwd_off, static_cast<float>(rows), b.s.K, e->param + b.g_off,
                              e->param + b.b_off, b.rm, b.rv, b.nb, c.bn_momentum, c.bn_eps, b.scale, b.shift, b.mean,
                              b.invstd, b.pivot, b.pivot_new, st);
}

// z = relu(BN_ci(y_ci) [+ res | + BN_cres(y_cres)]).  Training, one rank: the statistics are finalised inside the
// apply kernel.  Training, multi-rank SyncBatchNorm: exchange_forward() has finalised them, only the plain apply
// remains.  Evaluation uses the running statistics.
int bn_apply_step(PmdEngine* e, int ci, int cres, const bf16* res, bf16* z, long long rows, bool train, cudaStream_t st,
                  unsigned char* bits = nullptr) {
  if (!train) bits = nullptr;
  Conv& cv = e->convs[ci];
  const PmdEngineConfig& c = e->cfg;
  if (train && !sync_bn_active(e)) {
    const PmdBnFinalize f = bn_desc(e, ci);
    PmdBnFinalize fr{};
    if (cres >= 0) fr = bn_desc(e, cres);
    return pmd_bn_finalize_apply_bits(&f, cres >= 0 ? &fr : nullptr, static_cast<float>(rows), c.bn_momentum, c.bn_eps, cv.y,
                                      cres >= 0 ? e->convs[cres].y : res, 1, z, bits, rows, cv.s.K, st);
  }
  if (!train) {
    E_CHECK(pmd_bn_eval_affine(cv.s.K, e->param + cv.g_off, e->param + cv.b_off, cv.rm, cv.rv, c.bn_eps, cv.scale, cv.shift, st));
    if (cres >= 0) {
      Conv& cr = e->convs[cres];
      E_CHECK(pmd_bn_eval_affine(cr.s.K, e->param + cr.g_off, e->param + cr.b_off, cr.rm, cr.rv, c.bn_eps, cr.scale, cr.shift, st));
    }
  }
  if (cres >= 0) {
    Conv& cr = e->convs[cres];
    return pmd_bn_apply_bits(cv.y, cv.scale, cv.shift, cr.y, cr.scale, cr.shift, 1, z, bits, rows, cv.s.K, st);
  }
  return pmd_bn_apply_bits(cv.y, cv.scale, cv.shift, res, nullptr, nullptr, 1, z, bits, rows, cv.s.K, st);
}

// BatchNorm backward of conv `ci` (and, for a projection block's output, of the shortcut conv `cj`
// in the same two passes).  dz = gradient wrt the activation that followed; the ReLU mask comes from z
// (block outputs, where a residual was added before the ReLU) or is recomputed from y (z == nullptr).
// already_reduced: the dgrad that produced dz accumulated the sums (and masked dz) in its epilogue.
int bn_backward(PmdEngine* e, int ci, int cj, const bf16* dz, const bf16* z, long long rows, bf16* g_out,
                cudaStream_t st, bool already_reduced = false, const unsigned char* zbits = nullptr) {
  const bool sync = sync_bn_active(e);
  const int nb = cj >= 0 ? 2 : 1;
  PmdBnBranch br[2];
  const int idx[2] = {ci, cj};
  for (int b = 0; b < nb; ++b) {
    Conv& cv = e->convs[idx[b]];
    br[b] = PmdBnBranch{};
    br[b].y = cv.y;
    br[b].mean = cv.mean;
    br[b].invstd = cv.invstd;
    br[b].gamma = e->param + cv.g_off;
    br[b].sums = cv.sums;
    br[b].totals = sync ? cv.sums_total : nullptr;
    br[b].dy = cv.dy;
    br[b].dgamma = e->grad + cv.g_off;
    br[b].dbeta = e->grad + cv.b_off;
    br[b].pivot = cv.pivot;
    br[b].pivot_new = cv.pivot_new;
  }
  Conv& c0 = e->convs[ci];
  int mask = z != nullptr ? 1 : 2;
  if (zbits != nullptr) {        // the forward apply recorded which outputs passed the ReLU: 1/16 of the bytes of z
    mask = 3;
    z = reinterpret_cast<const bf16*>(zbits);
  }
  if (!already_reduced) E_CHECK(pmd_bn_bwd_reduce_multi(dz, z, c0.scale, c0.shift, mask, br, nb, rows, c0.s.K, st));
  if (sync) {
    // one push-exchange kernel; a projection block's two [2C] vectors leave in the same NVLink trip
    PmdSmallExchange x[2];
    for (int b = 0; b < nb; ++b) {
      Conv& cv = e->convs[idx[b]];
      x[b].local = cv.sums;
      x[b].n = 2 * cv.s.K;
      x[b].recv_off = cv.recv_bwd_off;
      x[b].total = cv.sums_total;
      x[b].slot = static_cast<int>(e->convs.size()) + idx[b];
    }
    E_CHECK(pmd_small_allreduce_push(&e->sym, x, nb, st));
  }
  const float count = static_cast<float>(rows) * (sync ? e->sym.world : 1);
  return pmd_bn_bwd_apply_multi(dz, z, c0.scale, c0.shift, mask, br, nb, count, g_out, rows, c0.s.K, st);
}

int run_forward(PmdEngine* e, Plan* pl, bool train, cudaStream_t st) {
  const PmdEngineConfig& c = e->cfg;
  const int n = pl->n;
  if (pl->stem_fused) {
    // gather / crop / flip / normalise + stem convolution + its statistics: one persistent kernel
    StemFusedParams sp = pl->stem_f;
    if (e->ds_on) {
      const int* idx = e->in_index + e->in_slot * (4 + c.max_batch);
      sp.ds = e->d_ds;
      sp.hdr = idx;
      sp.idx = idx + 4;
      sp.labels_out = e->in_labels;
      sp.valid_out = e->in_valid;
    } else {
      sp.u8 = e->in_u8;
      sp.aug = e->in_aug;
    }
    sp.x0 = train ? e->x0 : nullptr;      // only the weight gradient reads the im2col rows
    sp.stats = train ? e->convs[0].stats : nullptr;
    sp.stat_pivot = e->convs[0].pivot;
    sp.w_raw = e->wbf16 + e->convs[0].w_off;      // the optimizer's bf16 copy; padded inside the kernel
    sp.w_rsc = e->convs[0].real_rsc;
    E_CHECK(launch_stem_fused(sp, st));
  } else if (e->ds_on) {
    const int* idx = e->in_index + e->in_slot * (4 + c.max_batch);
    E_CHECK(pmd_augment_im2col_gather(e->d_ds, idx, idx + 4, n, c.image_hw, c.image_hw, 8, e->x0, e->in_labels, e->in_valid,
                                      st));
  } else
    E_CHECK(pmd_augment_im2col(e->in_u8, e->in_aug, n, c.image_hw, c.image_hw, 8, e->x0, st));
  auto conv_fwd = [&](int ci, cudaStream_t st) -> int {
    Conv& cv = e->convs[ci];
    if (cv.tc && pl->conv[ci].resident) {
      E_CHECK(launch_conv_resident(train ? pl->conv[ci].fprop_r : pl->conv[ci].fprop_eval_r, st));
    } else if (cv.tc) {
      E_CHECK(launch_conv_gemm(train ? pl->conv[ci].fprop : pl->conv[ci].fprop_eval, st));
    } else {
      PmdConvShape s = cv.s;
      s.N = n;
      E_CHECK(pmd_conv_fprop_direct(&s, cv.x_cstride, cv.x, e->wbf16 + cv.w_off, cv.y, st));
      if (train) E_CHECK(pmd_bn_stats_pivot(cv.y, 1LL * n * cv.Ho * cv.Wo, cv.s.K, cv.stats, cv.pivot, st));
    }
    if (train && sync_bn_active(e)) E_CHECK(exchange_forward(e, ci, 1LL * n * cv.Ho * cv.Wo, st));
    return 0;
  };
  if (!pl->stem_fused) E_CHECK(conv_fwd(0, st));
  else if (train && sync_bn_active(e)) E_CHECK(exchange_forward(e, 0, 1LL * n * e->convs[0].Ho * e->convs[0].Wo, st));
  {
    Conv& cv = e->convs[0];
    E_CHECK(bn_apply_step(e, 0, -1, nullptr, e->z_stem, 1LL * n * cv.Ho * cv.Wo, train, st));
  }
  for (Block& b : e->blocks) {
    cudaEvent_t sc_done = nullptr;
    if (b.sc >= 0) {
      // the projection shortcut only needs the block input: run it on the side stream next to the main path
      cudaEvent_t ev = e->wg_events[(e->wg_event_next++) % e->wg_events.size()];
      PMD_CHECK_CUDA(cudaEventRecord(ev, st));
      PMD_CHECK_CUDA(cudaStreamWaitEvent(e->wg_stream, ev, 0));
      E_CHECK(conv_fwd(b.sc, e->wg_stream));
      sc_done = e->wg_events[(e->wg_event_next++) % e->wg_events.size()];
      PMD_CHECK_CUDA(cudaEventRecord(sc_done, e->wg_stream));
    }
    for (int j = 0; j < b.n_main; ++j) {
      E_CHECK(conv_fwd(b.main[j], st));
      if (j < b.n_main - 1) {
        Conv& cv = e->convs[b.main[j]];
        E_CHECK(bn_apply_step(e, b.main[j], -1, nullptr, b.z_mid[j], 1LL * n * cv.Ho * cv.Wo, train, st));
      }
    }
    if (sc_done != nullptr) PMD_CHECK_CUDA(cudaStreamWaitEvent(st, sc_done, 0));
    Conv& last = e->convs[b.main[b.n_main - 1]];
    E_CHECK(bn_apply_step(e, b.main[b.n_main - 1], b.sc, b.sc >= 0 ? nullptr : b.z_in, b.z_out,
                          1LL * n * last.Ho * last.Wo, train, st, b.zbits));
  }
  Block& lb = e->blocks.back();
  E_CHECK(pmd_pool_fc_fwd(lb.z_out, n, lb.Hout * lb.Hout, lb.Cout, e->param + c.fc_w_off, e->param + c.fc_b_off,
                          c.num_classes, e->feat, e->logits, st));
  return 0;
}

int run_train(PmdEngine* e, Plan* pl, float lr, cudaStream_t st) {
  const PmdEngineConfig& c = e->cfg;
  const int n = pl->n;
  PMD_CHECK_CUDA(cudaMemsetAsync(e->grad, 0, c.slab_elems * 4, st));
  PMD_CHECK_CUDA(cudaMemsetAsync(e->hptr<char>(e->bnpart_off), 0, e->bnpart_bytes, st));
  E_CHECK(run_forward(e, pl, true, st));
  BucketTracker bt{e, st, e->bucket_params, 0, lr};
  Block& lb = e->blocks.back();
  // (head forward / loss / backward stay three launches: one kernel per image doing all three was 6 us per step slower)
  E_CHECK(pmd_ce_fwd_bwd(e->logits, e->in_labels, nullptr, n, c.num_classes, 1.0f / n, e->dlogits, e->metrics, st));
  E_CHECK(pmd_pool_fc_bwd(e->dlogits, e->feat, e->param + c.fc_w_off, n, lb.Hout * lb.Hout, lb.Cout, c.num_classes,
                          e->grad + c.fc_w_off, e->grad + c.fc_b_off, lb.dz_out, st));
  E_CHECK(bt.done(c.fc_b_bucket));
  E_CHECK(bt.done(c.fc_w_bucket));
  auto bn_done = [&](int ci) -> int {
    E_CHECK(bt.done(e->convs[ci].g_bkt));
    return bt.done(e->convs[ci].b_bkt);
  };
  // Weight gradients are off the critical path (nothing in backward consumes them), so they run on a
  // side stream: the L2-bound wgrad GEMMs overlap the DRAM-bound BatchNorm passes of the layer below.
  // dy_ready() marks the point on the main stream where a convolution's dy exists; wgrad(ci, ev) runs behind that
  // mark on the side stream.  The dgrad of the same layer is enqueued BETWEEN the two calls: it is on the critical
  // path, so its CTAs should reach the SMs first (PMD_DGRAD_FIRST=0 restores wgrad-first submission).
  static const bool dgrad_first = getenv("PMD_DGRAD_FIRST") == nullptr || atoi(getenv("PMD_DGRAD_FIRST")) != 0;
  cudaEvent_t ev_dy = nullptr;
  auto dy_ready = [&]() -> int {
    ev_dy = e->wg_events[(e->wg_event_next++) % e->wg_events.size()];
    PMD_CHECK_CUDA(cudaEventRecord(ev_dy, st));
    return 0;
  };
  auto wgrad = [&](int ci) -> int {
    Conv& cv = e->convs[ci];
    PMD_CHECK_CUDA(cudaStreamWaitEvent(e->wg_stream, ev_dy, 0));
    if (cv.tc) {
      E_CHECK(launch_wgrad(pl->conv[ci].wgrad, e->wg_stream));
    } else {
      PmdConvShape s = cv.s;
      s.N = n;
      E_CHECK(pmd_conv_wgrad_direct(&s, cv.x_cstride, cv.x, cv.dy, e->grad + cv.w_off, e->wg_stream));
    }
    return 0;
  };
  // A weight bucket may be handed to the fused all-reduce + SGD only once every reader of the CURRENT weights has
  // been enqueued on the main stream: the update rewrites the bf16 working copy in place, and the dgrad of the same
  // convolution (launched after its wgrad) still reads it.  done() records the main-stream event, so calling it
  // after the dgrad launch orders the update behind that kernel.
  auto w_done = [&](int ci) -> int { return bt.done(e->convs[ci].w_bkt); };
  {   // the wgrad stream joins the step after the gradient slab was cleared
    cudaEvent_t ev = e->wg_events[(e->wg_event_next++) % e->wg_events.size()];
    PMD_CHECK_CUDA(cudaEventRecord(ev, st));
    PMD_CHECK_CUDA(cudaStreamWaitEvent(e->wg_stream, ev, 0));
  }
  for (int bi = static_cast<int>(e->blocks.size()) - 1; bi >= 0; --bi) {
    Block& b = e->blocks[bi];
    const int last = b.main[b.n_main - 1];
    const long long rows_out = 1LL * n * b.Hout * b.Hout;
    E_CHECK(bn_backward(e, last, b.sc, b.dz_out, b.z_out, rows_out, pl->g_out_unused[bi] ? nullptr : b.g_out, st,
                        pl->out_reduced[bi] != 0, b.zbits));
    E_CHECK(bn_done(last));
    if (b.sc >= 0) E_CHECK(bn_done(b.sc));
    for (int j = b.n_main - 1; j >= 1; --j) {
      const int ci = b.main[j], prev = b.main[j - 1];
      E_CHECK(dy_ready());
      if (!dgrad_first) E_CHECK(wgrad(ci));
      if (pl->inner_resident[ci])
        E_CHECK(launch_conv_resident(pl->inner_dgrad_r[ci], st));
      else
        E_CHECK(launch_conv_gemm(pl->inner_dgrad[ci], st));
      if (dgrad_first) E_CHECK(wgrad(ci));
      E_CHECK(w_done(ci));
      Conv& pv = e->convs[prev];
      E_CHECK(bn_backward(e, prev, -1, b.dz_mid[j - 1], nullptr, 1LL * n * pv.Ho * pv.Wo, nullptr, st,
                          pl->inner_reduced[ci] != 0));
      E_CHECK(bn_done(prev));
    }
    E_CHECK(dy_ready());
    if (!dgrad_first) {
      E_CHECK(wgrad(b.main[0]));
      if (b.sc >= 0) E_CHECK(wgrad(b.sc));
    }
    if (b.sc >= 0 && e->convs[b.sc].s.stride != e->convs[b.main[0]].s.stride)
      E_CHECK(launch_conv_gemm(pl->sc_dgrad[bi], st));
    if (pl->block_resident[bi])
      E_CHECK(launch_conv_resident(pl->block_dgrad_r[bi], st));
    else
      E_CHECK(launch_conv_gemm(pl->block_dgrad[bi], st));
    if (dgrad_first) {
      E_CHECK(wgrad(b.main[0]));
      if (b.sc >= 0) E_CHECK(wgrad(b.sc));
    }
    E_CHECK(w_done(b.main[0]));
    if (b.sc >= 0) E_CHECK(w_done(b.sc));
  }
  {
    Conv& cv = e->convs[0];
    E_CHECK(bn_backward(e, 0, -1, e->dz_stem, nullptr, 1LL * n * cv.Ho * cv.Wo, nullptr, st, pl->stem_reduced));
    E_CHECK(bn_done(0));
    E_CHECK(dy_ready());
    E_CHECK(wgrad(0));
    E_CHECK(w_done(0));       // the stem has no dgrad
  }
  PMD_REQUIRE(bt.launched == c.n_buckets, "internal: not every gradient bucket was launched");
  PMD_CHECK_CUDA(cudaEventRecord(e->ev_join, e->comm_stream));
  PMD_CHECK_CUDA(cudaStreamWaitEvent(st, e->ev_join, 0));
  {
    cudaEvent_t ev = e->wg_events[(e->wg_event_next++) % e->wg_events.size()];
    PMD_CHECK_CUDA(cudaEventRecord(ev, e->wg_stream));
    PMD_CHECK_CUDA(cudaStreamWaitEvent(st, ev, 0));
  }
  if (!pl->stem_fused)      // the two-kernel stem reads a padded copy; the fused kernel pads in shared memory
    E_CHECK(pmd_pad_stem_weight(e->wbf16 + e->convs[0].w_off, e->stem_w_pad, 64, e->convs[0].real_rsc, st));
  return 0;
}

int run_eval(PmdEngine* e, Plan* pl, cudaStream_t st) {
  E_CHECK(run_forward(e, pl, false, st));
  return pmd_ce_fwd_bwd(e->logits, e->in_labels, e->in_valid, pl->n, e->cfg.num_classes, 0.f, nullptr, e->metrics, st);
}

int enter(PmdEngine* e, cudaStream_t user);
int leave(PmdEngine* e, cudaStream_t user);

// Eager (first call per batch size, or graphs off): the body runs on the engine's stream `st`, bracketed by events that
// order it after the caller's pending work and the caller's later work after it.  Replay: the graph was captured on
// `st` but is LAUNCHED ON THE CALLER'S STREAM -- its input copy, the step and the metric read-back then sit in one
// stream and no cross-stream event hop separates them (each hop costs microseconds of idle GPU per step).
template <typename Body>
int run_or_replay(PmdEngine* e, cudaGraphExec_t* graph, bool* warm, int* launches, cudaStream_t st, cudaStream_t user,
                  Body body) {
  if (!e->cfg.use_graph || !*warm) {
    E_CHECK(enter(e, user));
    // first call for this batch size runs eagerly: surfaces errors, sets function attributes, counts launches
    const unsigned long long before = g_launch_count;
    E_CHECK(body());
    *launches = static_cast<int>(g_launch_count - before);
    *warm = true;
    return leave(e, user);
  }
  if (*graph == nullptr) {
    cudaGraph_t g = nullptr;
    PMD_CHECK_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const unsigned long long before = g_launch_count;
    const int rc = body();
    g_launch_count = before;   // capture records nodes, nothing ran
    const cudaError_t ce = cudaStreamEndCapture(st, &g);
    if (rc != 0) {
      if (g != nullptr) cudaGraphDestroy(g);
      return -1;
    }
    PMD_CHECK_CUDA(ce);
    PMD_CHECK_CUDA(cudaGraphInstantiate(graph, g, 0));
    cudaGraphDestroy(g);
    PMD_CHECK_CUDA(cudaGraphUpload(*graph, user));   // device-side set-up of the graph now, not inside its first launch
  }
  PMD_CHECK_CUDA(cudaGraphLaunch(*graph, user));
  count_launch(*launches);
  return 0;
}

}  // namespace

extern "C" PMD_API int pmd_engine_create(const PmdEngineConfig* cfg, PmdComm* comm, PmdEngine** out) {
  PMD_REQUIRE(cfg != nullptr && comm != nullptr, "engine needs a config and a communicator");
  PmdEngine* e = new PmdEngine();
  *out = e;
  e->cfg = *cfg;
  e->comm = comm;
  E_CHECK(pmd_comm_info(comm, &e->info));
  e->sym.rank = e->info.rank;
  e->sym.world = e->info.world;
  e->sym.d_bases = reinterpret_cast<const unsigned long long*>(e->info.d_bases);
  e->sym.mc_base = e->info.mc_base;
  cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, e->info.device);
  e->bucket_start.assign(cfg->bucket_start, cfg->bucket_start + cfg->n_buckets);
  e->bucket_numel.assign(cfg->bucket_numel, cfg->bucket_numel + cfg->n_buckets);
  e->bucket_algo.assign(cfg->n_buckets, cfg->algo);
  e->bucket_blocks.assign(cfg->n_buckets, cfg->comm_blocks);
  if (cfg->bucket_algo != nullptr) e->bucket_algo.assign(cfg->bucket_algo, cfg->bucket_algo + cfg->n_buckets);
  if (cfg->bucket_blocks != nullptr) e->bucket_blocks.assign(cfg->bucket_blocks, cfg->bucket_blocks + cfg->n_buckets);
  PMD_REQUIRE(cfg->n_buckets >= 1 && 2 * cfg->n_convs + 8 + cfg->n_buckets * 128 + 200 < PMD_MAX_SLOTS,
              "too many gradient buckets for the signal pad (raise --bucket-mb)");
  E_CHECK(build_topology(e));
  E_CHECK(allocate(e));
  e->bucket_recv.assign(cfg->n_buckets, 0);
  for (int b = 0; b < cfg->n_buckets; ++b) {
    if (e->bucket_algo[b] != PMD_ALGO_LL) continue;
    if (e->sym.world == 1 || e->bucket_numel[b] > PMD_LL_MAX_ELEMS) {      // not applicable: fall back
      e->bucket_algo[b] = PMD_ALGO_ONESHOT;
      continue;
    }
    const unsigned long long bytes = PMD_LL_RECV_BYTES(e->sym.world, e->bucket_numel[b]);
    E_CHECK(e->halloc(&e->bucket_recv[b], bytes));
    PMD_CHECK_CUDA(cudaMemset(e->hptr<char>(e->bucket_recv[b]), 0, bytes));
  }
  e->bucket_params.assign(cfg->n_buckets, 0);
  for (Conv& cv : e->convs) {
    e->bucket_params[cv.w_bkt]++;
    e->bucket_params[cv.g_bkt]++;
    e->bucket_params[cv.b_bkt]++;
  }
  e->bucket_params[cfg->fc_w_bucket]++;
  e->bucket_params[cfg->fc_b_bucket]++;
  // copy the offset tables: the caller's arrays need not outlive this call
  e->cfg.conv_w_off = e->cfg.bn_g_off = e->cfg.bn_b_off = nullptr;
  e->cfg.bucket_start = e->cfg.bucket_numel = nullptr;
  e->cfg.bucket_algo = e->cfg.bucket_blocks = nullptr;
  {
    // The main stream carries the step's critical path (forward, BatchNorm backward, dgrad); the comm stream's
    // kernels are short and other ranks wait on them; weight gradients are consumed only by the bucket update.
    // So pending wgrad CTAs yield to the other two streams when SM resources free up (PMD_STREAM_PRIORITY=0: off).
    int least = 0, greatest = 0;
    PMD_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    const bool prio = getenv("PMD_STREAM_PRIORITY") == nullptr || atoi(getenv("PMD_STREAM_PRIORITY")) != 0;
    const int hi = prio ? greatest : 0, lo = prio ? least : 0;
    PMD_CHECK_CUDA(cudaStreamCreateWithPriority(&e->comm_stream, cudaStreamNonBlocking, hi));
    PMD_CHECK_CUDA(cudaStreamCreateWithPriority(&e->main_stream, cudaStreamNonBlocking, hi));
    PMD_CHECK_CUDA(cudaStreamCreateWithPriority(&e->wg_stream, cudaStreamNonBlocking, lo));
  }
  e->wg_events.resize(4 * (cfg->n_convs + cfg->n_buckets) + 16);
  for (auto& ev : e->wg_events) PMD_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  PMD_CHECK_CUDA(cudaEventCreateWithFlags(&e->ev_in, cudaEventDisableTiming));
  PMD_CHECK_CUDA(cudaEventCreateWithFlags(&e->ev_out, cudaEventDisableTiming));
  e->events.resize(cfg->n_buckets + 1);
  for (auto& ev : e->events) PMD_CHECK_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  PMD_CHECK_CUDA(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
  PMD_CHECK_CUDA(cudaDeviceSynchronize());
  return 0;
}

extern "C" PMD_API int pmd_engine_destroy(PmdEngine* e) {
  if (e == nullptr) return 0;
  cudaDeviceSynchronize();
  for (auto& kv : e->plans) {
    if (kv.second->train_graph) cudaGraphExecDestroy(kv.second->train_graph);
    if (kv.second->train_graph_alt) cudaGraphExecDestroy(kv.second->train_graph_alt);
    if (kv.second->eval_graph) cudaGraphExecDestroy(kv.second->eval_graph);
  }
  for (void* p : e->owned) cudaFree(p);
  for (auto& ev : e->events) cudaEventDestroy(ev);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  if (e->comm_stream) cudaStreamDestroy(e->comm_stream);
  if (e->main_stream) cudaStreamDestroy(e->main_stream);
  if (e->wg_stream) cudaStreamDestroy(e->wg_stream);
  for (auto& ev : e->wg_events) cudaEventDestroy(ev);
  if (e->ev_in) cudaEventDestroy(e->ev_in);
  if (e->ev_out) cudaEventDestroy(e->ev_out);
  delete e;
  return 0;
}

extern "C" PMD_API int pmd_engine_buffers(PmdEngine* e, PmdEngineBuffers* o) {
  memset(o, 0, sizeof(*o));
  o->param = reinterpret_cast<unsigned long long>(e->param);
  o->grad = reinterpret_cast<unsigned long long>(e->grad);
  o->mom = reinterpret_cast<unsigned long long>(e->mom);
  o->bf16 = reinterpret_cast<unsigned long long>(e->wbf16);
  o->in_u8 = reinterpret_cast<unsigned long long>(e->in_u8);
  o->in_aug = reinterpret_cast<unsigned long long>(e->in_aug);
  o->in_labels = reinterpret_cast<unsigned long long>(e->in_labels);
  o->in_valid = reinterpret_cast<unsigned long long>(e->in_valid);
  o->in_index = reinterpret_cast<unsigned long long>(e->in_index);
  o->metrics_snap = reinterpret_cast<unsigned long long>(e->metrics_snap);
  o->logits = reinterpret_cast<unsigned long long>(e->logits);
  o->metrics = reinterpret_cast<unsigned long long>(e->metrics);
  o->slab_elems = e->cfg.slab_elems;
  o->activation_bytes = e->activation_bytes;
  o->heap_bytes_used = static_cast<long long>(e->bnpart_off + e->bnpart_bytes);
  return 0;
}

extern "C" PMD_API int pmd_engine_sync_params(PmdEngine* e, int broadcast_from_rank0, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (broadcast_from_rank0 && e->sym.world > 1) {
    E_CHECK(pmd_broadcast(&e->sym, e->param_off, static_cast<unsigned long long>(e->cfg.slab_elems) * 4, 0, 64,
                          PMD_MAX_SLOTS - 128, st));
    E_CHECK(pmd_broadcast(&e->sym, e->mom_off, static_cast<unsigned long long>(e->cfg.slab_elems) * 4, 0, 64,
                          PMD_MAX_SLOTS - 128, st));
  }
  E_CHECK(pmd_cast_f32_to_bf16(e->param, e->wbf16, e->cfg.slab_elems, st));
  E_CHECK(pmd_pad_stem_weight(e->wbf16 + e->convs[0].w_off, e->stem_w_pad, 64, e->convs[0].real_rsc, st));
  PMD_CHECK_CUDA(cudaStreamSynchronize(st));
  return 0;
}

// The step runs on the engine's own stream (the caller's may be the legacy default stream, which
// cannot be captured); events order it after the caller's pending work (the H2D copies of this
// step's inputs) and order the caller's later work (metric reads) after it.
namespace {
int enter(PmdEngine* e, cudaStream_t user) {
  PMD_CHECK_CUDA(cudaEventRecord(e->ev_in, user));
  PMD_CHECK_CUDA(cudaStreamWaitEvent(e->main_stream, e->ev_in, 0));
  return 0;
}
int leave(PmdEngine* e, cudaStream_t user) {
  PMD_CHECK_CUDA(cudaEventRecord(e->ev_out, e->main_stream));
  PMD_CHECK_CUDA(cudaStreamWaitEvent(user, e->ev_out, 0));
  return 0;
}
}  // namespace

extern "C" PMD_API int pmd_engine_train_step(PmdEngine* e, int n, float lr, void* stream) {
  return pmd_engine_train_step_slot(e, n, lr, 0, stream);
}

extern "C" PMD_API int pmd_engine_train_step_slot(PmdEngine* e, int n, float lr, int slot, void* stream) {
  PMD_REQUIRE(n >= 1 && n <= e->cfg.max_batch, "batch size outside [1, max_batch]");
  PMD_REQUIRE(slot == 0 || slot == 1, "input slot must be 0 or 1");
  PMD_REQUIRE(slot == 0 || e->ds_on, "input slot 1 exists for the device-resident dataset path only");
  e->in_slot = slot;
  cudaStream_t user = static_cast<cudaStream_t>(stream), st = e->main_stream;
  Plan* pl = get_plan(e, n);
  if (pl == nullptr) return -1;
  if (lr != e->last_lr) {
    // rare (LR milestones): a synchronous copy keeps it ordered with whichever stream runs the next step
    PMD_CHECK_CUDA(cudaDeviceSynchronize());
    PMD_CHECK_CUDA(cudaMemcpy(e->d_lr, &lr, sizeof(float), cudaMemcpyHostToDevice));
    e->last_lr = lr;
  }
  // one captured graph per input slot (they differ in the index vector the first kernel reads and in the metric
  // snapshot the last node writes)
  return run_or_replay(e, slot == 0 ? &pl->train_graph : &pl->train_graph_alt, &pl->train_warm, &pl->train_launches, st, user,
                       [&]() {
                         E_CHECK(run_train(e, pl, lr, st));
                         PMD_CHECK_CUDA(cudaMemcpyAsync(e->metrics_snap + 4 * slot, e->metrics, 4 * sizeof(double),
                                                        cudaMemcpyDeviceToDevice, st));
                         return 0;
                       });
}

extern "C" PMD_API int pmd_engine_eval_step(PmdEngine* e, int n, void* stream) {
  PMD_REQUIRE(n >= 1 && n <= e->cfg.max_batch, "batch size outside [1, max_batch]");
  e->in_slot = 0;
  cudaStream_t user = static_cast<cudaStream_t>(stream), st = e->main_stream;
  Plan* pl = get_plan(e, n);
  if (pl == nullptr) return -1;
  return run_or_replay(e, &pl->eval_graph, &pl->eval_warm, &pl->eval_launches, st, user,
                       [&]() { return run_eval(e, pl, st); });
}

extern "C" PMD_API int pmd_engine_read_metrics(PmdEngine* e, int reduce_over_ranks, int reset, double* out4,
                                               void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const double* src = e->metrics;
  if (reduce_over_ranks && e->sym.world > 1) {
    E_CHECK(pmd_metric_allreduce(&e->sym, PMD_MAX_SLOTS - 160, e->metrics_off, e->metrics_total, st));
    src = e->metrics_total;
  }
  PMD_CHECK_CUDA(cudaMemcpyAsync(out4, src, 4 * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (reset) PMD_CHECK_CUDA(cudaMemsetAsync(e->metrics, 0, 4 * sizeof(double), st));
  PMD_CHECK_CUDA(cudaStreamSynchronize(st));
  return 0;
}

static int drop_plans(PmdEngine* e) {
  PMD_CHECK_CUDA(cudaDeviceSynchronize());
  for (auto& kv : e->plans) {       // plans (tensor maps, captured graphs) bake in the pointers and switches
    if (kv.second->train_graph) cudaGraphExecDestroy(kv.second->train_graph);
    if (kv.second->train_graph_alt) cudaGraphExecDestroy(kv.second->train_graph_alt);
    if (kv.second->eval_graph) cudaGraphExecDestroy(kv.second->eval_graph);
  }
  e->plans.clear();
  return 0;
}

extern "C" PMD_API int pmd_engine_set_dataset(PmdEngine* e, const void* d_images_u8, const long long* d_labels,
                                              long long count) {
  PMD_REQUIRE(d_images_u8 == nullptr || (d_labels != nullptr && count > 0 && count < (1LL << 31)),
              "a resident dataset needs labels and 0 < count < 2^31");
  const bool on = d_images_u8 != nullptr;
  if (on != e->ds_on) E_CHECK(drop_plans(e));    // the step's first kernel differs; WHICH dataset is read at run time
  if (on) {
    const unsigned long long desc[2] = {reinterpret_cast<unsigned long long>(d_images_u8),
                                        reinterpret_cast<unsigned long long>(d_labels)};
    // twice per epoch (train <-> test): synchronous, so it is ordered with every stream that has run or will run steps
    PMD_CHECK_CUDA(cudaDeviceSynchronize());
    PMD_CHECK_CUDA(cudaMemcpy(e->d_ds, desc, sizeof(desc), cudaMemcpyHostToDevice));
  }
  e->ds_on = on;
  e->ds_count = on ? count : 0;
  return 0;
}

extern "C" PMD_API int pmd_engine_set_comm_stub(PmdEngine* e, int mask) {
  PMD_REQUIRE(mask >= 0 && mask <= 3, "comm stub mask: bit 0 = BatchNorm statistics local, bit 1 = gradients local");
  if (mask == e->comm_stub) return 0;
  E_CHECK(drop_plans(e));
  e->comm_stub = mask;
  // the statistics pivots must be identical on every rank whenever sums are exchanged; ranks diverge under a stub
  for (Conv& cv : e->convs) PMD_CHECK_CUDA(cudaMemset(cv.pivot, 0, 2 * sizeof(float) * cv.s.K));
  return 0;
}

extern "C" PMD_API int pmd_engine_launches_per_step(PmdEngine* e, int n) {
  auto it = e->plans.find(n);
  return it == e->plans.end() ? 0 : it->second->train_launches;
}
