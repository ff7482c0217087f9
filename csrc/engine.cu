// Native step executor: forward + backward of the reference CNN (and its widened / MLP variants) as
// one C call that only enqueues kernels - no host sync, no allocation - so a whole training step can
// be captured in a CUDA graph.
//
// Replaces the TF1 graph + Session of mnist_sync/model/model.py:17-112: ``sess.run(self.grads)``
// (mnist_sync/worker.py:30) is ddl_engine_train_step, ``sess.run(self.accuracy)`` (worker.py:72) is
// ddl_engine_eval.  Parameters and gradients live in the caller's flat arenas (one offset per
// variable v0..v13), activations in a workspace the engine allocates once.
#include "common.cuh"
#include "../include/ddl_b200.h"

#include <map>
#include <stdlib.h>
#include <new>
#include <vector>

namespace {

inline int cdiv(int a, int b) { return (a + b - 1) / b; }


struct FwdPlans {                  // forward plans for one batch size
  int B = 0;
  DdlTcPlan* conv[3] = {nullptr, nullptr, nullptr};    // conv2..conv4
  DdlPatchPlan* pconv[3] = {nullptr, nullptr, nullptr}; // ... or the patch-resident plan of the layer (tc_patch.cu)
  DdlTcPlan* fc[2] = {nullptr, nullptr};                // fc1, fc2
  int conv_split[3] = {1, 1, 1};
  bool fused_pool[3] = {false, false, false};           // conv plan's epilogue is bias + ReLU + pool
  void destroy() {
    for (auto& p : conv) { ddl_tc_plan_destroy(p); p = nullptr; }
    for (auto& p : pconv) { ddl_patch_plan_destroy(p); p = nullptr; }
    for (auto& p : fc) { ddl_tc_plan_destroy(p); p = nullptr; }
  }
};

}  // namespace

struct DdlEngine {
  DdlEngineCfg cfg;
  int nconv;                       // 4 (cnn) or 0 (mlp)
  int S[5];                        // spatial size at the input of conv1..4 and after the last pool
  int C[5];                        // channels
  int F[4];                        // fc sizes: in, h1, h2, classes
  int maxB;
  // workspace
  float* pooled[4] = {};           // p1..p4  [B, S[i+1], S[i+1], C[i+1]]
  uint8_t* idx[4] = {};
  float* raw = nullptr;            // raw conv sums, reused by conv2..4  [B, S, S, Cout] max
  float* dfull[4] = {};            // full-resolution gradients dY of conv1..conv4 [B, S, S, Cout]: one buffer
                                   // per layer, because layer l's wgrad (side stream) still reads its dY while
                                   // the main stream is already producing layer l-1's
  float* dpool[4] = {};            // gradient wrt pooled outputs p1..p4
  float* h1 = nullptr;             // fc1 output after ReLU + dropout [B, F1]
  float* h2 = nullptr;             // fc2 output after dropout        [B, F2]
  float* dh1 = nullptr;
  float* dh2 = nullptr;
  float* logits = nullptr;
  float* dlogits = nullptr;
  DdlStepState* eval_st = nullptr; // scratch step state for evaluation
  std::map<int, FwdPlans> train_fwd, eval_fwd;
  // backward plans (train batch only)
  DdlTcPlan* wg_conv[3] = {};
  DdlTcPlan* dg_conv[3] = {};
  DdlPatchPlan* pdg_conv0 = nullptr;           // conv2's dgrad as a patch-resident plan (replaces dg_conv[0] when set)
  int dg_split[3] = {1, 1, 1};
  bool dg_unpool[3] = {false, false, false};   // dgrad plan's epilogue is the pool + ReLU backward of the layer below
  bool w1_pooled = false;                      // dg_conv[0] keeps conv1's gradient in the pooled layout (see build_bwd)
  DdlTcPlan* wg_fc[2] = {};
  DdlTcPlan* dg_fc[2] = {};        // dg_fc[0]: fc1 -> dp4 (cnn only), dg_fc[1]: fc2 -> dh1
  std::vector<void*> allocs;
  // fork/join: weight-gradient kernels run on a side stream next to the dgrad chain (they only feed the
  // optimizer); captured into the step graph as parallel branches
  // backward-overlapped push (gradient bucketing): argument blocks of the bucket launched on the side stream
  DdlComm ov_comm;
  DdlSyncArgs ov_args[8];
  int ov_n = 0, ov_kind = 2;
  cudaStream_t side = nullptr;
  cudaStream_t side2 = nullptr;    // the overlapped push/update/pull bucket: must not delay the remaining wgrads
  cudaEvent_t ev_fork[8] = {};
  cudaEvent_t ev_join = nullptr, ev_fork2 = nullptr, ev_join2 = nullptr, ev_fc_w = nullptr, ev_fc_m = nullptr;
  // one stream per conv weight-gradient kernel: they only depend on their own dY, and serialising them on one
  // stream made that stream the critical path of the backward (tools/timeline.py)
  cudaStream_t wside[2] = {nullptr, nullptr};       // conv3 / conv2 wgrads (conv4 + FC wgrads stay on `side`)
  cudaEvent_t ev_wjoin[2] = {nullptr, nullptr}, ev_w0 = nullptr, ev_main2 = nullptr, ev_pre0 = nullptr, ev_pre1 = nullptr;
  bool own_advance = false;                         // the engine launches ddl_step_advance itself (side stream)
  float adv_lr = 0.f, adv_b1 = 0.f, adv_b2 = 0.f;
  int adv_opt = 1;

  const float* P(int v) const { return cfg.params + cfg.var_offset[v]; }
  float* G(int v) const { return cfg.grads + cfg.var_offset[v]; }
};

namespace {

int alloc(DdlEngine* e, void** p, size_t bytes) {
  DDL_CHECK_CUDA(cudaMalloc(p, bytes < 256 ? 256 : bytes));
  e->allocs.push_back(*p);
  return 0;
}

// Tile-shape policy (measured on B200, profiles/tc_issue_latency.md): the TMA engine feeds one SM at
// ~80 B/clk and every k-iteration costs ~600 cycles of barrier + issue overhead, so (1) use the widest
// N tile the layer allows - operand bytes per MAC fall as 1/N - and (2) get the CTA count from a
// split over the reduction (taps x channel blocks), reduced with red.add, not from narrow N tiles.
void pick_tiling(int pix_tiles, int nout, int kiters, bool allow_split, int* ntile, int* ksplit) {
  int nt = 32;
  for (int c : {256, 128, 64, 32}) if (nout % c == 0) { nt = c; break; }
  const int ctas = pix_tiles * cdiv(nout, nt);
  int ks = 1;
  if (allow_split && ctas < 100) {
    ks = (140 + ctas / 2) / ctas;
    if (ks > kiters / 4) ks = kiters / 4 > 0 ? kiters / 4 : 1;
    if (ks > 16) ks = 16;
    if (ks < 1) ks = 1;
  }
  *ntile = nt; *ksplit = ks;
}

// conv plans with a staged (pool / un-pool) epilogue.  Measured (tools/pool_perf.py, profiles/pool_fusion.md): the
// fastest (N tile, cluster split) is the one whose grid comes closest to ~208 CTAs (two resident per SM on 104 SMs
// worth of clusters) from below, ties going to the wider cluster, then the wider tile.  Larger batches (no choice
// fits): widest tile, no split.  DDL_POOL_NT / DDL_POOL_CL override both (perf experiments).
void pick_staged_tiling(int pix_tiles, int nout, int kiters, int* ntile, int* cluster) {
  int best_nt = 0, best_cl = 1, best_ctas = -1;
  for (int nt : {256, 128, 64, 32}) {
    if (nout % nt) continue;
    if (best_nt == 0) best_nt = nt;                       // fallback: widest legal tile, unsplit
    for (int cl : {8, 4, 2, 1}) {
      if (kiters / cl < 4 || (nt / cl) % 4 || nt / cl < 8) continue;
      const int ctas = pix_tiles * (nout / nt) * cl;
      if (ctas > 210) continue;
      if (ctas > best_ctas) { best_ctas = ctas; best_nt = nt; best_cl = cl; }
    }
  }
  if (best_nt == 0) best_nt = 32;
  const int lab_nt = ddl_lab_int("DDL_POOL_NT", 0), lab_cl = ddl_lab_int("DDL_POOL_CL", 0);
  if (lab_nt > 0 && nout % lab_nt == 0) best_nt = lab_nt;
  if (lab_cl > 0) best_cl = lab_cl;
  *ntile = best_nt; *cluster = best_cl;
}

bool fuse_pool_enabled() {
  static const bool on = DDL_LAB_ENV("DDL_NO_POOL_FUSE") == nullptr;
  return on;
}

// FC layers (one pixel tile, M <= 128 rows): the work is streaming the weight matrix once, so spread it
// over ~a wave of SMs with narrow N tiles and a cluster split over K (the fused epilogue survives).
void pick_fc_tiling(int nout, int kiters, int* ntile, int* ksplit) {
  const int nt = 32;
  const int ctas = cdiv(nout, nt);
  int ks = ctas >= 120 ? 1 : (128 + ctas - 1) / ctas;
  if (ks > 8) ks = 8;
  while (ks > 1 && kiters / ks < 2) --ks;
  *ntile = nt; *ksplit = ks;
}

int pix_tiles_of(DdlTcPlan* pl) {
  int info[8];
  ddl_tc_plan_info(pl, info);
  return info[0];
}

// Forward plans for batch size B.  train=true adds dropout to the FC epilogues.
int build_fwd(DdlEngine* e, int B, bool train, FwdPlans* out) {
  const DdlEngineCfg& c = e->cfg;
  out->B = B;
  int v = 2;                                       // v0/v1 are conv1 (SIMT)
  for (int l = 1; l < e->nconv; ++l, v += 2) {     // conv2..conv4: raw sums; bias+ReLU+pool follow
    const int S = e->S[l], cin = e->C[l], cout = e->C[l + 1];
    // probe the pixel tiling with a throw-away plan to size ntile / ksplit
    DdlTcPlan* probe = ddl_tc_plan_act(0, e->pooled[l - 1], e->P(v), e->raw, B, S, S, cin, cout, 5, 32, 1, 0,
                                       nullptr, nullptr, 1.f, nullptr, 0, 1.f, nullptr, nullptr);
    if (!probe) return -1;
    int nt, ks;
    if (fuse_pool_enabled() && ddl_patch_conv_preferred(S, S, cin, cout)) {
      // large planes (conv2): one CTA per image, the padded activation patch resident in shared memory, every filter tap
      // a shifted descriptor - the activation crosses L2 once instead of 25 times (csrc/tc_patch.cu)
      ddl_tc_plan_destroy(probe);
      out->pconv[l - 1] = ddl_patch_plan_create(0, e->pooled[l - 1], e->P(v), e->P(v + 1), e->pooled[l], train ? e->idx[l] : nullptr,
                                                nullptr, B, S, S, cin, cout);
      if (!out->pconv[l - 1]) return -1;
      out->fused_pool[l - 1] = true;
      continue;
    }
    if (fuse_pool_enabled()) {
      // conv + bias + ReLU + pool in one kernel: no raw buffer, no zeroing, no pool launch
      ddl_tc_plan_destroy(probe);
      probe = ddl_tc_plan_conv_pool(e->pooled[l - 1], e->P(v), e->P(v + 1), e->pooled[l], nullptr, B, S, S, cin, cout, 5, 32, 1);
      if (!probe) return -1;
      pick_staged_tiling(pix_tiles_of(probe), cout, 25 * cdiv(cin, 32), &nt, &ks);
      ddl_tc_plan_destroy(probe);
      out->conv[l - 1] = ddl_tc_plan_conv_pool(e->pooled[l - 1], e->P(v), e->P(v + 1), e->pooled[l], train ? e->idx[l] : nullptr,
                                               B, S, S, cin, cout, 5, nt, ks);
      if (!out->conv[l - 1]) return -1;
      out->fused_pool[l - 1] = true;
      continue;
    }
    pick_tiling(pix_tiles_of(probe), cout, 25 * cdiv(cin, 32), true, &nt, &ks);
    ddl_tc_plan_destroy(probe);
    out->conv[l - 1] = ddl_tc_plan_act(0, e->pooled[l - 1], e->P(v), e->raw, B, S, S, cin, cout, 5, nt, ks, 0,
                                       nullptr, nullptr, 1.f, nullptr, 0, 1.f, nullptr, nullptr);
    if (!out->conv[l - 1]) return -1;
    out->conv_split[l - 1] = ddl_tc_plan_needs_zero(out->conv[l - 1]) ? 2 : 1;   // > 1: raw must be zeroed
  }
  const float* fc_in = e->nconv ? e->pooled[3] : nullptr;   // mlp: x itself (bound at run time)
  const int vf = 2 * e->nconv;
  const float kp = train ? c.keep_prob : 1.f;
  const int drop = (train && c.keep_prob < 1.f) ? 4 : 0;
  int fnt, fks;
  if (e->nconv) {
    pick_fc_tiling(e->F[1], cdiv(e->F[0], 32), &fnt, &fks);
    if (B > 128) fks = 1;                      // evaluation batches already fill the machine with pixel tiles
    out->fc[0] = ddl_tc_plan_act(0, fc_in, (train && c.fc_w_src[0]) ? c.fc_w_src[0] : e->P(vf), e->h1, B, 1, 1, e->F[0], e->F[1], 1, fnt, fks, 1 | 2 | drop,
                                 e->P(vf + 1), nullptr, 1.f, train ? c.st : nullptr, 0, kp, nullptr, train && c.fc_w_src[0] ? const_cast<float*>(e->P(vf)) : nullptr);
    if (!out->fc[0]) return -1;
  }
  pick_fc_tiling(e->F[2], cdiv(e->F[1], 32), &fnt, &fks);
  if (B > 128) fks = 1;
  out->fc[1] = ddl_tc_plan_act(0, e->h1, (train && c.fc_w_src[1]) ? c.fc_w_src[1] : e->P(vf + 2), e->h2, B, 1, 1, e->F[1], e->F[2], 1, fnt, fks, 1 | drop,
                               e->P(vf + 3), nullptr, 1.f, train ? c.st : nullptr, 1, kp, nullptr, train && c.fc_w_src[1] ? const_cast<float*>(e->P(vf + 2)) : nullptr);
  if (!out->fc[1]) return -1;
  return 0;
}

int build_bwd(DdlEngine* e) {
  const DdlEngineCfg& c = e->cfg;
  const int B = c.batch;
  const int vf = 2 * e->nconv;
  const float inv_keep = c.keep_prob < 1.f ? 1.f / c.keep_prob : 1.f;
  // fc2: dW2 = h1^T dh2 ; dh1 = (dh2 W2^T) * (h1 > 0) / keep
  e->wg_fc[1] = ddl_tc_plan_wgrad(e->dh2, e->h1, e->G(vf + 2), B, 1, 1, e->F[2], e->F[1], 1, 64, 1);
  // its epilogue also reduces the columns of dh1 into the fc1 bias gradient
  int fnt, fks;
  pick_fc_tiling(e->F[1], cdiv(e->F[2], 32), &fnt, &fks);
  e->dg_fc[1] = ddl_tc_plan_act(1, e->dh2, e->P(vf + 2), e->dh1, B, 1, 1, e->F[2], e->F[1], 1, fnt, fks, 8, nullptr,
                                e->h1, inv_keep, nullptr, 0, 1.f, e->G(vf + 1), nullptr);
  if (!e->wg_fc[1] || !e->dg_fc[1]) return -1;
  if (e->nconv) {
    e->wg_fc[0] = ddl_tc_plan_wgrad(e->dh1, e->pooled[3], e->G(vf), B, 1, 1, e->F[1], e->F[0], 1, 64, 1);
    pick_fc_tiling(e->F[0], cdiv(e->F[1], 32), &fnt, &fks);
    e->dg_fc[0] = ddl_tc_plan_act(1, e->dh1, e->P(vf), e->dpool[3], B, 1, 1, e->F[1], e->F[0], 1, fnt, fks, 0, nullptr,
                                  nullptr, 1.f, nullptr, 0, 1.f, nullptr, nullptr);
    if (!e->wg_fc[0] || !e->dg_fc[0]) return -1;
  }
  int v = 2;
  for (int l = 1; l < e->nconv; ++l, v += 2) {
    const int S = e->S[l], cin = e->C[l], cout = e->C[l + 1];
    const int nt_w = cin >= 128 ? 128 : (cin >= 64 ? 64 : 32);
    e->wg_conv[l - 1] = ddl_tc_plan_wgrad(e->dfull[l], e->pooled[l - 1], e->G(v), B, S, S, cout, cin, 5, nt_w, 0);
    if (!e->wg_conv[l - 1]) return -1;
    // dgrad: dp_{l-1}[pix, cin] = sum_{tap, cout} dY[pix - off, cout] * W[tap][cin][cout]
    DdlTcPlan* probe2 = ddl_tc_plan_act(1, e->dfull[l], e->P(v), e->dpool[l - 1], B, S, S, cout, cin, 5, 32, 1, 0,
                                        nullptr, nullptr, 1.f, nullptr, 0, 1.f, nullptr, nullptr);
    if (!probe2) return -1;
    int nt, ks;
    if (fuse_pool_enabled()) {
      // dgrad + un-pool + ReLU mask + bias gradient of layer l-1 in one kernel: writes dfull[l-1] directly
      pick_staged_tiling(pix_tiles_of(probe2), cin, 25 * cdiv(cout, 32), &nt, &ks);
      ddl_tc_plan_destroy(probe2);
      const int US = e->S[l - 1];
      if (l == 1 && e->C[0] == 1 && e->S[0] == 28 && cin % 32 == 0) {
        // conv2's dgrad: conv1 has one input channel and no data gradient, so the un-pooled plane is never built - the
        // epilogue keeps the gradient in the pooled layout (ReLU-dead windows zeroed, conv1's bias gradient summed) and
        // ddl_conv1_wgrad_pooled reads value + argmax slot per window (4x fewer bytes, no 75%-zero tensor)
        e->dg_conv[0] = ddl_tc_plan_dgrad_poolgrad(e->dfull[1], e->P(v), e->idx[0], e->dpool[0], e->G(1), B, S, S, cout, cin,
                                                   5, nt, ks);
        e->w1_pooled = e->dg_conv[0] != nullptr;
        // The patch-resident dgrad is faster alone (29 vs 35 us) but slower inside the step (0.226 vs 0.217 ms): it runs
        // beside conv2's wgrad, whose CTAs take whole SMs; one patch CTA per image also needs a whole SM each
        // (208 KB), so 12 of the 100 wait for a second wave, while the tap-by-tap plan's 100 KB CTAs pack two per SM
        // into what is free.  Kept for DDL_PATCH_DGRAD=1 experiments.
        static const bool patch_dgrad = DDL_LAB_ENV("DDL_PATCH_DGRAD") != nullptr;
        if (patch_dgrad && e->w1_pooled && ddl_patch_conv_preferred(S, S, cout, cin)) {
          e->pdg_conv0 = ddl_patch_plan_create(1, e->dfull[1], e->P(v), nullptr, e->dpool[0], e->idx[0], e->G(1), B, S, S, cout, cin);
          if (!e->pdg_conv0) return -1;
        }
      } else {
        e->dg_conv[l - 1] = ddl_tc_plan_dgrad_unpool(e->dfull[l], e->P(v), e->idx[l - 1], e->dfull[l - 1],
                                                     l - 1 >= 1 ? e->G(v - 1) : nullptr, B, S, S, US, US, cout, cin, 5, nt, ks);
      }
      if (!e->dg_conv[l - 1]) return -1;
      e->dg_unpool[l - 1] = true;
      continue;
    }
    pick_tiling(pix_tiles_of(probe2), cin, 25 * cdiv(cout, 32), true, &nt, &ks);
    ddl_tc_plan_destroy(probe2);
    e->dg_conv[l - 1] = ddl_tc_plan_act(1, e->dfull[l], e->P(v), e->dpool[l - 1], B, S, S, cout, cin, 5, nt, ks, 0,
                                        nullptr, nullptr, 1.f, nullptr, 0, 1.f, nullptr, nullptr);
    if (!e->dg_conv[l - 1]) return -1;
    e->dg_split[l - 1] = ddl_tc_plan_needs_zero(e->dg_conv[l - 1]) ? 2 : 1;
  }
  return 0;
}

FwdPlans* get_fwd(DdlEngine* e, int B, bool train) {
  auto& cache = train ? e->train_fwd : e->eval_fwd;
  auto it = cache.find(B);
  if (it != cache.end()) return &it->second;
  FwdPlans fp;
  if (build_fwd(e, B, train, &fp)) { fp.destroy(); return nullptr; }
  return &(cache[B] = fp);
}

// forward through fc2; x [B,784].  Writes pooled[], idx[] (train), h1, h2.
int forward(DdlEngine* e, FwdPlans* fp, const float* x, int B, bool train, cudaStream_t s, cudaEvent_t before_fc = nullptr) {
  int v = 0;
  ddl::no_pdl_once();        // first kernel of the pass: full dependency on whatever updated the weights before it
  if (e->nconv) {
    if (ddl_conv1_relu_pool(x, e->P(0), e->P(1), e->pooled[0], train ? e->idx[0] : nullptr, B, e->S[0], e->C[1], s))
      return -1;
    v = 2;
    for (int l = 1; l < e->nconv; ++l, v += 2) {
      const int S = e->S[l], cout = e->C[l + 1];
      if (fp->fused_pool[l - 1]) {
        if (fp->pconv[l - 1] ? ddl_patch_plan_run(fp->pconv[l - 1], s) : ddl_tc_plan_run(fp->conv[l - 1], s)) return -1;
        continue;
      }
      if (fp->conv_split[l - 1] > 1)
        DDL_CHECK_CUDA(cudaMemsetAsync(e->raw, 0, (size_t)B * S * S * cout * sizeof(float), s));
      if (ddl_tc_plan_run(fp->conv[l - 1], s)) return -1;
      if (ddl_maxpool_fwd(e->raw, e->P(v + 1), e->pooled[l], train ? e->idx[l] : nullptr, B, S, S, cout, s)) return -1;
    }
    if (before_fc) DDL_CHECK_CUDA(cudaStreamWaitEvent(s, before_fc, 0));    // dropout reads the step state
    if (ddl_tc_plan_run(fp->fc[0], s)) return -1;
  } else if (before_fc) {
    DDL_CHECK_CUDA(cudaStreamWaitEvent(s, before_fc, 0));
  }
  if (ddl_tc_plan_run(fp->fc[1], s)) return -1;
  return 0;
}

}  // namespace

extern "C" {

DdlEngine* ddl_engine_create(const DdlEngineCfg* cfg) {
  auto fail = [](DdlEngine* e, const char* why) -> DdlEngine* {
    if (why) ddl::set_error("ddl_engine_create: %s", why);
    if (e) ddl_engine_destroy(e);
    return nullptr;
  };
  if (cfg->kind != 0 && cfg->kind != 1) return fail(nullptr, "kind must be 0 (cnn) or 1 (mlp)");
  if (cfg->kind == 1) return fail(nullptr, "the mlp variant runs on the torch backend (fc1 input is not a plan-bound buffer)");
  DdlEngine* e = new (std::nothrow) DdlEngine();
  if (!e) return fail(nullptr, "out of host memory");
  e->cfg = *cfg;
  e->nconv = 4;
  const int wm = cfg->width_mult;
  e->C[0] = 1; e->C[1] = 32 * wm; e->C[2] = 64 * wm; e->C[3] = 128 * wm; e->C[4] = 256 * wm;
  e->S[0] = 28;
  for (int i = 1; i < 5; ++i) e->S[i] = (e->S[i - 1] + 1) / 2;
  e->F[0] = e->S[4] * e->S[4] * e->C[4]; e->F[1] = 1024 * wm; e->F[2] = 512 * wm; e->F[3] = 10;
  e->maxB = cfg->batch > cfg->max_eval_batch ? cfg->batch : cfg->max_eval_batch;
  const size_t mb = (size_t)e->maxB;
  size_t full_max = 0;
  for (int l = 0; l < 4; ++l) {
    const size_t pooled = mb * e->S[l + 1] * e->S[l + 1] * e->C[l + 1];
    const size_t full = mb * e->S[l] * e->S[l] * e->C[l + 1];
    if (l >= 1 && full > full_max) full_max = full;   // conv1 never touches `raw`
    if (alloc(e, (void**)&e->pooled[l], pooled * 4)) return fail(e, nullptr);
    if (alloc(e, (void**)&e->idx[l], (size_t)cfg->batch * e->S[l + 1] * e->S[l + 1] * e->C[l + 1])) return fail(e, nullptr);
    if (alloc(e, (void**)&e->dpool[l], (size_t)cfg->batch * e->S[l + 1] * e->S[l + 1] * e->C[l + 1] * 4)) return fail(e, nullptr);
  }
  if (alloc(e, (void**)&e->raw, full_max * 4)) return fail(e, nullptr);
  for (int l = 0; l < 4; ++l)
    if (alloc(e, (void**)&e->dfull[l], (size_t)cfg->batch * e->S[l] * e->S[l] * e->C[l + 1] * 4)) return fail(e, nullptr);
  if (cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking) != cudaSuccess) return fail(e, "side stream");
  if (cudaStreamCreateWithFlags(&e->side2, cudaStreamNonBlocking) != cudaSuccess) return fail(e, "side stream 2");
  if (cudaEventCreateWithFlags(&e->ev_fork2, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_join2, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_fc_w, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_fc_m, cudaEventDisableTiming) != cudaSuccess) return fail(e, "event");
  for (auto& ev : e->ev_fork)
    if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return fail(e, "event");
  if (cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming) != cudaSuccess) return fail(e, "event");
  for (int i = 0; i < 2; ++i) {
    if (cudaStreamCreateWithFlags(&e->wside[i], cudaStreamNonBlocking) != cudaSuccess) return fail(e, "wgrad stream");
    if (cudaEventCreateWithFlags(&e->ev_wjoin[i], cudaEventDisableTiming) != cudaSuccess) return fail(e, "event");
  }
  if (cudaEventCreateWithFlags(&e->ev_w0, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_main2, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_pre0, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_pre1, cudaEventDisableTiming) != cudaSuccess) return fail(e, "event");
  if (alloc(e, (void**)&e->h1, mb * e->F[1] * 4) || alloc(e, (void**)&e->h2, mb * e->F[2] * 4) ||
      alloc(e, (void**)&e->dh1, (size_t)cfg->batch * e->F[1] * 4) || alloc(e, (void**)&e->dh2, (size_t)cfg->batch * e->F[2] * 4) ||
      alloc(e, (void**)&e->logits, mb * 16 * 4) || alloc(e, (void**)&e->dlogits, (size_t)cfg->batch * 16 * 4) ||
      alloc(e, (void**)&e->eval_st, sizeof(DdlStepState)))
    return fail(e, nullptr);
  if (cudaMemset(e->eval_st, 0, sizeof(DdlStepState)) != cudaSuccess) return fail(e, "memset failed");
  if (!get_fwd(e, cfg->batch, true)) return fail(e, nullptr);
  if (build_bwd(e)) return fail(e, nullptr);
  return e;
}

void ddl_engine_destroy(DdlEngine* e) {
  if (!e) return;
  for (auto& kv : e->train_fwd) kv.second.destroy();
  for (auto& kv : e->eval_fwd) kv.second.destroy();
  for (auto& p : e->wg_conv) ddl_tc_plan_destroy(p);
  for (auto& p : e->dg_conv) ddl_tc_plan_destroy(p);
  ddl_patch_plan_destroy(e->pdg_conv0);
  for (auto& p : e->wg_fc) ddl_tc_plan_destroy(p);
  for (auto& p : e->dg_fc) ddl_tc_plan_destroy(p);
  for (void* p : e->allocs) cudaFree(p);
  for (auto& ev : e->ev_fork) if (ev) cudaEventDestroy(ev);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  if (e->ev_fork2) cudaEventDestroy(e->ev_fork2);
  if (e->ev_join2) cudaEventDestroy(e->ev_join2);
  if (e->ev_fc_w) cudaEventDestroy(e->ev_fc_w);
  if (e->ev_fc_m) cudaEventDestroy(e->ev_fc_m);
  if (e->side) cudaStreamDestroy(e->side);
  if (e->side2) cudaStreamDestroy(e->side2);
  for (int i = 0; i < 2; ++i) {
    if (e->ev_wjoin[i]) cudaEventDestroy(e->ev_wjoin[i]);
    if (e->wside[i]) cudaStreamDestroy(e->wside[i]);
  }
  if (e->ev_w0) cudaEventDestroy(e->ev_w0);
  if (e->ev_main2) cudaEventDestroy(e->ev_main2);
  if (e->ev_pre0) cudaEventDestroy(e->ev_pre0);
  if (e->ev_pre1) cudaEventDestroy(e->ev_pre1);
  delete e;
}

/* gradients of the mean cross-entropy of (x, y) into the gradient arena; loss_sum / correct into st.
 * The caller advances st (ddl_step_advance) first.  Enqueue-only. */
int ddl_engine_train_step(DdlEngine* e, const float* x, const int* y, cudaStream_t s) {
  const DdlEngineCfg& c = e->cfg;
  const int B = c.batch;
  FwdPlans* fp = get_fwd(e, B, true);
  DDL_REQUIRE(fp != nullptr, "ddl_engine_train_step: no forward plans");
  // the step-state advance (when the engine owns it) and the zeroing of the gradient arena run on the side stream
  // beside conv1..conv4; the first consumers are fc1 (dropout stream / seed) and the fc3 kernel (first gradient)
  DDL_CHECK_CUDA(cudaEventRecord(e->ev_pre0, s));
  DDL_CHECK_CUDA(cudaStreamWaitEvent(e->side, e->ev_pre0, 0));
  if (e->own_advance && ddl_step_advance(c.st, e->adv_lr, e->adv_b1, e->adv_b2, e->adv_opt, e->side)) return -1;
  DDL_CHECK_CUDA(cudaMemsetAsync(c.grads, 0, (size_t)c.arena_elems * sizeof(float), e->side));
  DDL_CHECK_CUDA(cudaEventRecord(e->ev_pre1, e->side));
  if (forward(e, fp, x, B, true, s, e->ev_pre1)) return -1;
  const int vf = 2 * e->nconv;
  // fc3 forward + loss + fc3 backward in one cluster kernel (also emits the fc2 bias gradient = column sums of dh2);
  // DDL_FC3_SPLIT=1 or a shape that does not fit: the two separate kernels
  static const bool fc3_split = DDL_LAB_ENV("DDL_FC3_SPLIT") != nullptr;
  if (!fc3_split && ddl_fc3_fused_fits(B, e->F[2], e->F[3])) {
    if (ddl_fc3_fused_st(e->h2, e->P(vf + 4), e->P(vf + 5), y, e->G(vf + 4), e->G(vf + 5), e->dh2, e->G(vf + 3), c.st, B,
                         e->F[2], e->F[3], 1.f / (float)B, 1, c.keep_prob, s)) return -1;
  } else {
    if (ddl_fc3_xent_fwd(e->h2, e->P(vf + 4), e->P(vf + 5), y, nullptr, e->dlogits, c.st, B, e->F[2], e->F[3],
                         1.f / (float)B, s)) return -1;
    if (ddl_fc3_bwd_st(e->h2, e->P(vf + 4), e->dlogits, e->G(vf + 4), e->G(vf + 5), e->dh2, e->G(vf + 3), B, e->F[2],
                       e->F[3], c.st, 1, c.keep_prob, s)) return -1;
  }
  // Main stream: the data-gradient chain.  Side stream: every weight-gradient GEMM, forked as soon as its
  // dY exists and joined once before the optimizer.
  cudaStream_t w = e->side;
  // stream of layer l's conv wgrad (l = 3, 2, 1 <-> conv4, conv3, conv2)
  auto wstream = [&](int l) -> cudaStream_t { return l == 3 ? w : e->wside[2 - l]; };
  int fk = 0;
  auto fork = [&](cudaStream_t to) -> int {
    DDL_CHECK_CUDA(cudaEventRecord(e->ev_fork[fk], s));
    DDL_CHECK_CUDA(cudaStreamWaitEvent(to, e->ev_fork[fk], 0));
    ++fk;
    return 0;
  };
  // Overlapped push buckets (each has its own signal row, so their kernels need no order among themselves):
  //   bucket 0 = the FC layers (60% of the bytes), stream side2: once their weight gradients (stream w) exist and fc1's
  //              dgrad - the last reader of a weight this bucket updates in place - has run on the main stream;
  //   bucket 1 = conv4, behind its weight gradient on stream w, once conv4's dgrad has run;
  //   bucket 2 = conv3, behind its weight gradient on wside[0], once conv3's dgrad has run.
  // The bias gradients come from main-stream kernels that have run by then.  The caller launches the last bucket
  // (conv1 / conv2) after the step.
  auto launch_bucket = [&](int bucket, cudaStream_t bs) -> int {
    for (int i = 0; i < e->ov_n; ++i)
      if (e->ov_args[i].bucket == bucket && ddl_sync_push_update_pull(&e->ov_comm, &e->ov_args[i], e->ov_kind, bs)) return -1;
    return 0;
  };
  auto launch_overlap_fc = [&]() -> int {
    DDL_CHECK_CUDA(cudaEventRecord(e->ev_fc_w, w));
    DDL_CHECK_CUDA(cudaStreamWaitEvent(e->side2, e->ev_fc_w, 0));
    DDL_CHECK_CUDA(cudaEventRecord(e->ev_fc_m, s));
    DDL_CHECK_CUDA(cudaStreamWaitEvent(e->side2, e->ev_fc_m, 0));
    if (launch_bucket(0, e->side2)) return -1;
    DDL_CHECK_CUDA(cudaEventRecord(e->ev_join2, e->side2));
    return 0;
  };
  // bucket `bucket` on the stream `ws` of the layer's weight gradient, after the main stream's latest kernel (the
  // layer's dgrad: the last reader of the weights the bucket updates in place)
  auto launch_overlap_conv = [&](int bucket, cudaStream_t ws, cudaEvent_t ev) -> int {
    DDL_CHECK_CUDA(cudaEventRecord(ev, s));
    DDL_CHECK_CUDA(cudaStreamWaitEvent(ws, ev, 0));
    return launch_bucket(bucket, ws);
  };
  if (fork(w)) return -1;
  if (ddl_tc_plan_run(e->wg_fc[1], w)) return -1;
  if (ddl_tc_plan_run(e->dg_fc[1], s)) return -1;      // + fc1 bias gradient in its epilogue
  if (fork(w)) return -1;
  if (ddl_tc_plan_run(e->wg_fc[0], w)) return -1;
  if (ddl_tc_plan_needs_zero(e->dg_fc[0]))             // plain split-K output: partial sums are red.add-ed
    DDL_CHECK_CUDA(cudaMemsetAsync(e->dpool[3], 0, (size_t)B * e->F[0] * sizeof(float), s));
  if (ddl_tc_plan_run(e->dg_fc[0], s)) return -1;
  if (e->ov_n > 0 && e->nconv && launch_overlap_fc()) return -1;
  int v = 6;
  const bool unpool = e->dg_unpool[0] && e->dg_unpool[1] && e->dg_unpool[2];
  if (unpool) {
    // dp4 -> dY of conv4 (the only stand-alone pool backward left); every conv dgrad then writes the dY of the layer
    // below itself
    if (ddl_maxpool_bwd(e->dpool[3], e->idx[3], e->dfull[3], e->G(7), B, e->S[3], e->S[3], e->C[4], s)) return -1;
    for (int l = 3; l >= 1; --l) {
      cudaStream_t wl = wstream(l);
      if (fork(wl)) return -1;
      if (ddl_tc_plan_run(e->wg_conv[l - 1], wl)) return -1;
      if ((l == 1 && e->pdg_conv0) ? ddl_patch_plan_run(e->pdg_conv0, s) : ddl_tc_plan_run(e->dg_conv[l - 1], s)) return -1;
      if (l == 3 && e->ov_n > 0 && launch_overlap_conv(1, wl, e->ev_fork2)) return -1;
      if (l == 2 && e->ov_n > 0 && launch_overlap_conv(2, wl, e->ev_main2)) return -1;
    }
  }
  for (int l = unpool ? 0 : 3; l >= 1; --l, v -= 2) {           // conv4, conv3, conv2
    const int S = e->S[l], cout = e->C[l + 1], cin = e->C[l];
    if (e->dg_split[l - 1] > 1)
      DDL_CHECK_CUDA(cudaMemsetAsync(e->dpool[l - 1], 0, (size_t)B * S * S * cin * sizeof(float), s));
    if (ddl_maxpool_bwd(e->dpool[l], e->idx[l], e->dfull[l], e->G(v + 1), B, S, S, cout, s)) return -1;
    cudaStream_t wl = wstream(l);
    if (fork(wl)) return -1;
    if (ddl_tc_plan_run(e->wg_conv[l - 1], wl)) return -1;
    if (ddl_tc_plan_run(e->dg_conv[l - 1], s)) return -1;
    if (l == 3 && e->ov_n > 0 && launch_overlap_conv(1, wl, e->ev_fork2)) return -1;
    if (l == 2 && e->ov_n > 0 && launch_overlap_conv(2, wl, e->ev_main2)) return -1;
  }
  if (!unpool && ddl_maxpool_bwd(e->dpool[0], e->idx[0], e->dfull[0], nullptr, B, e->S[0], e->S[0], e->C[1], s)) return -1;
  if (unpool && e->w1_pooled) {
    // no programmatic launch here: this kernel's only early work is a 3 KB image load, and launched early its 200 blocks
    // sat on the SMs next to conv2's dgrad for its whole duration (measured: 0.2222 -> 0.2179 ms per step without)
    ddl::no_pdl_once();
    if (ddl_conv1_wgrad_pooled(x, e->dpool[0], e->idx[0], e->G(0), nullptr, B, e->S[0], e->C[1], s)) return -1;
  } else if (ddl_conv1_wgrad(x, e->dfull[0], e->G(0), e->G(1), B, e->S[0], e->C[1], s)) {
    return -1;
  }
  DDL_CHECK_CUDA(cudaEventRecord(e->ev_join, w));
  DDL_CHECK_CUDA(cudaStreamWaitEvent(s, e->ev_join, 0));
  for (int i = 0; i < 2; ++i) {
    DDL_CHECK_CUDA(cudaEventRecord(e->ev_wjoin[i], e->wside[i]));
    DDL_CHECK_CUDA(cudaStreamWaitEvent(s, e->ev_wjoin[i], 0));
  }
  if (e->ov_n > 0) DDL_CHECK_CUDA(cudaStreamWaitEvent(s, e->ev_join2, 0));
  return 0;
}

/* host-only planning helpers, exposed for tests: the (N tile, cluster) choice of the staged conv plans and of the FC
 * plans (ntile_cluster[0] = N tile, [1] = cluster / K split) */
int ddl_plan_pick_staged(int pix_tiles, int nout, int kiters, int* ntile_cluster) {
  pick_staged_tiling(pix_tiles, nout, kiters, &ntile_cluster[0], &ntile_cluster[1]);
  return 0;
}
int ddl_plan_pick_fc(int nout, int kiters, int* ntile_cluster) {
  pick_fc_tiling(nout, kiters, &ntile_cluster[0], &ntile_cluster[1]);
  return 0;
}

int ddl_engine_set_advance(DdlEngine* e, float lr, float b1, float b2, int advance_opt) {
  DDL_REQUIRE(e != nullptr, "ddl_engine_set_advance: null engine");
  e->own_advance = true;
  e->adv_lr = lr; e->adv_b1 = b1; e->adv_b2 = b2; e->adv_opt = advance_opt;
  return 0;
}

int ddl_engine_set_overlap(DdlEngine* e, const DdlComm* c, const DdlSyncArgs* args, int nargs, int kind) {
  DDL_REQUIRE(e != nullptr && nargs >= 0 && nargs <= 8, "ddl_engine_set_overlap: bad arguments");
  if (nargs > 0) e->ov_comm = *c;
  for (int i = 0; i < nargs; ++i) e->ov_args[i] = args[i];
  e->ov_n = nargs;
  e->ov_kind = kind;
  return 0;
}

/* forward only (keep_prob = 1) over n <= max_eval_batch rows; adds the number of correct rows to
 * *correct (device) and optionally writes logits [n,10]. */
int ddl_engine_eval(DdlEngine* e, const float* x, const int* y, int n, unsigned* correct, float* logits_out,
                    cudaStream_t s) {
  DDL_REQUIRE(n >= 1 && n <= e->maxB, "ddl_engine_eval: n=%d exceeds max_eval_batch=%d", n, e->maxB);
  FwdPlans* fp = get_fwd(e, n, false);
  DDL_REQUIRE(fp != nullptr, "ddl_engine_eval: could not build plans for n=%d: %s", n, ddl_last_error());
  if (forward(e, fp, x, n, false, s)) return -1;
  const int vf = 2 * e->nconv;
  float* lg = logits_out ? logits_out : e->logits;
  if (ddl_fc3_xent_fwd(e->h2, e->P(vf + 4), e->P(vf + 5), nullptr, lg, nullptr, nullptr, n, e->F[2], e->F[3], 1.f, s))
    return -1;
  if (y && correct) return ddl_accuracy(lg, y, correct, n, e->F[3], s);
  return 0;
}

/* debug / test access to intermediate buffers: 0..3 pooled p1..p4, 4 h1, 5 h2, 6 dh1, 7 dh2,
 * 8..11 dpool p1..p4, 12 dfull, 13 raw */
const float* ddl_engine_buffer(const DdlEngine* e, int which) {
  if (which >= 0 && which < 4) return e->pooled[which];
  if (which == 4) return e->h1;
  if (which == 5) return e->h2;
  if (which == 6) return e->dh1;
  if (which == 7) return e->dh2;
  if (which >= 8 && which < 12) return e->dpool[which - 8];
  if (which == 12) return e->dfull[0];
  if (which == 13) return e->raw;
  return nullptr;
}

}  // extern "C"
