/* ddl_b200 native library - C ABI.
 *
 * Every entry point returns 0 on success and a negative code on failure; ddl_last_error()
 * then holds a human-readable message.  Pointers are device pointers unless noted.  All work
 * is enqueued on the given stream; nothing synchronises unless documented.
 *
 * The reference (epikjjh/DIstributed-Deep-Learning) contains no native code: each entry point
 * cites the TensorFlow / mpi4py call site whose work it takes over.
 */
#ifndef DDL_B200_H_
#define DDL_B200_H_

#include <stdint.h>
#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDL_MAX_PEERS 16

/* Device-resident per-step scalars (one 64-byte line).  Kept on the device so a whole training
 * step can be replayed from a CUDA graph with no host-written kernel arguments. */
typedef struct DdlStepState {
  uint32_t t;             /* optimizer steps applied to my shard (sync: == worker steps)        */
  float lr_t;             /* TF1 Adam step size for step t                                      */
  float loss_sum;         /* sum of per-sample cross-entropy of the current batch               */
  uint32_t correct;       /* correctly classified samples of the current batch                  */
  uint32_t seed;          /* dropout seed                                                       */
  uint32_t stream;        /* dropout stream of this step's first dropout layer (+1 for the next) */
  uint32_t wstep;         /* worker steps started                                                */
  uint32_t stream_base;   /* 2 * worker index                                                    */
  uint32_t stream_stride; /* 2 * number of workers: stream = base + (wstep-1) * stride           */
  float loss_mean;        /* loss_sum scaled by the kernels' `scale` argument (1 / batch): the step's mean loss   */
  uint32_t pad[6];
} DdlStepState;

const char* ddl_last_error(void);
int ddl_version(void);
/* kernels launched by this library in this process so far (a launch recorded during stream capture counts once) */
unsigned long long ddl_launch_count(void);
int ddl_memcpy_async(void* dst, const void* src, size_t bytes, cudaStream_t s);

/* ---- step state / optimizers (ApplyAdam: mnist_sync/parameter_server.py:21-27; model.py:106) */
/* Start of a worker step: wstep += 1, dropout stream, loss/correct reset; when advance_opt != 0
 * also t += 1 and lr_t (sync mode, where the shard steps in lock-step with the workers). */
int ddl_step_advance(DdlStepState* st, float lr, float b1, float b2, int advance_opt, cudaStream_t s);
/* kind 0 sgd | 1 momentum | 2 adam(TF1).  p, m, v, g are 16B-aligned ranges of n floats.
 * step_size is lr (sgd, momentum) or lr_t (adam; read from st->lr_t when st != NULL).
 * a = momentum or beta1, b = beta2.  g is multiplied by gscale first (sum/mean choice, Q3). */
int ddl_optimizer_step(int kind, float* p, float* m, float* v, const float* g, int64_t n,
                       float step_size, float a, float b, float eps, float gscale,
                       const DdlStepState* st, cudaStream_t s);

/* ---- SIMT model ops ------------------------------------------------------------------------ */
/* conv1 (Cin=1) + bias + ReLU + 2x2 SAME max-pool (model.py:24-31).  idx may be NULL (eval). */
int ddl_conv1_relu_pool(const float* x, const float* w, const float* bias, float* y, uint8_t* idx,
                        int B, int S, int C, cudaStream_t s);
int ddl_conv1_wgrad(const float* x, const float* dy, float* dw, float* db, int B, int S, int C,
                    cudaStream_t s);
/* the same gradients from the POOLED layout: after pool + ReLU backward only the argmax pixel of each 2x2 window is
 * non-zero, so the sum runs over the windows (idx [B,S/2,S/2,C], slots 0..3, 4 = dead) - a quarter of the work */
int ddl_conv1_wgrad_sparse(const float* x, const float* dy, const uint8_t* idx, float* dw, float* db, int B, int S, int C,
                           cudaStream_t s);
/* conv1 weight gradient from the pooled gradient: gp [B, S/2, S/2, C] (ReLU-dead windows already 0), idx the argmax
 * slots; dw[25][C] += sum over windows of gp * x[argmax pixel + tap offset]; db (nullable) += column sums of gp.
 * x is read before griddepcontrol.wait: it must not be produced by the kernel launched just before.  S == 28. */
int ddl_conv1_wgrad_pooled(const float* x, const float* gp, const uint8_t* idx, float* dw, float* db, int B, int S, int C,
                           cudaStream_t s);
/* (bias +) ReLU + tf.nn.max_pool2d 2x2/2 SAME on NHWC conv sums and its gradient (model.py:31,42,53,64);
 * the backward also accumulates the bias gradient of the producing conv into db (may be NULL). */
int ddl_maxpool_fwd(const float* a, const float* bias, float* y, uint8_t* idx, int B, int H, int W, int C,
                    cudaStream_t s);
int ddl_maxpool_bwd(const float* dp, const uint8_t* idx, float* dy, float* db, int B, int H, int W, int C,
                    cudaStream_t s);
int ddl_colsum(const float* x, float* out, int R, int C, cudaStream_t s);
/* fc3 + softmax_cross_entropy_with_logits + reduce_mean (model.py:85-92) and accuracy
 * (model.py:104-105).  dlogits = (softmax - onehot) * scale.  logits / dlogits / labels may be NULL. */
int ddl_fc3_xent_fwd(const float* h, const float* w, const float* bias, const int* labels, float* logits,
                     float* dlogits, DdlStepState* st, int B, int K, int NC, float scale, cudaStream_t s);
int ddl_fc3_bwd(const float* h, const float* w, const float* dl, float* dw, float* db, float* dh, int B, int K,
                int NC, uint32_t seed, uint32_t stream, float keep_prob, cudaStream_t s);
/* same, dropout seed/stream read from the device step state (graph-replay safe) */
/* dbh (may be NULL): += column sums of dh = bias gradient of the layer that produced h */
int ddl_fc3_bwd_st(const float* h, const float* w, const float* dl, float* dw, float* db, float* dh, float* dbh,
                   int B, int K, int NC, const DdlStepState* st, uint32_t stream_off, float keep_prob,
                   cudaStream_t s);
/* training path: fc3 forward + softmax-xent + fc3 backward (+ fc2 bias gradient) in ONE launch of a thread-block
 * cluster of 8 CTAs (model.py:85-92 forward, the matching gradients of model.py:101): each CTA computes the dlogits of
 * its rows and stores them into every CTA's shared memory (distributed shared memory), then owns K/8 columns of the
 * backward.  dw / db / dbh are accumulated (+=).  W3, bias and labels are read before griddepcontrol.wait: they must
 * not be outputs of the kernel launched just before this one.  ddl_fc3_fused_fits() tells whether a shape fits. */
int ddl_fc3_fused_st(const float* h, const float* w, const float* bias, const int* labels, float* dw, float* db,
                     float* dh, float* dbh, DdlStepState* st, int B, int K, int NC, float scale, uint32_t stream_off,
                     float keep_prob, cudaStream_t s);
int ddl_fc3_fused_fits(int B, int K, int NC);
int ddl_accuracy(const float* logits, const int* labels, unsigned* correct, int B, int NC, cudaStream_t s);

/* ---- tcgen05 implicit-GEMM plans (csrc/tc_gemm.cu) ------------------------------------------
 * conv2..4 forward (tf.nn.conv2d, model.py:39,50,61), their backprop-input / backprop-filter,
 * and the fc1 / fc2 matmuls forward and backward (model.py:70,79).  A plan owns the TMA
 * descriptors for fixed pointers/shapes; running it is one kernel launch.
 *   act plan, dgrad=0: out[pix,n] = epi(sum_{tap,c} act[pix+off(tap), c] * w[tap][c][n])   (w: HWIO)
 *   act plan, dgrad=1: out[pix,n] = epi(sum_{tap,c} act[pix-off(tap), c] * w[tap][n][c])
 *   wgrad plan:        out[tap][b][a] += sum_pix g[pix,a] * x[pix+off(tap), b]
 * flags: 1 bias, 2 relu, 4 dropout (seed/stream from st), 8 multiply by (mask_act > 0) * mask_scale.
 * ksplit > 1 without epilogue flags: raw partial sums are added with red.add into a pre-zeroed out
 * (ddl_tc_plan_needs_zero() returns 1).  ksplit in [2,8] WITH epilogue flags / colsum: the split CTAs of a
 * tile form a thread-block cluster and reduce through plan-owned scratch, keeping the fused epilogue.
 * colsum (may be NULL): += per-column sums of the final output values (a bias gradient).
 * Launch-order contract: every plan kernel starts under programmatic dependent launch and loads its WEIGHT operand
 * (`w`) before griddepcontrol.wait, i.e. possibly while earlier kernels of the stream still run.  `w` must therefore
 * not be written by a kernel of this library launched shortly before in the same stream without an intervening
 * full dependency (the engine launches the first kernel of every pass without the attribute for this reason);
 * outputs of other libraries' kernels (no early-launch trigger) are always safe.
 * w_store_dst (may be NULL; fprop only) - K4, the pull fused into the consumer GEMM: `w` may then point
 * into ANOTHER rank's parameter arena (a peer-mapped pointer): the weight tiles are TMA-loaded over
 * NVLink straight into the MMA's shared-memory stage and written through to w_store_dst (this rank's
 * copy, same layout) by a TMA store, so no separate all-gather of that variable is needed
 * (SS/ps.py:126 + SS/worker.py:94,97 + model.py:70,79 in one kernel). */
typedef struct DdlTcPlan DdlTcPlan;
DdlTcPlan* ddl_tc_plan_act(int dgrad, const float* act, const float* w, float* out, int N, int H, int W,
                           int cred, int nout, int ksize, int ntile, int ksplit, int flags, const float* bias,
                           const float* mask_act, float mask_scale, const DdlStepState* st,
                           uint32_t drop_stream_off, float keep_prob, float* colsum, float* w_store_dst);
DdlTcPlan* ddl_tc_plan_wgrad(const float* g, const float* x, float* out, int N, int H, int W, int ca, int cb,
                             int ksize, int ntile, int psplit);
/* Convolutions with the layer's pooling fused into the GEMM epilogue (model.py:39-42 / 50-53 / 61-64 in one kernel
 * forward; the pool + ReLU backward of the layer below in the dgrad kernel).  cluster in {1, 2, 4, 8}: the K loop is
 * split over a thread-block cluster; its CTAs reduce-scatter their partial accumulator tiles through distributed
 * shared memory (rank r sums and finishes columns [r * ntile / cluster, (r + 1) * ntile / cluster)), so a split
 * reduction keeps its fused epilogue and needs neither zeroed outputs nor atomics.
 *   conv_pool:    pooled[N, ceil(H/2), ceil(W/2), cout] = maxpool2x2_SAME(relu(conv5x5_SAME(act, w) + bias));
 *                 idx (may be NULL): argmax slot of every window, 0..3, 4 = the window's maximum is <= 0 (no gradient)
 *   dgrad_unpool: dfull[N, UH, UW, cin] = un-pool(idx, conv dgrad(dy, w)), dbias (may be NULL) += column sums */
DdlTcPlan* ddl_tc_plan_conv_pool(const float* act, const float* w, const float* bias, float* pooled, uint8_t* idx,
                                 int N, int H, int W, int cin, int cout, int ksize, int ntile, int cluster);
DdlTcPlan* ddl_tc_plan_dgrad_unpool(const float* dy, const float* w, const uint8_t* idx, float* dfull, float* dbias,
                                    int N, int H, int W, int UH, int UW, int cout, int cin, int ksize, int ntile,
                                    int cluster);
/* conv dgrad + the ReLU half of the pool backward, gradient kept in the POOLED layout: gp[N, H, W, cin] = dgrad value
 * where idx < 4 (window alive), 0 where the window is ReLU-dead; dbias += column sums.  Used for conv2's dgrad: conv1
 * (model.py:24-31) has no data gradient, and its weight gradient (ddl_conv1_wgrad_pooled) needs only value + argmax
 * slot per window - the un-pooled plane (4x the bytes, 75% zeros) is never materialised. */
DdlTcPlan* ddl_tc_plan_dgrad_poolgrad(const float* dy, const float* w, const uint8_t* idx, float* gp, float* dbias,
                                      int N, int H, int W, int cout, int cin, int ksize, int ntile, int cluster);
/* host-only planning helpers (no device needed; used by the CPU tests): pixel box of a plan, (N tile, cluster) choice
 * of the staged conv plans and of the FC plans */
int ddl_plan_choose_box(int N, int H, int W, int max_rows, int row_mult, int even_windows, int* box3);
int ddl_plan_pick_staged(int pix_tiles, int nout, int kiters, int* ntile_cluster);
int ddl_plan_pick_fc(int nout, int kiters, int* ntile_cluster);
int ddl_tc_plan_run(const DdlTcPlan* pl, cudaStream_t s);
void ddl_tc_plan_destroy(DdlTcPlan* pl);
int ddl_tc_plan_info(const DdlTcPlan* pl, int* out8);
int ddl_tc_plan_needs_zero(const DdlTcPlan* pl);

/* ---- patch-resident 5x5 convolution (csrc/tc_patch.cu): one CTA per image, the zero-padded activation patch is loaded
 * once and every filter tap reads it through a shifted shared-memory descriptor (conv2d + bias + relu + max_pool2d of
 * model.py:35-42 and its backprop-input).  dgrad = 0: out [N,ceil(H/2),ceil(W/2),nout], idx nullable; dgrad = 1: out
 * [N,H,W,nout] = dgrad masked by idx < 4, dbias += column sums.  ddl_patch_conv_fits says whether a shape qualifies. */
typedef struct DdlPatchPlan DdlPatchPlan;
int ddl_patch_conv_fits(int H, int W, int cred, int nout);
int ddl_patch_conv_preferred(int H, int W, int cred, int nout);   /* fits AND >= 70% of its M rows are real pixels */
DdlPatchPlan* ddl_patch_plan_create(int dgrad, const float* act, const float* w, const float* bias, float* out, uint8_t* idx,
                                    float* dbias, int N, int H, int W, int cred, int nout);
int ddl_patch_plan_run(const DdlPatchPlan* pl, cudaStream_t s);
void ddl_patch_plan_destroy(DdlPatchPlan* pl);
int ddl_patch_plan_set_trace(DdlPatchPlan* pl, unsigned long long* trace);   /* lab build (ops/build.py --lab) only */

/* ---- allocation-free convolution block (csrc/tc_gemm.cu).  No counterpart in the reference. --------
 * For host programs that must not call cudaMalloc: ones that have given the whole device to another
 * allocator (a caching pool that owns all free memory, an MPS / MIG slice with a hard quota), where a
 * further cudaMalloc fails, and latency-critical loops that cannot take cudaMalloc's device-wide
 * synchronisation.  y = maxpool2x2_SAME(relu(conv5x5_SAME(x, w) + bias)); x [N,H,W,cin], w [5,5,cin,cout],
 * y [N,ceil(H/2),ceil(W/2),cout], idx (may be NULL) the pool arg-max bytes.  The raw convolution sums live
 * in an array built into the library (DDL_STATIC_SCRATCH_BYTES on every GPU that loads it, held until
 * exit): N*H*W*cout*4 bytes must fit or the call is refused with an error naming the limit.  The first
 * call on a device does a one-time set-up and should be made before any stream capture starts.  All calls
 * share that array (it keeps the last call's contents; this function zeroes what it accumulates into), so
 * calls that overlap on the GPU race - serialise them.  Calls no cudaMalloc / cudaFree / allocating routine. */
#define DDL_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
int ddl_conv5x5_relu_pool_static(const float* x, const float* w, const float* bias, float* y, uint8_t* idx,
                                 int N, int H, int W, int cin, int cout, cudaStream_t s);

/* ---- peer-memory communication (csrc/comm.cu) -----------------------------------------------
 * Replaces the reference's mpi4py layer (SURVEY 2.3 call-site table). */
#define DDL_MAX_SKIP 4
#define DDL_MAX_SUB 16
#define DDL_STALE_BINS 16
/* signal pad layout (uint32 slots, one pad per rank) */
#define DDL_PAD_BARRIER 16   /* [16,32): host-requested device barrier epochs, one slot per source */
#define DDL_PAD_PUSHES 32    /* async: complete pushes landed in this owner's accumulator          */
#define DDL_PAD_VERSION 33   /* async: optimizer applications done by this owner                   */
#define DDL_PAD_PUSHES_W 48  /* [48,64): async: complete pushes of worker w landed in this owner's plane w      */
#define DDL_PAD_SYNC 64      /* [64, 64 + 16*8): sync-bucket barrier epochs, one row of 16 slots per bucket  */
#define DDL_PAD_SLOTS 256

typedef struct DdlComm {
  int rank, world;
  float* grads[DDL_MAX_PEERS];     /* every rank's gradient arena (peer-mapped)   */
  float* params[DDL_MAX_PEERS];    /* every rank's parameter arena                */
  float* accum[DDL_MAX_PEERS];     /* every rank's async gradient accumulator     */
  uint32_t* pads[DDL_MAX_PEERS];   /* every rank's signal pad                     */
  /* NVLS (NVSwitch multicast) views, NULL when unavailable: one address that names the SAME offset of every
   * rank's arena.  multimem.ld_reduce on mc_grads returns the in-switch sum over all ranks (owner ingress
   * 1x instead of Wx); multimem.st on mc_params writes every rank's copy at once (owner egress 1x). */
  float* mc_grads;
  float* mc_params;
} DdlComm;

typedef struct DdlSyncArgs {
  int64_t begin, end;              /* my owned arena range (elements, multiples of 4) */
  float* m; float* v;              /* optimizer state, indexed like the arena          */
  DdlStepState* st;
  uint32_t* done_counter;          /* device scratch word, zero between launches        */
  uint32_t worker_mask;            /* ranks whose gradients are summed                  */
  uint32_t dest_mask;              /* ranks that receive the updated range              */
  float lr, a, b, eps, gscale;
  int64_t skip_begin[DDL_MAX_SKIP], skip_end[DDL_MAX_SKIP];   /* not broadcast (K4 pulls them) */
  /* Gradient bucketing (backward-overlapped push): a step may be served by `nbuckets` launches, each over
   * `nsub` sub-ranges of [begin, end) (nsub == 0: the whole range).  Bucket b of step t uses barrier epochs
   * 2*nbuckets*t + 2b + 1 / + 2, so every rank must launch the same buckets in the same order. */
  int nsub, bucket, nbuckets;
  int64_t sub_begin[DDL_MAX_SUB], sub_end[DDL_MAX_SUB];
} DdlSyncArgs;

typedef struct DdlAsyncArgs {
  int num_shards;
  int owner_rank[DDL_MAX_PEERS];
  int64_t shard_begin[DDL_MAX_PEERS], shard_end[DDL_MAX_PEERS];
  int direct_sgd;                  /* 1: red.add(-lr*g) straight into the owners' parameters */
  float lr;
  uint32_t* done_counter;
  uint32_t* pulled_version;        /* [num_shards] owner version seen at the last pull    */
  uint32_t* stale_hist;            /* [DDL_STALE_BINS] staleness histogram of my pushes   */
  int slot;                        /* my worker index = my accumulator plane on every owner */
  int64_t slot_stride;             /* elements between planes (the arena size)             */
} DdlAsyncArgs;

/* owner-side bookkeeping of async mode (device memory, zero-initialised) */
typedef struct DdlAsyncState {
  uint32_t applied[DDL_MAX_PEERS]; /* pushes of worker w already consumed                              */
  uint32_t mask, nsteps;           /* latest apply: workers consumed, optimizer steps taken            */
  uint32_t folded;                 /* pushes that shared an optimizer step with another push           */
  uint32_t pad;
  float lr_t[DDL_MAX_PEERS];       /* TF1-Adam step size of each step of the latest apply              */
} DdlAsyncState;

/* NVLink probes behind profiles/nvlink_peak.json: dependent-load / store+fence round trips of one thread (out4: 4 x
 * uint64 device words = total ns, first, max, checksum), a polling kernel and a float4 copy kernel for bandwidth */
int ddl_p2p_latency(const void* src, int words, int iters, unsigned idle_ns, int store_fence, unsigned long long* out4,
                    cudaStream_t s);
int ddl_p2p_spin(const void* flag, int blocks, unsigned long long ns, unsigned long long* out1, cudaStream_t s);
int ddl_p2p_copy(void* dst, const void* src, int64_t bytes, int blocks, cudaStream_t s);
int ddl_dev_malloc(void** p, size_t bytes);
int ddl_dev_free(void* p);
int ddl_ipc_export(void* p, void* handle64);
int ddl_ipc_open(const void* handle64, void** p);
int ddl_ipc_close(void* p);
int ddl_set_device(int dev);
int ddl_can_access_peer(int dev, int peer);
/* K1 + K3, sync mode (SS/worker.py:36 + SS/ps.py:112-126 + SS/worker.py:90-97). */
int ddl_sync_push_update_pull(const DdlComm* c, const DdlSyncArgs* a, int kind, cudaStream_t s);
/* K2, async mode (AS/worker.py:36 + AS/ps.py:99-111). */
int ddl_async_push(const DdlComm* c, const DdlAsyncArgs* a, cudaStream_t s);
/* Owner side: consume what has landed since the last call.  Default (fold = 0): ONE optimizer step per worker with a
 * pending push, in worker order, each on that worker's own accumulator plane - mnist_async/parameter_server.py:55-69
 * applies Adam once per received gradient.  fold = 1: one step on the sum of all pending planes (round-1 behaviour). */
int ddl_async_apply(const DdlComm* c, const DdlSyncArgs* a, int kind, DdlAsyncState* state, int nworkers,
                    int64_t slot_stride, int fold, float b1, float b2, cudaStream_t s);
/* Resident owner of a dedicated async PS GPU (mnist_async/parameter_server.py:52-69 as ONE kernel that stays on the
 * device): polls the per-worker push counters, applies one optimizer step per push as they land (fold = 1: one per
 * round), publishes the shard version, until the host writes a non-zero word to *host_stop (pinned host memory). */
typedef struct DdlResident {
  uint32_t round, stop, done, rounds_served;   /* device words, zero-initialised */
  const uint32_t* host_stop;                   /* pinned host word the kernel polls */
} DdlResident;
int ddl_async_owner_start(const DdlComm* c, const DdlSyncArgs* a, int kind, DdlAsyncState* state, DdlResident* resident,
                          int nworkers, int64_t slot_stride, int fold, float b1, float b2, cudaStream_t s);
/* K3 stand-alone (initial pull, async pull, pre-eval refresh). */
int ddl_pull_allgather(const DdlComm* c, const DdlAsyncArgs* a, cudaStream_t s);
int ddl_device_barrier(const DdlComm* c, uint32_t* epoch_counter, cudaStream_t s);

/* ---- native step executor (csrc/engine.cu) ---------------------------------------------------
 * The TF1 graph + Session of model.py:17-112 as enqueue-only C calls (CUDA-graph capturable). */
typedef struct DdlEngineCfg {
  int kind;                 /* 0 cnn (the reference model, optionally widened) */
  int width_mult;
  int batch;                /* training batch size (reference: 100)            */
  int max_eval_batch;       /* largest evaluation chunk                         */
  float keep_prob;
  float* params;            /* flat parameter arena                             */
  float* grads;             /* flat gradient arena                              */
  int64_t arena_elems;
  int64_t var_offset[16];   /* element offset of v0..v13 in both arenas          */
  DdlStepState* st;
  const float* fc_w_src[2]; /* K4: where the training forward fetches the fc1 / fc2 weight matrices from -
                               the OWNER's arena (peer pointer); NULL = this rank's own copy            */
} DdlEngineCfg;
typedef struct DdlEngine DdlEngine;
DdlEngine* ddl_engine_create(const DdlEngineCfg* cfg);
void ddl_engine_destroy(DdlEngine* e);
/* sess.run(self.grads, keep_prob=.5) (mnist_sync/worker.py:30): x [B,784], y [B] class ids */
int ddl_engine_train_step(DdlEngine* e, const float* x, const int* y, cudaStream_t s);
/* After this call ddl_engine_train_step launches ddl_step_advance(st, lr, b1, b2, advance_opt) itself, on its side
 * stream beside the convolutions (the first reader of the step state is fc1's dropout); the caller must then NOT
 * advance the state before the step. */
int ddl_engine_set_advance(DdlEngine* e, float lr, float b1, float b2, int advance_opt);
/* Backward-overlapped push: the engine launches ddl_sync_push_update_pull for these argument blocks while the main
 * stream still runs the rest of the backward - blocks with bucket == 0 (the FC layers, 60% of the bytes) right after
 * fc1's dgrad, bucket == 1 (conv4) after conv4's dgrad and wgrad, bucket == 2 (conv3) after conv3's.  The caller
 * launches the last bucket (conv1 / conv2) after the step.  nargs <= 8. */
int ddl_engine_set_overlap(DdlEngine* e, const DdlComm* c, const DdlSyncArgs* args, int nargs, int kind);
/* sess.run(self.accuracy, keep_prob=1) (mnist_sync/worker.py:72) over n rows */
int ddl_engine_eval(DdlEngine* e, const float* x, const int* y, int n, unsigned* correct, float* logits_out,
                    cudaStream_t s);
const float* ddl_engine_buffer(const DdlEngine* e, int which);

#ifdef __cplusplus
}
#endif
#endif /* DDL_B200_H_ */
