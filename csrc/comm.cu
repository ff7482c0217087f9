// Peer-memory communication kernels: the product's replacement for the reference's whole MPI layer.
//
// Reference call sites taken over: worker push ``comm.Send([grad, FLOAT], dest, tag)``
// x14 (mnist_sync_sharding/worker.py:36), PS receive loop + numpy sum + ApplyAdam
// (mnist_sync_sharding/parameter_server.py:112-114, 78-82), PS fan-out ``comm.Send`` / ``Bcast``
// (parameter_server.py:124-126, mnist_sync/parameter_server.py:69), worker ``Recv`` + ``tf.assign``
// (worker.py:90-97); async: ``Recv(ANY_SOURCE)`` + apply + reply
// (mnist_async_sharding/parameter_server.py:99-111).
//
// Every rank maps every other rank's gradient arena, parameter arena, accumulator and signal pad
// (cudaIpc handles over NVLink / NVSwitch).  Kernels then move the data themselves:
//   K1+K3  ddl_sync_push_update_pull : device barrier -> owner pulls its range from all workers'
//          gradient arenas (fixed order: bitwise reproducible), scales, applies SGD / momentum /
//          TF1-Adam in place, and stores the new values into every worker's parameter arena ->
//          device barrier.  One launch replaces 28 blocking MPI calls + 16 Session.run calls.
//   K2     ddl_async_push            : vectorised red.global.add of a worker's gradient into the
//          owners' accumulators (or, for SGD, of -lr*g straight into the parameters) + a
//          release-ordered push counter.
//          ddl_async_apply           : owner consumes its accumulator exactly once (atomicExch),
//          applies the optimizer, bumps the shard version.
//   K3     ddl_pull_allgather        : worker copies the owners' ranges (peer loads).
// All spins are time-bounded (DDL_WAIT_TIMEOUT_NS) and trap instead of hanging when a peer dies.
#include "common.cuh"
#include "../include/ddl_b200.h"
#include <string.h>
#include <stdlib.h>

namespace {

// 128-thread blocks (~5k registers each) fit beside two resident GEMM CTAs on every SM; the 512-thread blocks of round 1
// only found room on SMs that hosted a single GEMM CTA, so an overlapped bucket crawled until the backward was over
constexpr int kCommThreads = 128;

// perf experiments: globaltimer stamps of block 0 / thread 0 of the sync kernel (tools/sync_phase.py)
static int g_cstamps_host = 0;                // passed as a kernel argument: no device-side flag load on the hot path
__device__ unsigned long long g_cstamps[16];
#define DDL_CSTAMP(i) do { if (cst) g_cstamps[i] = ddl_globaltimer(); } while (0)

DDL_DEVICE float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// all ranks: write `epoch` into slot [me] of every peer's pad, then wait until every slot of my own
// pad has reached `epoch`.  Called by ONE thread block; thread r talks to rank r.
DDL_DEVICE void pad_signal(const DdlComm& c, uint32_t base, uint32_t epoch) {
  if ((int)threadIdx.x < c.world) st_release_sys(c.pads[threadIdx.x] + base + c.rank, epoch);
}
// Same signal without the release: a st.release.sys costs a MEMBAR.SYS (~2 us at the head of a kernel, measured with
// tools/sync_phase.py).  Two uses, both already ordered by other means:
//  * barrier A - what the signal publishes (this rank's gradients, and "I am done reading my parameters") was written
//    by EARLIER kernels of this stream; a kernel boundary completes and flushes them to L2, the point of coherence a
//    peer's NVLink access reads, before this kernel's first instruction;
//  * barrier B - the caller is the single thread that has just executed the fences that order this grid's stores.
DDL_DEVICE void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DDL_DEVICE void pad_signal_relaxed(const DdlComm& c, uint32_t base, uint32_t epoch) {
  if ((int)threadIdx.x < c.world) st_relaxed_sys(c.pads[threadIdx.x] + base + c.rank, epoch);
}
// signal rows of the sync buckets: one row of DDL_MAX_PEERS slots per bucket, so buckets never see each other's epochs
// and their kernels need no ordering among themselves
DDL_DEVICE uint32_t sync_pad_base(int bucket) { return (uint32_t)(DDL_PAD_SYNC + DDL_MAX_PEERS * bucket); }
DDL_DEVICE void pad_wait(const DdlComm& c, uint32_t base, uint32_t epoch) {
  if ((int)threadIdx.x < c.world) {
    const uint32_t* slot = c.pads[c.rank] + base + threadIdx.x;
    uint32_t spins = 0;
    uint64_t t0 = 0;
    while ((int32_t)(ld_acquire_sys(slot) - epoch) < 0) {
      if ((++spins & 0x3ff) == 0) {
        const uint64_t now = ddl_globaltimer();
        if (t0 == 0) t0 = now;
        else if (now - t0 > 5 * DDL_WAIT_TIMEOUT_NS) {
          printf("[ddl] rank %d: peer %d never reached epoch %u\n", c.rank, (int)threadIdx.x, epoch);
          __trap();
        }
      }
    }
  }
}

// Consume 16 bytes of a worker's accumulator plane: read, then subtract exactly what was read with ONE vector red.add.
// The pushes counted by the snapshot have fully landed (their counter bump is release-ordered behind their reds), so the
// value read holds all of them; adds of a LATER push that race with this survive the subtraction (x + a - x), which is what
// "every push applied exactly once" needs - four scalar atomicExch per float4 did the same at 4x the atomic traffic
// (61 us for the apply of a 2.66 M parameter shard, profiles/ncu_full_r2_step_summary.csv).
DDL_DEVICE float4 take_plane(float* pl) {
  const float4 g = __ldcg(reinterpret_cast<const float4*>(pl));
  if (g.x != 0.f || g.y != 0.f || g.z != 0.f || g.w != 0.f) red_add_v4_f32(pl, make_float4(-g.x, -g.y, -g.z, -g.w));
  return g;
}

// optimizer arithmetic on registers: returns the new parameters, updates mv / vv in place
template <int KIND>
DDL_DEVICE float4 opt_math(float4 pv, float4 gv, float4& mv, float4& vv, float step, float a, float b, float eps) {
  float pp[4] = {pv.x, pv.y, pv.z, pv.w};
  float gg[4] = {gv.x, gv.y, gv.z, gv.w};
  float mm[4] = {mv.x, mv.y, mv.z, mv.w};
  float vs[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (KIND == 0) {
      pp[k] -= step * gg[k];
    } else if (KIND == 1) {
      mm[k] = a * mm[k] + gg[k];
      pp[k] -= step * mm[k];
    } else {
      mm[k] = a * mm[k] + (1.f - a) * gg[k];
      vs[k] = b * vs[k] + (1.f - b) * gg[k] * gg[k];
      pp[k] -= step * mm[k] / (sqrtf(vs[k]) + eps);
    }
  }
  mv = make_float4(mm[0], mm[1], mm[2], mm[3]);
  vv = make_float4(vs[0], vs[1], vs[2], vs[3]);
  return make_float4(pp[0], pp[1], pp[2], pp[3]);
}

template <int KIND>
DDL_DEVICE float4 opt_apply(float4 pv, float4 gv, float* m, float* v, int64_t i, float step, float a, float b,
                            float eps) {
  float4 mv = make_float4(0.f, 0.f, 0.f, 0.f), vv = mv;
  if (KIND >= 1) mv = *reinterpret_cast<float4*>(m + i);
  if (KIND == 2) vv = *reinterpret_cast<float4*>(v + i);
  const float4 pn = opt_math<KIND>(pv, gv, mv, vv, step, a, b, eps);
  if (KIND >= 1) *reinterpret_cast<float4*>(m + i) = mv;
  if (KIND == 2) *reinterpret_cast<float4*>(v + i) = vv;
  return pn;
}

// ------------------------------------------------------------------------------------------------
// K1 + K3 (sync): reduce-scatter + scale + optimizer + all-gather in one kernel.
// m / v are indexed like the arena (only [begin, end) is touched).  `skip` lists arena ranges whose
// new values are NOT broadcast because their first consumer GEMM pulls them itself (K4).
// ------------------------------------------------------------------------------------------------
// U float4 per thread per trip: every load of a trip is issued before the first use, so a thread keeps
// U x (workers) NVLink round trips (~2 us each) in flight.  U = 1 for latency-bound small ranges (more SMs, one
// trip), U = 4 for bandwidth-bound large ones.
template <int KIND, int U>
__global__ void __launch_bounds__(kCommThreads)
sync_push_update_pull_kernel(const DdlComm c, const DdlSyncArgs a, const int stamps_on, const int skip_a) {
  pdl_enter();
  __shared__ uint32_t s_last;
#ifdef DDL_TC_LAB
  const bool cst = stamps_on && blockIdx.x == 0 && threadIdx.x == 0 && a.bucket == stamps_on - 1;
#else
  const bool cst = false;                          // phase stamps exist only in the lab build (ops/build.py --lab)
#endif
  DDL_CSTAMP(0);
  const uint32_t t = a.st->t;                       // same on every rank: all ranks advance together
  const uint32_t nb = a.nbuckets > 1 ? (uint32_t)a.nbuckets : 1u;
  const uint32_t epochA = 2u * nb * t + 2u * (uint32_t)a.bucket + 1u, epochB = epochA + 1u;
  const uint32_t pbase = sync_pad_base(a.bucket);
  // ---- barrier A: my gradients are complete (previous kernels on this stream) -> tell everyone (see
  // pad_signal_relaxed).  A single rank has nobody to wait for: no signals, no system fences.
  const bool solo = c.world == 1;
  float* myp = c.params[c.rank];
  const int nsub = a.nsub > 0 ? a.nsub : 1;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // One flat index space over all sub-ranges of the bucket (float4 units): a thread's trips are spread over the whole
  // bucket, not restarted per sub-range - a bucket of four small variables is ONE NVLink round trip, not four
  // dependent ones.
  int64_t n4 = 0;
  for (int sr = 0; sr < nsub; ++sr) n4 += ((a.nsub > 0 ? a.sub_end[sr] : a.end) - (a.nsub > 0 ? a.sub_begin[sr] : a.begin)) / 4;
  auto elem = [&](int64_t q) -> int64_t {          // arena element index of flat float4 index q (q < n4)
    if (a.nsub <= 0) return a.begin + q * 4;
    int64_t i = a.sub_begin[0] + q * 4;
#pragma unroll 1
    for (int sr = 0; sr < a.nsub; ++sr) {
      const int64_t len = (a.sub_end[sr] - a.sub_begin[sr]) / 4;
      if (q < len) { i = a.sub_begin[sr] + q * 4; break; }
      q -= len;
    }
    return i;
  };
  if (!solo && !skip_a) {                           // skip_a: sync_wait_kernel (previous launch on this stream) did it
    if (blockIdx.x == 0) pad_signal_relaxed(c, pbase, epochA);
    DDL_CSTAMP(1);
    pad_wait(c, pbase, epochA);                     // every block polls its own (local) pad
    __syncthreads();
  }
  DDL_CSTAMP(2);

  constexpr int WCH = (U == 1) ? DDL_MAX_PEERS : 4;  // peers loaded per batch: WCH * U float4 in flight
  const float step = (KIND == 2) ? a.st->lr_t : a.lr;
  for (int64_t q0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q0 < n4; q0 += stride * U) {
    int64_t ei[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t q = q0 + u * stride;
      ei[u] = elem(q < n4 ? q : q0);
    }
    // optimizer state first: local HBM loads that overlap the NVLink round trip of the gradient loads below (and all U
    // chunks before the first store: m / v / p may alias for the compiler)
    float4 pv[U], mv[U], vv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t i = ei[u];
      pv[u] = *reinterpret_cast<const float4*>(myp + i);
      mv[u] = vv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (KIND >= 1) mv[u] = *reinterpret_cast<const float4*>(a.m + i);
      if (KIND == 2) vv[u] = *reinterpret_cast<const float4*>(a.v + i);
    }
    float4 g[U];
    if (c.mc_grads != nullptr) {
      // NVLS: the switch sums this offset of every rank's gradient arena (ranks that are not workers hold
      // zeros) - one load, owner ingress 1x instead of Wx.  Summation order inside the switch is unspecified.
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t q = q0 + u * stride;
        g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q < n4) g[u] = multimem_ld_reduce_add_v4(c.mc_grads + ei[u]);
      }
    } else {
      // all peer loads of a batch first, then a fixed-order sum (ascending rank): loads overlap instead of
      // W dependent round trips, and the result is bitwise reproducible
#pragma unroll
      for (int u = 0; u < U; ++u) g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int w0 = 0; w0 < c.world; w0 += WCH) {
        float4 x[WCH][U];
#pragma unroll
        for (int wi = 0; wi < WCH; ++wi) {
          const int w = w0 + wi;
          const bool on = w < c.world && ((a.worker_mask >> w) & 1u);
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int64_t q = q0 + u * stride;
            x[wi][u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (on && q < n4) x[wi][u] = ldcg4(c.grads[w] + ei[u]);
          }
        }
#pragma unroll
        for (int wi = 0; wi < WCH; ++wi) {
#pragma unroll
          for (int u = 0; u < U; ++u) {
            g[u].x += x[wi][u].x; g[u].y += x[wi][u].y; g[u].z += x[wi][u].z; g[u].w += x[wi][u].w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t q = q0 + u * stride;
      if (q >= n4) continue;
      const int64_t i = ei[u];
      float4 gs = g[u];
      gs.x *= a.gscale; gs.y *= a.gscale; gs.z *= a.gscale; gs.w *= a.gscale;
      const float4 pn = opt_math<KIND>(pv[u], gs, mv[u], vv[u], step, a.a, a.b, a.eps);
      if (KIND >= 1) *reinterpret_cast<float4*>(a.m + i) = mv[u];
      if (KIND == 2) *reinterpret_cast<float4*>(a.v + i) = vv[u];
      *reinterpret_cast<float4*>(myp + i) = pn;
      bool bcast = true;
#pragma unroll
      for (int k = 0; k < DDL_MAX_SKIP; ++k) bcast = bcast && !(i >= a.skip_begin[k] && i < a.skip_end[k]);
      if (bcast) {
        if (c.mc_params != nullptr) {
          multimem_st_v4(c.mc_params + i, pn);      // NVLS: one store, the switch writes every rank's copy
        } else {
#pragma unroll 1
          for (int d = 0; d < c.world; ++d) {
            if (d == c.rank || !((a.dest_mask >> d) & 1u)) continue;
            __stcg(reinterpret_cast<float4*>(c.params[d] + i), pn);   // peer store: the pull, fused
          }
        }
      }
    }
  }
  // ---- barrier B: last block to finish publishes "my range is delivered" and waits for the others.
  // One fence per block, not per thread: bar.sync makes the block's stores observed by thread 0, whose
  // (cumulative) fence then orders them before the counter bump - the cooperative-groups grid-sync idiom.
  // A system-scope fence is only needed when stores went to peers.
  DDL_CSTAMP(3);
  if (solo) return;                                // the kernel boundary orders everything for the next kernel
  __syncthreads();
  DDL_CSTAMP(4);
  if (threadIdx.x == 0) {
    __threadfence_system();
    DDL_CSTAMP(5);
    s_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    DDL_CSTAMP(6);
  }
  __syncthreads();
  if (s_last) {
    // Every block executed fence.sys (its remote stores are performed system-wide) and then bumped the counter; thread 0
    // of this block saw the last bump.  A gpu-scope fence turns that relaxed observation into an acquire (it pairs with
    // the other blocks' fences), and the same thread then issues the "delivered" signals - fence, then flag, from one
    // thread: the message-passing pattern without a second MEMBAR.SYS (1.9 us, tools/sync_phase.py).
    if (threadIdx.x == 0) {
      *a.done_counter = 0;
      __threadfence();
      for (int r = 0; r < c.world; ++r) st_relaxed_sys(c.pads[r] + pbase + c.rank, epochB);
    }
    pad_wait(c, pbase, epochB);
  }
  DDL_CSTAMP(7);
}

// Barrier A of an OVERLAPPED bucket as its own one-block launch: "my gradients of this bucket are complete" to every
// rank, then wait for everybody's.  The update kernel behind it is launched with a full (non-programmatic) dependency
// and skip_a = 1, so none of its blocks is resident - holding threads and registers next to the backward's GEMM
// kernels - while peers are still computing (round 1 parked 148 x 512 spinning threads there).
__global__ void __launch_bounds__(32) sync_wait_kernel(const DdlComm c, const DdlSyncArgs a) {
  pdl_enter();
  const uint32_t nb = a.nbuckets > 1 ? (uint32_t)a.nbuckets : 1u;
  const uint32_t epochA = 2u * nb * a.st->t + 2u * (uint32_t)a.bucket + 1u;
  pad_signal_relaxed(c, sync_pad_base(a.bucket), epochA);
  pad_wait(c, sync_pad_base(a.bucket), epochA);
}

// ------------------------------------------------------------------------------------------------
// K2 (async): worker pushes its whole gradient arena to the owners with remote atomics.
// direct_sgd: params[owner][i] += -lr * g[i]  (Hogwild, no owner-side work at all);
// otherwise:  accum[owner][slot][i] += g[i], slot = this worker's index: every worker has its own accumulator plane on
//             every owner, so the owner can apply ONE optimizer step PER PUSH (ddl_async_apply), as the reference's
//             PS does (mnist_async/parameter_server.py:55-69), instead of one step on the sum of whatever has landed.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCommThreads)
async_push_kernel(const DdlComm c, const DdlAsyncArgs a) {
  pdl_enter();
  __shared__ uint32_t s_last;
  const float* g = c.grads[c.rank];
  for (int o = 0; o < a.num_shards; ++o) {
    const int owner = a.owner_rank[o];
    float* dst = a.direct_sgd ? c.params[owner] : c.accum[owner] + (int64_t)a.slot * a.slot_stride;
    const float scale = a.direct_sgd ? -a.lr : 1.f;
    const int64_t b = a.shard_begin[o], n4 = (a.shard_end[o] - b) / 4;
    for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < n4; q += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = b + q * 4;
      float4 x = *reinterpret_cast<const float4*>(g + i);
      if (x.x == 0.f && x.y == 0.f && x.z == 0.f && x.w == 0.f) continue;
      x.x *= scale; x.y *= scale; x.z *= scale; x.w *= scale;
      red_add_v4_f32(dst + i, x);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();                      // this block's remote reds are performed (cumulative after bar.sync)
    s_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x == 0) { *a.done_counter = 0; __threadfence_system(); }
    __syncthreads();
    // one release-ordered counter bump per owner: "a complete push of worker `slot` has landed"
    if ((int)threadIdx.x < a.num_shards) {
      const int owner = a.owner_rank[threadIdx.x];
      const uint32_t seen = ld_acquire_sys(c.pads[owner] + DDL_PAD_VERSION);
      atom_add_release_sys(c.pads[owner] + DDL_PAD_PUSHES_W + a.slot, 1u);
      atom_add_release_sys(c.pads[owner] + DDL_PAD_PUSHES, 1u);
      // staleness of this push = optimizer steps the owner has taken since our last pull
      const uint32_t stale = seen - a.pulled_version[threadIdx.x];
      atomicAdd(&a.stale_hist[stale < DDL_STALE_BINS - 1 ? stale : DDL_STALE_BINS - 1], 1u);
    }
  }
}

// owner, 1 thread: which workers have pushes pending, how many optimizer steps that makes, and the Adam step size of
// each.  Parity mode: one step per worker with a pending push, in worker order (a worker that pushed twice since the
// last apply has both gradients summed in its plane - counted in `folded`).  fold mode: one step on the sum of all.
__global__ void async_snapshot_kernel(const DdlComm c, DdlStepState* st, DdlAsyncState* as, int nworkers, int fold,
                                      float lr, float b1, float b2) {
  pdl_enter();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    uint32_t mask = 0, nsteps = 0;
    uint32_t t = st->t;
    for (int w = 0; w < nworkers; ++w) {
      const uint32_t cnt = ld_acquire_sys(c.pads[c.rank] + DDL_PAD_PUSHES_W + w);
      const uint32_t n = cnt - as->applied[w];
      if (n) {
        as->applied[w] = cnt;
        mask |= 1u << w;
        as->folded += n - 1;
        if (!fold || nsteps == 0) {
          ++t;
          as->lr_t[nsteps] = (float)((double)lr * sqrt(1.0 - pow((double)b2, (double)t)) / (1.0 - pow((double)b1, (double)t)));
          ++nsteps;
        } else {
          as->folded += 1;                       // fold mode: this worker's push shares the step of an earlier one
        }
      }
    }
    as->mask = mask;
    as->nsteps = nsteps;
    if (nsteps) { st->t = t; st->lr_t = as->lr_t[nsteps - 1]; }
  }
}

template <int KIND>
__global__ void __launch_bounds__(kCommThreads)
async_apply_kernel(const DdlComm c, const DdlSyncArgs a, const DdlAsyncState* as, int nworkers, int64_t slot_stride, int fold) {
  pdl_enter();
  __shared__ uint32_t s_last;
  const uint32_t mask = as->mask;
  if (mask == 0) return;
  float* myp = c.params[c.rank];
  float* acc = c.accum[c.rank];
  const int64_t n4 = (a.end - a.begin) / 4;
  for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < n4; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = a.begin + q * 4;
    float4 pv = *reinterpret_cast<float4*>(myp + i);
    float4 mv = make_float4(0.f, 0.f, 0.f, 0.f), vv = mv;
    if (KIND >= 1) mv = *reinterpret_cast<float4*>(a.m + i);
    if (KIND == 2) vv = *reinterpret_cast<float4*>(a.v + i);
    // every pending plane is consumed exactly once (atomicExch), even against adds still in flight for a LATER push;
    // all exchanges are issued before the first optimizer step
    float4 g[DDL_MAX_PEERS];
#pragma unroll
    for (int w = 0; w < DDL_MAX_PEERS; ++w) {
      g[w] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (w < nworkers && ((mask >> w) & 1u)) {
        float* pl = acc + (int64_t)w * slot_stride + i;
        g[w] = take_plane(pl);
      }
    }
    if (fold) {
      float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < DDL_MAX_PEERS; ++w) { gs.x += g[w].x; gs.y += g[w].y; gs.z += g[w].z; gs.w += g[w].w; }
      gs.x *= a.gscale; gs.y *= a.gscale; gs.z *= a.gscale; gs.w *= a.gscale;
      pv = opt_math<KIND>(pv, gs, mv, vv, KIND == 2 ? as->lr_t[0] : a.lr, a.a, a.b, a.eps);
    } else {
      int k = 0;
#pragma unroll
      for (int w = 0; w < DDL_MAX_PEERS; ++w) {
        if (w < nworkers && ((mask >> w) & 1u)) {
          float4 gs = g[w];
          gs.x *= a.gscale; gs.y *= a.gscale; gs.z *= a.gscale; gs.w *= a.gscale;
          pv = opt_math<KIND>(pv, gs, mv, vv, KIND == 2 ? as->lr_t[k] : a.lr, a.a, a.b, a.eps);
          ++k;
        }
      }
    }
    if (KIND >= 1) *reinterpret_cast<float4*>(a.m + i) = mv;
    if (KIND == 2) *reinterpret_cast<float4*>(a.v + i) = vv;
    *reinterpret_cast<float4*>(myp + i) = pv;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    s_last = (atomicAdd(a.done_counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    *a.done_counter = 0;
    __threadfence_system();
    atom_add_release_sys(c.pads[c.rank] + DDL_PAD_VERSION, as->nsteps);
  }
}

// ------------------------------------------------------------------------------------------------
// Resident owner (dedicated async PS GPU): ONE launch that stays on the device and serves pushes as they land - the
// reference's ``while`` loop of mnist_async/parameter_server.py:52-69, without a host in it.  Round r:
//   block 0 / thread 0 polls the per-worker push counters (and the host's stop word); when something is pending it
//   takes the snapshot (same rules as async_snapshot_kernel), publishes round r with a release store;
//   every block waits for round r, applies its slice (one optimizer step per pending push, planes consumed with
//   atomicExch), bumps the done counter; block 0 collects the bumps, publishes the shard version (+ nsteps) and goes
//   back to polling.
// All blocks must be co-resident (the host sizes the grid from the occupancy API).  Idle polling is bounded:
// DDL_WAIT_TIMEOUT_NS x 16 without a push or a stop traps.
// ------------------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(kCommThreads)
async_owner_resident_kernel(const DdlComm c, const DdlSyncArgs a, DdlAsyncState* as, DdlResident* rs, const int nworkers,
                            const int64_t slot_stride, const int fold, const float b1, const float b2) {
  __shared__ uint32_t s_stop;
  float* myp = c.params[c.rank];
  float* acc = c.accum[c.rank];
  const int64_t n4 = (a.end - a.begin) / 4;
  for (uint32_t round = 1;; ++round) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      uint32_t spins = 0, mask = 0, nsteps = 0, stop = 0;
      uint64_t t0 = 0;
      for (;;) {
        uint32_t t = a.st->t;
        mask = 0; nsteps = 0;
        for (int w = 0; w < nworkers; ++w) {
          const uint32_t cnt = ld_acquire_sys(c.pads[c.rank] + DDL_PAD_PUSHES_W + w);
          const uint32_t n = cnt - as->applied[w];
          if (n) {
            as->applied[w] = cnt;
            mask |= 1u << w;
            as->folded += n - 1;
            if (!fold || nsteps == 0) {
              ++t;
              as->lr_t[nsteps] = (float)((double)a.lr * sqrt(1.0 - pow((double)b2, (double)t)) / (1.0 - pow((double)b1, (double)t)));
              ++nsteps;
            } else {
              as->folded += 1;
            }
          }
        }
        if (mask) { a.st->t = t; a.st->lr_t = as->lr_t[nsteps - 1]; break; }
        if ((++spins & 0x3f) == 0) {
          uint32_t hv;                                   // the host's stop word lives in pinned host memory
          asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(hv) : "l"(rs->host_stop) : "memory");
          if (hv) { stop = 1; break; }
          const uint64_t now = ddl_globaltimer();
          if (t0 == 0) t0 = now;
          else if (now - t0 > 16 * DDL_WAIT_TIMEOUT_NS) { printf("[ddl] resident owner %d: idle too long\n", c.rank); __trap(); }
        }
      }
      as->mask = mask;
      as->nsteps = nsteps;
      rs->stop = stop;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&rs->round), "r"(round) : "memory");
    }
    if (threadIdx.x == 0) {
      uint32_t seen;
      do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(&rs->round) : "memory"); } while (seen != round);
      s_stop = rs->stop;
    }
    __syncthreads();
    if (s_stop) return;
    const uint32_t mask = as->mask;
    for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < n4; q += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = a.begin + q * 4;
      float4 pv = *reinterpret_cast<float4*>(myp + i);
      float4 mv = make_float4(0.f, 0.f, 0.f, 0.f), vv = mv;
      if (KIND >= 1) mv = *reinterpret_cast<float4*>(a.m + i);
      if (KIND == 2) vv = *reinterpret_cast<float4*>(a.v + i);
      float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
      int k = 0;
#pragma unroll 1
      for (int w = 0; w < nworkers; ++w) {
        if (!((mask >> w) & 1u)) continue;
        float* pl = acc + (int64_t)w * slot_stride + i;
        float4 g = take_plane(pl);
        if (fold) { gs.x += g.x; gs.y += g.y; gs.z += g.z; gs.w += g.w; continue; }
        g.x *= a.gscale; g.y *= a.gscale; g.z *= a.gscale; g.w *= a.gscale;
        pv = opt_math<KIND>(pv, g, mv, vv, KIND == 2 ? as->lr_t[k] : a.lr, a.a, a.b, a.eps);
        ++k;
      }
      if (fold) {
        gs.x *= a.gscale; gs.y *= a.gscale; gs.z *= a.gscale; gs.w *= a.gscale;
        pv = opt_math<KIND>(pv, gs, mv, vv, KIND == 2 ? as->lr_t[0] : a.lr, a.a, a.b, a.eps);
      }
      if (KIND >= 1) *reinterpret_cast<float4*>(a.m + i) = mv;
      if (KIND == 2) *reinterpret_cast<float4*>(a.v + i) = vv;
      *reinterpret_cast<float4*>(myp + i) = pv;
    }
    // grid barrier: every block's stores are fenced before its bump; block 0 publishes the new version after the last one
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      atomicAdd(&rs->done, 1u);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      uint32_t seen;
      do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(&rs->done) : "memory"); } while (seen != gridDim.x * round);
      __threadfence_system();
      atom_add_release_sys(c.pads[c.rank] + DDL_PAD_VERSION, as->nsteps);
      rs->rounds_served = round;
    }
  }
}

// K3: worker copies every owner's range into its own parameter arena (peer loads, no barrier).
__global__ void __launch_bounds__(kCommThreads)
pull_allgather_kernel(const DdlComm c, const DdlAsyncArgs a) {
  pdl_enter();
  float* myp = c.params[c.rank];
  for (int o = 0; o < a.num_shards; ++o) {
    const int owner = a.owner_rank[o];
    if (blockIdx.x == 0 && threadIdx.x == 0 && a.pulled_version)
      a.pulled_version[o] = ld_acquire_sys(c.pads[owner] + DDL_PAD_VERSION);
    if (owner == c.rank) continue;
    const float* src = c.params[owner];
    const int64_t b = a.shard_begin[o], n4 = (a.shard_end[o] - b) / 4;
    // 4 peer loads in flight per thread per trip
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t q0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q0 < n4; q0 += 4 * stride) {
      float4 x[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u * stride < n4) x[u] = ldcg4(src + b + (q0 + u * stride) * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u) if (q0 + u * stride < n4) *reinterpret_cast<float4*>(myp + b + (q0 + u * stride) * 4) = x[u];
    }
  }
}

__global__ void device_barrier_kernel(const DdlComm c, uint32_t* epoch_counter) {
  pdl_enter();
  __shared__ uint32_t e;
  if (threadIdx.x == 0) e = ++(*epoch_counter);
  __syncthreads();
  __threadfence_system();
  pad_signal(c, DDL_PAD_BARRIER, e);
  pad_wait(c, DDL_PAD_BARRIER, e);
}

// NVLink probe (tools/p2p_probe.py): one thread issues `iters` DEPENDENT 4-byte loads from `src` (a peer-mapped or local
// pointer; the value loaded selects the next address, so no two are in flight), sleeping `idle_ns` before each one.
// out[0] = total ns spent in the loads, out[1] = ns of the first load, out[2] = max, out[3] = checksum.
__global__ void p2p_latency_kernel(const uint32_t* src, int words, int iters, unsigned idle_ns, unsigned long long* out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  unsigned long long total = 0, first = 0, worst = 0;
  uint32_t idx = 0, sum = 0;
  for (int i = 0; i < iters; ++i) {
    if (idle_ns) __nanosleep(idle_ns);
    const unsigned long long t0 = ddl_globaltimer();
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src + idx) : "memory");
    idx = (idx + 97u * 32u + (v & 1u)) % (uint32_t)words;       // depends on v: the next load cannot start early
    sum += v;
    const unsigned long long dt = ddl_globaltimer() - t0;
    total += dt;
    if (i == 0) first = dt;
    if (dt > worst) worst = dt;
  }
  out[0] = total; out[1] = first; out[2] = worst; out[3] = sum;
}

// one thread: `iters` 4-byte stores to `dst`, each followed by fence.sys (returns once the store is performed at its
// destination): the store + acknowledgement round trip that a "delivered" signal has to wait for
__global__ void p2p_store_fence_kernel(uint32_t* dst, int words, int iters, unsigned idle_ns, unsigned long long* out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  unsigned long long total = 0, first = 0, worst = 0;
  for (int i = 0; i < iters; ++i) {
    if (idle_ns) __nanosleep(idle_ns);
    const unsigned long long t0 = ddl_globaltimer();
    st_relaxed_sys(dst + (i * 97 * 32) % words, (uint32_t)i);
    __threadfence_system();
    const unsigned long long dt = ddl_globaltimer() - t0;
    total += dt;
    if (i == 0) first = dt;
    if (dt > worst) worst = dt;
  }
  out[0] = total; out[1] = first; out[2] = worst; out[3] = 0;
}

// probe: `gridDim.x` blocks, two threads each, poll a flag with ld.acquire.sys for `ns` nanoseconds - what the blocks of a
// sync kernel do while they wait for a barrier; run next to the latency probes to see what the polling costs others
__global__ void p2p_spin_kernel(const uint32_t* flag, unsigned long long ns, unsigned long long* out) {
  if (threadIdx.x >= 2) return;
  const unsigned long long t0 = ddl_globaltimer();
  unsigned long long polls = 0;
  while (ddl_globaltimer() - t0 < ns) { if (ld_acquire_sys(flag + threadIdx.x) == 0xffffffffu) break; ++polls; }
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = polls;
}

// bandwidth probe: grid-stride float4 copy dst <- src (either may be a peer pointer), 4 loads in flight per thread
__global__ void __launch_bounds__(256) p2p_copy_kernel(float4* __restrict__ dst, const float4* __restrict__ src, int64_t n4) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t q0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q0 < n4; q0 += 4 * stride) {
    float4 x[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) if (q0 + u * stride < n4) x[u] = __ldcg(src + q0 + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) if (q0 + u * stride < n4) __stcg(dst + q0 + u * stride, x[u]);
  }
}

int comm_grid(int64_t elems) {
  // ~2 float4 per thread: the update is a latency-bound stream (7 arena-sized passes), so it wants many threads in
  // flight rather than a long grid-stride loop
  int64_t g = (elems / 4 + 2 * kCommThreads - 1) / (2 * kCommThreads);
  if (g < 1) g = 1;
  if (g > 148 * 16) g = 148 * 16;
  return (int)g;
}

}  // namespace

extern "C" {

/* perf experiments: enable (on = bucket + 1, 0 = off) / read the globaltimer stamps of block 0 of that bucket's sync kernel */
int ddl_comm_stamps(int on, unsigned long long* out, int n) {
  g_cstamps_host = on;
  if (out && n > 0) DDL_CHECK_CUDA(cudaMemcpyFromSymbol(out, g_cstamps, sizeof(unsigned long long) * (n > 16 ? 16 : n)));
  return 0;
}

/* NVLink probes (tools/p2p_probe.py -> profiles/nvlink_peak.json); out4 is device memory (4 x uint64) */
int ddl_p2p_latency(const void* src, int words, int iters, unsigned idle_ns, int store_fence, unsigned long long* out4,
                    cudaStream_t s) {
  DDL_REQUIRE(words >= 4096 && iters >= 1, "ddl_p2p_latency: need >= 4096 words and >= 1 iteration");
  if (store_fence) p2p_store_fence_kernel<<<1, 32, 0, s>>>((uint32_t*)src, words, iters, idle_ns, out4);
  else p2p_latency_kernel<<<1, 32, 0, s>>>((const uint32_t*)src, words, iters, idle_ns, out4);
  DDL_LAUNCH_CHECK();
  return 0;
}
int ddl_p2p_spin(const void* flag, int blocks, unsigned long long ns, unsigned long long* out1, cudaStream_t s) {
  p2p_spin_kernel<<<blocks, 32, 0, s>>>((const uint32_t*)flag, ns, out1);
  DDL_LAUNCH_CHECK();
  return 0;
}
int ddl_p2p_copy(void* dst, const void* src, int64_t bytes, int blocks, cudaStream_t s) {
  DDL_REQUIRE(bytes % 16 == 0 && blocks >= 1, "ddl_p2p_copy: bytes %% 16 == 0, blocks >= 1");
  p2p_copy_kernel<<<blocks, 256, 0, s>>>((float4*)dst, (const float4*)src, bytes / 16);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_dev_malloc(void** p, size_t bytes) {
  DDL_CHECK_CUDA(cudaMalloc(p, bytes));
  DDL_CHECK_CUDA(cudaMemset(*p, 0, bytes));
  return 0;
}
int ddl_dev_free(void* p) { DDL_CHECK_CUDA(cudaFree(p)); return 0; }
int ddl_ipc_export(void* p, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
  DDL_CHECK_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), p));
  return 0;
}
int ddl_ipc_open(const void* handle64, void** p) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  DDL_CHECK_CUDA(cudaIpcOpenMemHandle(p, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int ddl_ipc_close(void* p) { DDL_CHECK_CUDA(cudaIpcCloseMemHandle(p)); return 0; }
int ddl_set_device(int dev) { DDL_CHECK_CUDA(cudaSetDevice(dev)); return 0; }
int ddl_can_access_peer(int dev, int peer) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) return 0;
  return ok;
}

int ddl_sync_push_update_pull(const DdlComm* c, const DdlSyncArgs* a, int kind, cudaStream_t s) {
  DDL_REQUIRE(c->world >= 1 && c->world <= DDL_MAX_PEERS, "world %d out of range", c->world);
  DDL_REQUIRE(a->begin % 4 == 0 && a->end % 4 == 0, "shard range must be float4 aligned");
  DDL_REQUIRE(a->nsub >= 0 && a->nsub <= DDL_MAX_SUB, "nsub %d out of range", a->nsub);
  int64_t elems = a->end - a->begin;
  if (a->nsub > 0) {
    elems = 0;
    for (int i = 0; i < a->nsub; ++i) {
      DDL_REQUIRE(a->sub_begin[i] % 4 == 0 && a->sub_end[i] % 4 == 0, "sub-range must be float4 aligned");
      elems += a->sub_end[i] - a->sub_begin[i];
    }
  }
  int grid = comm_grid(elems);
  // a bucket that overlaps the backward shares the SMs with the GEMM kernels of the critical chain: keep it narrow
  static const int cap = ddl_lab_int("DDL_COMM_CAP", 592);
  const bool overlapped = a->nbuckets > 1 && a->bucket < a->nbuckets - 1;
  if (overlapped && grid > cap) grid = cap;
  // bandwidth-bound ranges (>= 16 float4 per thread at the chosen grid): 4 float4 per thread per trip; smaller
  // ranges are latency-bound, and multimem loads measured faster with one (profiles/comm_roofline.md)
  static const int force_u = ddl_lab_int("DDL_COMM_UNROLL", 0);
  const bool wide = force_u ? force_u > 1 : (c->mc_grads == nullptr && elems / 4 >= (int64_t)grid * kCommThreads * 16);
  int skip_a = 0;
  if (overlapped && c->world > 1) {
    DDL_LAUNCH(sync_wait_kernel, 1, 32, 0, s, *c, *a);
    ddl::no_pdl_once();                    // full dependency: no block of the update kernel is resident during the wait
    skip_a = 1;
  }
  if (wide) {
    if (kind == 0) DDL_LAUNCH((sync_push_update_pull_kernel<0, 4>), grid, kCommThreads, 0, s, *c, *a, g_cstamps_host, skip_a);
    else if (kind == 1) DDL_LAUNCH((sync_push_update_pull_kernel<1, 4>), grid, kCommThreads, 0, s, *c, *a, g_cstamps_host, skip_a);
    else DDL_LAUNCH((sync_push_update_pull_kernel<2, 4>), grid, kCommThreads, 0, s, *c, *a, g_cstamps_host, skip_a);
  } else {
    if (kind == 0) DDL_LAUNCH((sync_push_update_pull_kernel<0, 1>), grid, kCommThreads, 0, s, *c, *a, g_cstamps_host, skip_a);
    else if (kind == 1) DDL_LAUNCH((sync_push_update_pull_kernel<1, 1>), grid, kCommThreads, 0, s, *c, *a, g_cstamps_host, skip_a);
    else DDL_LAUNCH((sync_push_update_pull_kernel<2, 1>), grid, kCommThreads, 0, s, *c, *a, g_cstamps_host, skip_a);
  }
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_async_push(const DdlComm* c, const DdlAsyncArgs* a, cudaStream_t s) {
  int64_t total = 0;
  for (int o = 0; o < a->num_shards; ++o) total += a->shard_end[o] - a->shard_begin[o];
  DDL_LAUNCH(async_push_kernel, comm_grid(total), kCommThreads, 0, s, *c, *a);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_async_apply(const DdlComm* c, const DdlSyncArgs* a, int kind, DdlAsyncState* state, int nworkers,
                    int64_t slot_stride, int fold, float b1, float b2, cudaStream_t s) {
  DDL_REQUIRE(nworkers >= 1 && nworkers <= DDL_MAX_PEERS, "ddl_async_apply: %d workers out of range", nworkers);
  DDL_LAUNCH(async_snapshot_kernel, 1, 32, 0, s, *c, a->st, state, nworkers, fold, a->lr, b1, b2);
  const int grid = comm_grid(a->end - a->begin);
  if (kind == 0) DDL_LAUNCH(async_apply_kernel<0>, grid, kCommThreads, 0, s, *c, *a, (const DdlAsyncState*)state, nworkers, slot_stride, fold);
  else if (kind == 1) DDL_LAUNCH(async_apply_kernel<1>, grid, kCommThreads, 0, s, *c, *a, (const DdlAsyncState*)state, nworkers, slot_stride, fold);
  else DDL_LAUNCH(async_apply_kernel<2>, grid, kCommThreads, 0, s, *c, *a, (const DdlAsyncState*)state, nworkers, slot_stride, fold);
  DDL_LAUNCH_CHECK();
  return 0;
}

/* Starts the resident owner kernel on stream s (a stream of its own: it runs until *host_stop != 0).  The grid is sized
 * so that every block is resident at once (occupancy API), at most two blocks per SM. */
int ddl_async_owner_start(const DdlComm* c, const DdlSyncArgs* a, int kind, DdlAsyncState* state, DdlResident* resident,
                          int nworkers, int64_t slot_stride, int fold, float b1, float b2, cudaStream_t s) {
  DDL_REQUIRE(nworkers >= 1 && nworkers <= DDL_MAX_PEERS, "ddl_async_owner_start: %d workers out of range", nworkers);
  int dev = 0, sms = 0, per_sm = 0;
  DDL_CHECK_CUDA(cudaGetDevice(&dev));
  DDL_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const void* fn = kind == 0 ? (const void*)async_owner_resident_kernel<0>
                 : kind == 1 ? (const void*)async_owner_resident_kernel<1> : (const void*)async_owner_resident_kernel<2>;
  DDL_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kCommThreads, 0));
  DDL_REQUIRE(per_sm >= 1, "ddl_async_owner_start: the resident kernel does not fit an SM");
  int grid = sms * (per_sm > 2 ? 2 : per_sm);
  const int want = comm_grid(a->end - a->begin);
  if (grid > want) grid = want;
  if (kind == 0) async_owner_resident_kernel<0><<<grid, kCommThreads, 0, s>>>(*c, *a, state, resident, nworkers, slot_stride, fold, b1, b2);
  else if (kind == 1) async_owner_resident_kernel<1><<<grid, kCommThreads, 0, s>>>(*c, *a, state, resident, nworkers, slot_stride, fold, b1, b2);
  else async_owner_resident_kernel<2><<<grid, kCommThreads, 0, s>>>(*c, *a, state, resident, nworkers, slot_stride, fold, b1, b2);
  DDL_LAUNCH_CHECK();
  ddl::note_launch();
  return 0;
}

int ddl_pull_allgather(const DdlComm* c, const DdlAsyncArgs* a, cudaStream_t s) {
  int64_t total = 0;
  for (int o = 0; o < a->num_shards; ++o) total += a->shard_end[o] - a->shard_begin[o];
  DDL_LAUNCH(pull_allgather_kernel, comm_grid(total), kCommThreads, 0, s, *c, *a);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_device_barrier(const DdlComm* c, uint32_t* epoch_counter, cudaStream_t s) {
  DDL_LAUNCH(device_barrier_kernel, 1, 32, 0, s, *c, epoch_counter);
  DDL_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
