// tcgen05 / TMEM / TMA implicit-GEMM kernels for sm_100a: conv fprop, conv dgrad, conv wgrad and the
// FC matmuls (fwd, dX, dW) of the reference CNN - every GEMM-shaped op with >= 32 reduction channels.
//
// Reference ops replaced: tf.nn.conv2d x3 (mnist_sync/model/model.py:39,50,61) and their
// backprop-input / backprop-filter ops (implied by compute_gradients, model.py:101), tf.matmul
// fc1/fc2 forward and backward (model.py:70,79).  conv1 (Cin = 1) and fc3 (N = 10) are SIMT kernels
// in elementwise.cu: their shapes cannot feed a 128xNx8 UMMA usefully (and a [*,10] fp32 row is
// not a legal 16-byte TMA stride).
//
// Design (B200-first, not a port of anything):
//   * data stays fp32 in HBM; the tensor core consumes it as kind::tf32 (fp32 accumulate in TMEM);
//   * the im2col operand is never materialised: one 4-D TMA box [32 ch, bw, bh, bn] of the NHWC
//     activation, shifted by the filter tap, IS the 128-row K-major A tile of that tap.  Out-of-
//     bounds rows/columns are zero-filled by TMA, which is exactly SAME padding;
//   * weights (HWIO) are read in place: [Cin, Cout] per tap is an MN-major B tile for fprop and a
//     K-major B tile for dgrad - no transposed copies exist anywhere;
//   * wgrad reduces over pixels, so both operands (dY, shifted X) are MN-major boxes; partial sums
//     from pixel splits go straight into the gradient arena with red.global.add;
//   * warp-specialised CTA: warp 0 = TMA producer, warp 1 = MMA issuer (one lane) + TMEM owner,
//     warps 2..5 = epilogue (tcgen05.ld -> bias / ReLU / dropout / mask -> global);
//   * SWIZZLE_128B everywhere: a 32-float row is one 128-byte swizzle span.
#include "common.cuh"
#include "../include/ddl_b200.h"

#include <mutex>
#include <stdlib.h>
#include <string.h>
#include <new>


// Lab instrumentation (clock64 / globaltimer traces, skip-MMA / skip-TMA switches used by tools/tc_*.py and
// tools/wgrad_trace.py) is compiled in only with -DDDL_TC_LAB (ops/build.py --lab -> libddl_b200_lab.so); in the
// product library both macros are compile-time constants and every instrumented branch disappears.
#ifdef DDL_TC_LAB
#define LAB_DEBUG(p) ((p).debug)
#define LAB_TRACE(p) ((p).trace)
#else
#define LAB_DEBUG(p) 0
#define LAB_TRACE(p) (static_cast<unsigned long long*>(nullptr))
#endif

namespace {

constexpr int kTileM = 128;        // UMMA M (TMEM lanes)
constexpr int kKB = 32;            // fp32 elements per K-block = one 128-byte swizzle row
constexpr int kUmmaK = 8;          // tf32: 32 bytes of K per instruction
constexpr int kThreadsTc = 192;    // 6 warps
constexpr int kMaxStages = 8;

struct ActParams {
  // pixel space of the activation operand and of the output
  int N, H, W;
  int bw, bh, bn;                  // TMA pixel box; rows per tile = bw*bh*bn <= 128
  int tiles_w, tiles_h, tiles_n;
  int ksize, flip;                 // filter size (5 or 1); flip = dgrad (offset negated)
  int cblocks;                     // ceil(reduction channels / 32)
  int kiters;                      // ksize*ksize*cblocks
  int ksplit;                      // gridDim.z
  int nout, ntile;                 // output channels, UMMA N
  int stages;
  int kps;                         // K blocks (32 reduction channels each) per pipeline stage: 1 or 2.  Two halve the
                                   // barrier round trips (wait + commit ~175 of ~600 cycles per K block, tc_issue_latency.md)
  int b_mn_major;                  // 1: weights [tap][cred][nout] (fprop), 0: [tap][nout][cred] (dgrad)
  int flags;
  int cluster;                     // > 1: the ksplit CTAs of a tile form a thread-block cluster; ranks 1.. park their
                                   // partial accumulators in `scratch`, rank 0 adds them and runs the fused epilogue
  float* scratch;                  // [tiles][cluster-1][ntile/4][128] float4 partial tiles (plan-owned)
  int pull;                        // K4: weight tiles come from another rank's arena (map_w on a peer pointer);
                                   // pixel-tile 0 CTAs write every tile through to the local arena (map_wst)
  int debug;                       // perf experiments only: 1 = skip MMA issue, 2 = skip TMA loads
  unsigned long long* trace;       // perf experiments only: per-iteration clock64 stamps of CTA (0,0,0)
  float* out;                      // [N*H*W, nout]
  const float* bias;               // [nout]
  float* colsum;                   // [nout]; EPI_COLSUM adds the column sums of the final values (bias gradient)
  const float* mask_act;           // [N*H*W, nout]; EPI_MASKPOS multiplies by (mask_act > 0) * mask_scale
  float mask_scale;
  const DdlStepState* st;          // dropout seed / stream
  uint32_t drop_stream_off;
  uint32_t drop_thresh24;
  float drop_inv_keep;
  // EPI_POOL / EPI_UNPOOL (staged epilogues, see the kernel)
  int bars_off;                    // byte offset of PipeBars (past the pipeline stages AND the staging areas)
  float* pool_out;                 // EPI_POOL: [N, ceil(H/2), ceil(W/2), nout]; EPI_UNPOOL: [N, UH, UW, nout]
  uint8_t* pool_idx;               // argmax slots: written by EPI_POOL (nullable), read by EPI_UNPOOL
  int UH, UW;                      // EPI_UNPOOL: un-pooled plane (H = ceil(UH/2), W = ceil(UW/2))
  float* dbias;                    // EPI_UNPOOL: += column sums of the routed gradient (nullable)
};

struct WgradParams {
  int N, H, W;
  int bw, bh, bn;                  // pixel box, rows R = bw*bh*bn, R % 8 == 0
  int tiles_w, tiles_h, tiles_n;
  int ksize;
  int ca, cb;                      // channels of the gradient operand (M side) and of the input operand (N side)
  int mtiles, ntiles, ntile;
  int tpc, tgroups;                // filter taps handled by one CTA (concatenated along the MMA N dimension)
  int a_alloc;                     // A boxes that own shared memory (>= live boxes)
  int psplit;
  int stages;
  float* out;                      // [taps][cb][ca], accumulated with red.add
  unsigned long long* trace;       // perf experiments (DDL_TC_TRACE): per-iteration clock64 stamps of CTA (0,0,0)
};

enum { EPI_BIAS = 1, EPI_RELU = 2, EPI_DROPOUT = 4, EPI_MASKPOS = 8, EPI_ATOMIC = 16, EPI_COLSUM = 32,
       EPI_POOL = 64, EPI_UNPOOL = 128,
       EPI_STAGED = 256,
       EPI_POOLGRAD = 512 };   // cluster split-K with a fused bias / ReLU / dropout / mask / colsum epilogue: the partial tiles are
                             // reduce-scattered through distributed shared memory and every CTA finishes its own columns
                             // (instead of parking them in global scratch for the cluster leader)

struct alignas(8) PipeBars {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full;
  uint32_t tmem_addr;
  uint32_t pad;
};

// 4x4 transpose across a quad of lanes (two butterfly rounds, 4 shuffles): in r[c] = (this lane, column c),
// out r[k] = (quad lane k, column ql).
__device__ __forceinline__ void quad_transpose4(float r[4], int ql) {
  const bool b0 = ql & 1, b1 = ql & 2;
  float sx = b0 ? r[0] : r[1], sy = b0 ? r[2] : r[3];
  float rx = __shfl_xor_sync(0xffffffffu, sx, 1), ry = __shfl_xor_sync(0xffffffffu, sy, 1);
  if (b0) { r[0] = rx; r[2] = ry; } else { r[1] = rx; r[3] = ry; }
  sx = b1 ? r[0] : r[2]; sy = b1 ? r[1] : r[3];
  rx = __shfl_xor_sync(0xffffffffu, sx, 2); ry = __shfl_xor_sync(0xffffffffu, sy, 2);
  if (b1) { r[0] = rx; r[1] = ry; } else { r[2] = rx; r[3] = ry; }
}

// Column position inside a staging row of cw floats: channel quads are de-interleaved (c -> (c % 4) * cw/4 + c/4) so
// that a thread finishing 4 consecutive channels reads 4 conflict-free words (lanes = consecutive quads).
__device__ __forceinline__ int spos(int c, int cw) { return (c & 3) * (cw >> 2) + (c >> 2); }

__device__ __forceinline__ uint32_t tmem_cols_for(int n) {
  return n <= 32 ? 32u : n <= 64 ? 64u : n <= 128 ? 128u : 256u;
}

// ------------------------------------------------------------------------------------------------
// act-GEMM: out[pix, n] = epi( sum_{tap, c} act[pix + off(tap), c] * Wt[tap][c][n] )
// grid = (pixel tiles, n tiles, ksplit)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreadsTc, 1)
tc_act_gemm_kernel(const __grid_constant__ CUtensorMap map_act, const __grid_constant__ CUtensorMap map_w,
                   const __grid_constant__ CUtensorMap map_wst, const ActParams p) {
  pdl_launch_dependents();                 // the next kernel may start its prologue now
  const bool ptrace = (LAB_DEBUG(p) & 16) && LAB_TRACE(p) && blockIdx.x + blockIdx.y + blockIdx.z == 0;
  if (ptrace && threadIdx.x == 0) LAB_TRACE(p)[0] = ddl_globaltimer();
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the __shared__ array (an integer round trip would turn every later
  // access into a generic LD.E / ST.E instead of LDS / STS)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int a_bytes = kTileM * 128;
  const int b_bytes = p.ntile * 128;
  const int stage_bytes = p.kps * (a_bytes + b_bytes);    // [A x kps | B x kps]
  const int sb_off = p.kps * a_bytes;                     // first B block of a stage
  PipeBars* bars = (PipeBars*)(smem + (size_t)p.bars_off);

  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int tw = tile % p.tiles_w, th = (tile / p.tiles_w) % p.tiles_h, tn = tile / (p.tiles_w * p.tiles_h);
  const int w0 = tw * p.bw, h0 = th * p.bh, n0 = tn * p.bn;
  const int nt0 = blockIdx.y * p.ntile;
  const int per = (p.kiters + p.ksplit - 1) / p.ksplit;
  const int k_begin = blockIdx.z * per;
  const int k_end = min(p.kiters, k_begin + per);
  const int niter = max(0, k_end - k_begin);              // K blocks of this CTA
  const int nst = (niter + p.kps - 1) / p.kps;            // pipeline stage fills (kps K blocks each, the last may be short)
  const uint32_t ncols = tmem_cols_for(p.ntile);
  const bool creduce = p.cluster > 1;
  const bool staged = (p.flags & (EPI_POOL | EPI_UNPOOL | EPI_STAGED)) != 0;
  const uint32_t crank = creduce ? cluster_ctarank() : 0u;
  const int tile_lin = blockIdx.y * gridDim.x + blockIdx.x;

  // K4 write-through: one CTA per weight tile (the first pixel tile) persists what it pulled
  const bool store_through = p.pull && blockIdx.x == 0 && p.b_mn_major;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], store_through ? 2 : 1);      // MMA commit (+ the store warp's arrive)
    }
    mbar_init(&bars->tmem_full, 1);
    fence_barrier_init();
    tma_prefetch_desc(&map_act);
    tma_prefetch_desc(&map_w);
    if (store_through) tma_prefetch_desc(&map_wst);
  }
  if (warp == 1) tmem_alloc(&bars->tmem_addr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_addr;
  if (ptrace && threadIdx.x == 0) LAB_TRACE(p)[1] = ddl_globaltimer();
  // everything above overlapped the previous kernel; its data (the activation operand) is needed from here.  The
  // weight operand is not produced by the preceding kernels (see ddl::no_pdl_once), so the B producer (warp 3) starts
  // filling the pipeline right away and only joins the wait before its epilogue duty.
  const bool early_b = !p.pull;
  if (!(early_b && warp == 3)) pdl_wait();
  if (ptrace && threadIdx.x == 0) LAB_TRACE(p)[2] = ddl_globaltimer();

  if (warp == 2 && store_through) {
    // ================= K4 store warp: smem weight tile -> local arena (TMA store), before its epilogue duty
    int s = 0;
    uint32_t ph = 0;
    int tap = k_begin / p.cblocks, cb = k_begin % p.cblocks;
    for (int i = 0; i < niter; ++i) {
      mbar_wait(&bars->full[s], ph);
      if (elect_one()) {
        uint8_t* sb = smem + (size_t)s * stage_bytes + a_bytes;
        for (int j = 0; j < p.ntile / 32; ++j)
          tma_store_4d(&map_wst, sb + j * 4096, nt0 + 32 * j, cb * kKB, tap, 0);
        tma_store_commit();
        tma_store_wait_read0();                  // smem may be overwritten once the store has read it
        mbar_arrive(&bars->empty[s]);
      }
      __syncwarp();
      if (++cb == p.cblocks) { cb = 0; ++tap; }
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
    if (elect_one()) tma_store_wait_all0();      // writes complete before the kernel can end
    __syncwarp();
  }

  if (warp == 3 && !(LAB_DEBUG(p) & 2)) {
    // ================= second TMA producer: the weight (B) boxes.  One UTMALDG of a 128-row box keeps the
    // issuing lane busy for ~200 cycles, so A and B are issued from two warps in parallel; the transaction
    // bytes of both are announced by warp 0's expect_tx (a complete_tx that lands first just drives the
    // count negative until then - the phase cannot complete before warp 0's arrive).
    int s = 0;
    uint32_t ph = 0;
    int tap = k_begin / p.cblocks, cb = k_begin % p.cblocks;
    for (int i = 0; i < nst; ++i) {
      const int cnt = min(p.kps, niter - i * p.kps);
      mbar_wait(&bars->empty[s], ph ^ 1u);
      const bool el = elect_one();
      for (int u = 0; u < cnt; ++u) {
        if (el) {
          uint8_t* sb = smem + (size_t)s * stage_bytes + sb_off + u * b_bytes;
          if (p.b_mn_major) {
            for (int j = 0; j < p.ntile / 32; ++j)
              tma_load_4d(sb + j * 4096, &map_w, &bars->full[s], nt0 + 32 * j, cb * kKB, tap, 0);
          } else {
            tma_load_4d(sb, &map_w, &bars->full[s], cb * kKB, nt0, tap, 0);
          }
        }
        if (++cb == p.cblocks) { cb = 0; ++tap; }
      }
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
  }
  if (early_b && warp == 3) pdl_wait();

  if (warp == 0) {
    // ================= TMA producer (warp-uniform loop, one elected lane issues) =================
    const int box_rows = p.bw * p.bh * p.bn;
    const uint32_t tx1 = (uint32_t)(box_rows * 128 + b_bytes);          // bytes of one K block (A box + B boxes)
    const int half = p.ksize / 2;
    int s = 0;
    uint32_t ph = 0;
    int tap = k_begin / p.cblocks, cb = k_begin % p.cblocks;
    for (int i = 0; i < nst; ++i) {
      const int cnt = min(p.kps, niter - i * p.kps);
      const bool ptr = (LAB_DEBUG(p) & 8) && LAB_TRACE(p) && blockIdx.x + blockIdx.y + blockIdx.z == 0 && i < 512;
      if (ptr && lane == 0) LAB_TRACE(p)[i * 4 + 0] = clock64();
      mbar_wait(&bars->empty[s], ph ^ 1u);
      const bool el = elect_one();
      if (el) {
        if (ptr) LAB_TRACE(p)[i * 4 + 1] = clock64();
        if (!(LAB_DEBUG(p) & 24) && LAB_TRACE(p) && blockIdx.x + blockIdx.y + blockIdx.z == 0 && i < 512) LAB_TRACE(p)[i * 4 + 0] = clock64();
        if (LAB_DEBUG(p) & 2) mbar_arrive(&bars->full[s]);
        else mbar_expect_tx(&bars->full[s], tx1 * (uint32_t)cnt);   // covers the A boxes (here) and the B boxes (warp 3)
        if (ptr) LAB_TRACE(p)[i * 4 + 2] = clock64();
      }
      for (int u = 0; u < cnt; ++u) {
        if (el && !(LAB_DEBUG(p) & 2)) {
          int dh = tap / p.ksize - half, dw = tap % p.ksize - half;
          if (p.flip) { dh = -dh; dw = -dw; }
          tma_load_4d(smem + (size_t)s * stage_bytes + u * a_bytes, &map_act, &bars->full[s], cb * kKB, w0 + dw, h0 + dh, n0);
        }
        if (++cb == p.cblocks) { cb = 0; ++tap; }
      }
      if (el && ptr) LAB_TRACE(p)[i * 4 + 3] = clock64();
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (warp-uniform loop, one elected lane issues) =================
    const uint32_t idesc = umma_idesc_tf32(kTileM, p.ntile, 0, p.b_mn_major);
    const uint32_t smem_base = smem_u32(smem);
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < nst; ++i) {
      const int cnt = min(p.kps, niter - i * p.kps);
      const bool tr = !(LAB_DEBUG(p) & 24) && LAB_TRACE(p) && blockIdx.x + blockIdx.y + blockIdx.z == 0 && i < 512 && lane == 0;
      if (tr) LAB_TRACE(p)[i * 4 + 1] = clock64();
      mbar_wait(&bars->full[s], ph);
      if (ptrace && lane == 0 && i == 0) LAB_TRACE(p)[3] = ddl_globaltimer();
      if (tr) LAB_TRACE(p)[i * 4 + 2] = clock64();
      tc_fence_after();
      if (elect_one()) {
        const uint32_t st0 = smem_base + (uint32_t)(s * stage_bytes);
        if (!(LAB_DEBUG(p) & 1)) {
          for (int u = 0; u < cnt; ++u) {
            const uint32_t sa = st0 + (uint32_t)(u * a_bytes);
            const uint32_t sb = st0 + (uint32_t)(sb_off + u * b_bytes);
#pragma unroll
            for (int k = 0; k < kKB / kUmmaK; ++k) {
              const uint64_t ad = umma_smem_desc(sa + k * 32, 16, 1024, DDL_UMMA_SW128);
              const uint64_t bd = p.b_mn_major ? umma_smem_desc(sb + k * 1024, 4096, 512, DDL_UMMA_SW128_BASE32B)
                                               : umma_smem_desc(sb + k * 32, 16, 1024, DDL_UMMA_SW128);
              umma_tf32(tmem_base, ad, bd, idesc, (i > 0 || u > 0 || k > 0) ? 1u : 0u);
            }
          }
        }
        umma_commit(&bars->empty[s]);            // smem slot reusable once these MMAs have read it
        if (i == nst - 1) umma_commit(&bars->tmem_full);   // accumulator complete
        if (ptrace && i == nst - 1) LAB_TRACE(p)[4] = ddl_globaltimer();
      }
      __syncwarp();
      if (tr) LAB_TRACE(p)[i * 4 + 3] = clock64();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
    if (niter == 0 && elect_one()) umma_commit(&bars->tmem_full);
    __syncwarp();
  } else if (staged) {
    // ================= staged epilogue, phase 1 (warps 2..5): accumulator -> shared memory.
    // The pipeline stages are dead once tmem_full fires (every MMA has read its operands), so they are reused:
    //   cluster > 1: D[col][row] (column-major, 128 rows) - the partial tile the cluster's CTAs read from each other;
    //   cluster == 1: straight into S[row][cw + 1] with the bias added.
    const int q = warp & 3;
    const int row = q * 32 + lane;
    if (lane == 0) mbar_wait(&bars->tmem_full, 0);
    __syncwarp();
    tc_fence_after();
    float* D = reinterpret_cast<float*>(smem);
    const int sp = p.ntile / p.cluster + 1;
    for (int c0 = 0; c0 < p.ntile; c0 += 16) {
      float v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      if (niter == 0) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0.f;
      }
      if (creduce) {
#pragma unroll
        for (int j = 0; j < 16; ++j) D[(c0 + j) * 128 + row] = v[j];       // a warp writes 128 contiguous bytes
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = nt0 + c0 + j;
          const float b = ((p.flags & EPI_POOL) && col < p.nout) ? __ldg(p.bias + col) : 0.f;
          D[row * sp + spos(c0 + j, p.ntile)] = v[j] + b;                   // odd pitch: conflict-free
        }
      }
    }
  } else if (creduce && crank != 0) {
    // ================= split-K peer (warps 2..5): park the partial accumulator for the cluster leader
    const int q = warp & 3;
    const int row = q * 32 + lane;
    if (lane == 0) mbar_wait(&bars->tmem_full, 0);
    __syncwarp();
    tc_fence_after();
    float4* part = reinterpret_cast<float4*>(p.scratch) +
                   ((size_t)(tile_lin * (p.cluster - 1) + (crank - 1)) * (p.ntile / 4)) * 128 + row;
    for (int c0 = 0; c0 < p.ntile; c0 += 16) {
      float v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      if (niter == 0) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0.f;
      }
#pragma unroll
      for (int j = 0; j < 16; j += 4)       // [col/4][row]: a warp writes 512 contiguous bytes per store
        __stcg(part + (size_t)((c0 + j) / 4) * 128, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
    }
    __threadfence();
  }
  if (ptrace && threadIdx.x == 64) LAB_TRACE(p)[5] = ddl_globaltimer();
  if (creduce) {                             // every thread of every CTA of the cluster
    cluster_arrive_release();
    cluster_wait_acquire();
  }
  if (ptrace && threadIdx.x == 64) LAB_TRACE(p)[6] = ddl_globaltimer();
  if (staged) {
    if (warp >= 2) {
      // ================= staged epilogue, phases 2 + 3 (warps 2..5, 128 threads)
      // phase 2 (cluster > 1): reduce-scatter over distributed shared memory - rank r sums columns
      //   [r * cw, (r + 1) * cw) of all CTAs' D (fixed order: reproducible), + bias, into its own S[row][cw + 1];
      // phase 3: EPI_POOL   - 2x2 stride-2 SAME max-pool with the running max started at 0 (= ReLU) -> pool_out, pool_idx
      //          EPI_UNPOOL - route each value to its window's argmax slot of the un-pooled plane, zeros elsewhere
      //                       (ReLU-dead windows: idx 4 -> all zeros), column sums -> dbias.
      const int te = (int)threadIdx.x - 64;
      const int cw = p.ntile / p.cluster, sp = cw + 1;
      const int colbase = nt0 + (int)crank * cw;
      float* S = reinterpret_cast<float*>(smem) + (creduce ? (size_t)p.ntile * 128 : 0);
      if (creduce) {
        const int rg = te & 31, cq = te >> 5;
        uint32_t peer[8];
        const uint32_t d0 = smem_u32(smem);
#pragma unroll
        for (int k = 0; k < 8; ++k) peer[k] = cluster_map_shared(d0, (uint32_t)(k < p.cluster ? k : 0));
        for (int c = cq; c < cw; c += 4) {
          const uint32_t off = (uint32_t)((((int)crank * cw + c) * 128 + 4 * rg) * 4);
          float4 t[8];
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (k < p.cluster) t[k] = ld_dsmem_v4(peer[k] + off);
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (k < p.cluster) { acc.x += t[k].x; acc.y += t[k].y; acc.z += t[k].z; acc.w += t[k].w; }
          const bool addb = (p.flags & EPI_POOL) || ((p.flags & EPI_STAGED) && (p.flags & EPI_BIAS));
          const float b = (addb && colbase + c < p.nout) ? __ldg(p.bias + colbase + c) : 0.f;
          const int cp = spos(c, cw);
          S[(4 * rg + 0) * sp + cp] = acc.x + b;
          S[(4 * rg + 1) * sp + cp] = acc.y + b;
          S[(4 * rg + 2) * sp + cp] = acc.z + b;
          S[(4 * rg + 3) * sp + cp] = acc.w + b;
        }
      }
      named_bar_sync(1, 128);
      const int cq4 = cw >> 2;                   // channel quads per staging row; a thread finishes one quad per item
      float* __restrict__ pout = p.pool_out;
      if (p.flags & EPI_STAGED) {
        // generic fused epilogue on my columns: (bias already added) ReLU, dropout, ReLU/dropout-backward mask, column
        // sums, float4 stores
        const int R = p.bw * p.bh * p.bn;
        const int items = R * cq4;
        const bool fixed_col = (128 % cq4) == 0;
        uint32_t seed = 0, stream = 0;
        if (p.flags & EPI_DROPOUT) { seed = p.st->seed; stream = p.st->stream + p.drop_stream_off; }
        float csum[4] = {0.f, 0.f, 0.f, 0.f};
        for (int it = te; it < items; it += 128) {
          const int c4 = it % cq4, r = it / cq4;
          const int rw = r % p.bw, rh = (r / p.bw) % p.bh, rn = r / (p.bw * p.bh);
          const int gw = w0 + rw, gh = h0 + rh, gn = n0 + rn;
          const int col = colbase + 4 * c4;
          if (gn >= p.N || gh >= p.H || gw >= p.W || col >= p.nout) continue;
          const int64_t pix = ((int64_t)gn * p.H + gh) * p.W + gw;
          const float* sr = S + r * sp + c4;
          float x[4];
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) x[q4] = sr[q4 * cq4];
          if (p.flags & EPI_RELU) {
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) x[q4] = fmaxf(x[q4], 0.f);
          }
          if (p.flags & EPI_DROPOUT) {
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const uint32_t e = (uint32_t)(pix * p.nout + col + q4);
              x[q4] = ddl_dropout_keep(seed, stream, e, p.drop_thresh24) ? x[q4] * p.drop_inv_keep : 0.f;
            }
          }
          if (p.flags & EPI_MASKPOS) {
            const float4 m = *reinterpret_cast<const float4*>(p.mask_act + pix * p.nout + col);
            x[0] = m.x > 0.f ? x[0] * p.mask_scale : 0.f;
            x[1] = m.y > 0.f ? x[1] * p.mask_scale : 0.f;
            x[2] = m.z > 0.f ? x[2] * p.mask_scale : 0.f;
            x[3] = m.w > 0.f ? x[3] * p.mask_scale : 0.f;
          }
          *reinterpret_cast<float4*>(p.out + pix * p.nout + col) = make_float4(x[0], x[1], x[2], x[3]);
          if (p.flags & EPI_COLSUM) {
            if (fixed_col) { csum[0] += x[0]; csum[1] += x[1]; csum[2] += x[2]; csum[3] += x[3]; }
            else {
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) if (x[q4] != 0.f) atomicAdd(p.colsum + col + q4, x[q4]);
            }
          }
        }
        if ((p.flags & EPI_COLSUM) && fixed_col) {
          const bool pow2 = (cq4 & (cq4 - 1)) == 0;
          if (pow2) {
            for (int off = cq4; off < 32; off <<= 1) {
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) csum[q4] += __shfl_xor_sync(0xffffffffu, csum[q4], off);
            }
          }
          const int col = colbase + 4 * (te % cq4);
          if (!pow2 || cq4 >= 32 || lane < cq4) {
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4)
              if (csum[q4] != 0.f && col + q4 < p.nout) atomicAdd(p.colsum + col + q4, csum[q4]);
          }
        }
      } else if (p.flags & EPI_POOL) {
        uint8_t* __restrict__ pidx = p.pool_idx;
        const int PH = (p.H + 1) / 2, PW = (p.W + 1) / 2;
        const int pbw = (p.bw + 1) / 2, pbh = (p.bh + 1) / 2;
        const int items = pbw * pbh * p.bn * cq4;
        for (int it = te; it < items; it += 128) {
          const int c4 = it % cq4, pp = it / cq4;
          const int prw = pp % pbw, prh = (pp / pbw) % pbh, prn = pp / (pbw * pbh);
          const int gpw = w0 / 2 + prw, gph = h0 / 2 + prh, gn = n0 + prn;
          const int col = colbase + 4 * c4;
          if (gn >= p.N || gph >= PH || gpw >= PW || col >= p.nout) continue;
          float best[4] = {0.f, 0.f, 0.f, 0.f};
          int bi[4] = {4, 4, 4, 4};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int lh = 2 * prh + (k >> 1), lw = 2 * prw + (k & 1);
            if (lh < p.bh && lw < p.bw && h0 + lh < p.H && w0 + lw < p.W) {
              const float* sr = S + ((prn * p.bh + lh) * p.bw + lw) * sp + c4;
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) {
                const float x = sr[q4 * cq4];
                if (x > best[q4]) { best[q4] = x; bi[q4] = k; }
              }
            }
          }
          const int64_t o = (((int64_t)gn * PH + gph) * PW + gpw) * p.nout + col;
          *reinterpret_cast<float4*>(pout + o) = make_float4(best[0], best[1], best[2], best[3]);
          if (pidx) *reinterpret_cast<uchar4*>(pidx + o) = make_uchar4(bi[0], bi[1], bi[2], bi[3]);
        }
      } else {
        const uint8_t* __restrict__ pidx = p.pool_idx;
        const int R = p.bw * p.bh * p.bn;
        const int items = R * cq4;
        const bool fixed_col = (128 % cq4) == 0;     // then it % cq4 is the same on every trip of a thread
        float dbsum[4] = {0.f, 0.f, 0.f, 0.f};
        // two items per trip, every global load (argmax slots) issued before the first store
        for (int it0 = te; it0 < items; it0 += 256) {
          uchar4 bi[2];
          float g[2][4];
          int64_t obase[2];
          int gh[2], gw[2], colv[2];
          bool ok[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int it = it0 + u * 128;
            const int c4 = it % cq4, r = it / cq4;
            const int rw = r % p.bw, rh = (r / p.bw) % p.bh, rn = r / (p.bw * p.bh);
            const int gn = n0 + rn;
            gw[u] = w0 + rw; gh[u] = h0 + rh;
            colv[u] = colbase + 4 * c4;
            ok[u] = it < items && gn < p.N && gh[u] < p.H && gw[u] < p.W && colv[u] < p.nout;
            bi[u] = make_uchar4(4, 4, 4, 4);
            if (ok[u]) {
              bi[u] = *reinterpret_cast<const uchar4*>(pidx + (((int64_t)gn * p.H + gh[u]) * p.W + gw[u]) * p.nout + colv[u]);
              const float* sr = S + r * sp + c4;
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) g[u][q4] = sr[q4 * cq4];
            }
            obase[u] = (int64_t)gn * p.UH;
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            if (!ok[u]) continue;
            const int b4[4] = {bi[u].x, bi[u].y, bi[u].z, bi[u].w};
            if (p.flags & EPI_POOLGRAD) {
              // compact variant (plan sets UH = H, UW = W): the gradient stays in the pooled layout, ReLU-dead windows zeroed (the
              // layer below has no data gradient, its weight gradient reads value + argmax slot: conv1_wgrad_pooled)
              *reinterpret_cast<float4*>(pout + ((obase[u] + gh[u]) * p.UW + gw[u]) * p.nout + colv[u]) =
                  make_float4(b4[0] < 4 ? g[u][0] : 0.f, b4[1] < 4 ? g[u][1] : 0.f, b4[2] < 4 ? g[u][2] : 0.f,
                              b4[3] < 4 ? g[u][3] : 0.f);
            } else
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int uh = 2 * gh[u] + (k >> 1), uw = 2 * gw[u] + (k & 1);
              if (pout && uh < p.UH && uw < p.UW)
                *reinterpret_cast<float4*>(pout + ((obase[u] + uh) * p.UW + uw) * p.nout + colv[u]) =
                    make_float4(b4[0] == k ? g[u][0] : 0.f, b4[1] == k ? g[u][1] : 0.f, b4[2] == k ? g[u][2] : 0.f,
                                b4[3] == k ? g[u][3] : 0.f);
            }
            if (p.dbias) {
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) {
                if (b4[q4] < 4) {
                  if (fixed_col) dbsum[q4] += g[u][q4];
                  else if (g[u][q4] != 0.f) atomicAdd(p.dbias + colv[u] + q4, g[u][q4]);
                }
              }
            }
          }
        }
        if (p.dbias && fixed_col) {
          // lanes cq4 apart hold the same channel quad: fold them with shuffles, one atomic per warp and channel
          const bool pow2 = (cq4 & (cq4 - 1)) == 0;
          if (pow2) {
            for (int off = cq4; off < 32; off <<= 1) {
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) dbsum[q4] += __shfl_xor_sync(0xffffffffu, dbsum[q4], off);
            }
          }
          const int col = colbase + 4 * (te % cq4);
          if (!pow2 || cq4 >= 32 || lane < cq4) {
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4)
              if (dbsum[q4] != 0.f && col + q4 < p.nout) atomicAdd(p.dbias + col + q4, dbsum[q4]);
          }
        }
      }
    }
    if (creduce) {                           // nobody leaves (or frees its shared memory) while a peer may still read D
      cluster_arrive_release();
      cluster_wait_acquire();
    }
  } else if (warp >= 2 && (!creduce || crank == 0)) {
    // ================= epilogue (warps 2..5) =================
    const int q = warp & 3;                      // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;
    const int rw = row % p.bw, rh = (row / p.bw) % p.bh, rn = row / (p.bw * p.bh);
    const int pw = w0 + rw, phh = h0 + rh, pn = n0 + rn;
    const bool valid = (rn < p.bn) && pw < p.W && phh < p.H && pn < p.N && (niter > 0 || creduce);
    const int64_t pix = ((int64_t)pn * p.H + phh) * p.W + pw;
    float* orow = p.out + pix * p.nout + nt0;
    const float* mrow = (p.flags & EPI_MASKPOS) ? p.mask_act + pix * p.nout + nt0 : nullptr;
    uint32_t seed = 0, stream = 0;
    if (p.flags & EPI_DROPOUT) { seed = p.st->seed; stream = p.st->stream + p.drop_stream_off; }
    if (lane == 0) mbar_wait(&bars->tmem_full, 0);      // one poller per warp: 128 threads hammering the
    __syncwarp();                                        // same mbarrier slow the producer / MMA handoffs
    if (ptrace && threadIdx.x == 64) LAB_TRACE(p)[7] = ddl_globaltimer();
    tc_fence_after();
    const float4* parts = creduce ? reinterpret_cast<const float4*>(p.scratch) +
                                        ((size_t)(tile_lin * (p.cluster - 1)) * (p.ntile / 4)) * 128 + row
                                  : nullptr;
    for (int c0 = 0; c0 < p.ntile; c0 += 16) {
      float v[16];
      if (LAB_DEBUG(p) & 64) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = (float)j;
      } else {
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      }
      if (ptrace && threadIdx.x == 64 && c0 == 0) LAB_TRACE(p)[10] = ddl_globaltimer();
      if (creduce) {                         // add the peers' partial tiles (L2-resident, coalesced)
        if (niter == 0) {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = 0.f;
        }
        // four peers per batch, every load of a batch issued before the first add: one L2 round trip per batch
        // instead of one per peer (the rolled loop cost (cluster - 1) dependent latencies per chunk, tools/tc_phase.py)
        for (int r0 = 0; r0 < p.cluster - 1; r0 += 4) {
          float4 t[4][4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              t[k][j] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (r0 + k < p.cluster - 1) t[k][j] = __ldcg(parts + ((size_t)(r0 + k) * (p.ntile / 4) + c0 / 4 + j) * 128);
            }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              v[4 * j] += t[k][j].x; v[4 * j + 1] += t[k][j].y; v[4 * j + 2] += t[k][j].z; v[4 * j + 3] += t[k][j].w;
            }
          }
        }
      }
      if (p.flags & EPI_ATOMIC) {
        if (valid) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            if (nt0 + c0 + j < p.nout) red_add_v4_f32(orow + c0 + j, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
          }
        }
        continue;
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int col = nt0 + c0 + j;
        float x = v[j];
        if (p.flags & EPI_BIAS) x += (col < p.nout) ? __ldg(p.bias + col) : 0.f;
        if (p.flags & EPI_RELU) x = fmaxf(x, 0.f);
        if (p.flags & EPI_DROPOUT) {
          const uint32_t e = (uint32_t)(pix * p.nout + col);
          x = ddl_dropout_keep(seed, stream, e, p.drop_thresh24) ? x * p.drop_inv_keep : 0.f;
        }
        v[j] = x;
      }
      if ((p.flags & EPI_MASKPOS) && valid) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          if (nt0 + c0 + j < p.nout) {
            const float4 m = *reinterpret_cast<const float4*>(mrow + c0 + j);
            v[j] = m.x > 0.f ? v[j] * p.mask_scale : 0.f;
            v[j + 1] = m.y > 0.f ? v[j + 1] * p.mask_scale : 0.f;
            v[j + 2] = m.z > 0.f ? v[j + 2] * p.mask_scale : 0.f;
            v[j + 3] = m.w > 0.f ? v[j + 3] * p.mask_scale : 0.f;
          }
        }
      }
      if (p.flags & EPI_COLSUM) {              // warp-wide column sums of this warp's 32 rows
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float sv = valid ? v[j] : 0.f;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) sv += __shfl_xor_sync(0xffffffffu, sv, o);
          if (lane == j && nt0 + c0 + j < p.nout) atomicAdd(p.colsum + nt0 + c0 + j, sv);
        }
      }
      if (ptrace && threadIdx.x == 64 && c0 == 0) LAB_TRACE(p)[11] = ddl_globaltimer();
      if (valid && !(LAB_DEBUG(p) & 32)) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          if (nt0 + c0 + j < p.nout)
            *reinterpret_cast<float4*>(orow + c0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
      }
      if (ptrace && threadIdx.x == 64 && c0 == 0) LAB_TRACE(p)[12] = ddl_globaltimer();
    }
  }
  if (ptrace && threadIdx.x == 64) LAB_TRACE(p)[8] = ddl_globaltimer();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ncols);
  if (ptrace && threadIdx.x == 32) LAB_TRACE(p)[9] = ddl_globaltimer();
}

// ------------------------------------------------------------------------------------------------
// wgrad: out[tap][b][a] += sum_{pix} G[pix, a] * X[pix + off(tap), b]
// grid = (tap groups, mtiles * ntiles, psplit)
//
// Both operands are MN-major (the reduction runs over pixels).  An MN-major B operand is a row of
// 32-column atoms one LBO apart - so the shifted X boxes of SEVERAL taps, laid out back to back, are
// one wide B operand: a single N <= 256 MMA multiplies the dY tile with up to 256/Cin taps at once and
// drops tap t into TMEM columns [t*ntile, (t+1)*ntile).  dY is fetched once per pixel tile for all of
// them (vs once per tap), and the MMA count falls by the same factor.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreadsTc, 1)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_x,
                const WgradParams p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the __shared__ array (an integer round trip would turn every later
  // access into a generic LD.E / ST.E instead of LDS / STS)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int R = p.bw * p.bh * p.bn;
  const int box_bytes = R * 128;
  const int b_boxes = p.ntile / 32;                       // atoms per tap
  const int stage_bytes = (p.a_alloc + p.tpc * b_boxes) * box_bytes;
  PipeBars* bars = (PipeBars*)(smem + (size_t)p.stages * stage_bytes);

  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  const int taps = p.ksize * p.ksize;
  const int tap0 = blockIdx.x * p.tpc;
  const int ntap = min(p.tpc, taps - tap0);
  const int mt = blockIdx.y % p.mtiles, nt = blockIdx.y / p.mtiles;
  const int a0 = mt * kTileM, b0 = nt * p.ntile;
  const int half = p.ksize / 2;
  const int ptiles = p.tiles_w * p.tiles_h * p.tiles_n;
  const int per = (ptiles + p.psplit - 1) / p.psplit;
  const int t_begin = blockIdx.z * per;
  const int t_end = min(ptiles, t_begin + per);
  const int niter = max(0, t_end - t_begin);
  const int ctot = ntap * p.ntile;                        // accumulator columns in use
  const uint32_t ncols = ctot <= 32 ? 32u : ctot <= 64 ? 64u : ctot <= 128 ? 128u : ctot <= 256 ? 256u : 512u;
  const bool wtrace = LAB_TRACE(p) != nullptr && blockIdx.x + blockIdx.y + blockIdx.z == 0;
  if (wtrace && threadIdx.x == 0) LAB_TRACE(p)[520] = clock64();
  const int a_live = min(kTileM / 32, (p.ca - a0 + 31) / 32);     // boxes that hold real channels
  const int b_live = min(b_boxes, (p.cb - b0 + 31) / 32);

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    mbar_init(&bars->tmem_full, 1);
    fence_barrier_init();
    tma_prefetch_desc(&map_g);
    tma_prefetch_desc(&map_x);
  }
  if (warp == 1) tmem_alloc(&bars->tmem_addr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_addr;
  pdl_wait();                              // everything above overlapped the previous kernel; its data is needed from here

  // ================= TMA producers: warps 0, 2, 3, 4, 5 (the epilogue warps are idle until the accumulator is done).
  // One UTMALDG keeps its issuing lane busy for ~290 cycles per 56-row box (tools/wgrad_trace.py): a single producer
  // spent ~4300 cycles issuing the 15 boxes of a conv2 stage while the MMA warp needed ~1200, so the boxes of a stage
  // are dealt round-robin to five issuing warps.  Warp 0 announces the stage's total bytes; a complete_tx that lands
  // first just drives the count negative until then - the phase cannot complete before warp 0's arrive.
  const int pidx = warp == 0 ? 0 : (warp >= 2 ? warp - 1 : -1);
  if (pidx >= 0) {
    constexpr int kProducers = 5;
    const int nboxes = a_live + ntap * b_live;
    const uint32_t tx = (uint32_t)(nboxes * box_bytes);
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < niter; ++i) {
      mbar_wait(&bars->empty[s], ph ^ 1u);
      if (elect_one()) {
        if (wtrace && pidx == 0 && i < 64) LAB_TRACE(p)[i * 8 + 0] = clock64();
        const int t = t_begin + i;
        const int tw = t % p.tiles_w, th = (t / p.tiles_w) % p.tiles_h, tn = t / (p.tiles_w * p.tiles_h);
        const int w0 = tw * p.bw, h0 = th * p.bh, n0 = tn * p.bn;
        uint8_t* sa = smem + (size_t)s * stage_bytes;
        uint8_t* sb = sa + p.a_alloc * box_bytes;
        if (pidx == 0) mbar_expect_tx(&bars->full[s], tx);
        for (int bx = pidx; bx < nboxes; bx += kProducers) {
          if (bx < a_live) {
            tma_load_4d(sa + bx * box_bytes, &map_g, &bars->full[s], a0 + 32 * bx, w0, h0, n0);
          } else {
            const int tt = (bx - a_live) / b_live, j = (bx - a_live) % b_live;
            const int tap = tap0 + tt;
            const int dh = tap / p.ksize - half, dw = tap % p.ksize - half;
            tma_load_4d(sb + (tt * b_boxes + j) * box_bytes, &map_x, &bars->full[s], b0 + 32 * j, w0 + dw, h0 + dh, n0);
          }
        }
        if (wtrace && pidx == 0 && i < 64) LAB_TRACE(p)[i * 8 + 1] = clock64();
      }
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
  }
  if (warp == 0) {
  } else if (warp == 1) {
    const int ksteps = R / kUmmaK;
    const uint32_t smem_base = smem_u32(smem);
    const int n_first = min(ctot, 256), n_rest = ctot - n_first;       // one or two MMAs per K step
    const uint32_t idesc0 = umma_idesc_tf32(kTileM, n_first, 1, 1);
    const uint32_t idesc1 = umma_idesc_tf32(kTileM, n_rest > 0 ? n_rest : 32, 1, 1);
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < niter; ++i) {
      if (wtrace && i < 64 && lane == 0) LAB_TRACE(p)[i * 8 + 2] = clock64();
      mbar_wait(&bars->full[s], ph);
      tc_fence_after();
      if (elect_one()) {
        if (wtrace && i < 64) LAB_TRACE(p)[i * 8 + 3] = clock64();
        const uint32_t sa = smem_base + (uint32_t)(s * stage_bytes);
        const uint32_t sb = sa + (uint32_t)(p.a_alloc * box_bytes);
        for (int k = 0; k < ksteps; ++k) {
          const uint32_t acc = (i > 0 || k > 0) ? 1u : 0u;
          const uint64_t ad = umma_smem_desc(sa + k * 1024, (uint32_t)box_bytes, 512, DDL_UMMA_SW128_BASE32B);
          const uint64_t bd0 = umma_smem_desc(sb + k * 1024, (uint32_t)box_bytes, 512, DDL_UMMA_SW128_BASE32B);
          umma_tf32(tmem_base, ad, bd0, idesc0, acc);
          if (n_rest > 0) {
            const uint64_t bd1 = umma_smem_desc(sb + 8 * box_bytes + k * 1024, (uint32_t)box_bytes, 512,
                                                DDL_UMMA_SW128_BASE32B);
            umma_tf32(tmem_base + 256, ad, bd1, idesc1, acc);
          }
        }
        umma_commit(&bars->empty[s]);
        if (i == niter - 1) umma_commit(&bars->tmem_full);
        if (wtrace && i < 64) LAB_TRACE(p)[i * 8 + 4] = clock64();
      }
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
    if (niter == 0 && elect_one()) umma_commit(&bars->tmem_full);
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int a = a0 + q * 32 + lane;
    if (lane == 0) mbar_wait(&bars->tmem_full, 0);
    __syncwarp();
    tc_fence_after();
    if (wtrace && threadIdx.x == 64) LAB_TRACE(p)[521] = clock64();
    // out[tap][b][a]: `a` (this lane's TMEM row) is the contiguous index.  A 4x4 register transpose inside
    // each quad of lanes turns 16 scalar reds per lane into 4 x red.global.add.v4.f32 on 4 consecutive a.
    const int ql = lane & 3;
    const int aq = a0 + q * 32 + (lane & ~3);               // first a of this quad
    const bool vec_ok = (p.ca % 4) == 0;
    for (int tt = 0; tt < ntap; ++tt) {
      float* otap = p.out + (int64_t)(tap0 + tt) * p.cb * p.ca;
      for (int c0 = 0; c0 < p.ntile; c0 += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(tt * p.ntile + c0), v);
        if (niter == 0) continue;
        if (!vec_ok) {
          if (a < p.ca) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int bcol = b0 + c0 + j;
              if (bcol < p.cb) atomicAdd(otap + (int64_t)bcol * p.ca + a, v[j]);
            }
          }
          continue;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {                       // registers 4g..4g+3 <-> columns b = c0+4g..+3
          float r[4] = {v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]};
          quad_transpose4(r, ql);                           // r[k] = value of quad lane k for column 4g + ql
          const int bcol = b0 + c0 + 4 * g + ql;
          if (bcol < p.cb && aq < p.ca)
            red_add_v4_f32(otap + (int64_t)bcol * p.ca + aq, make_float4(r[0], r[1], r[2], r[3]));
        }
      }
    }
  }
  tc_fence_before();
  if (wtrace && threadIdx.x == 64) LAB_TRACE(p)[522] = clock64();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ncols);
  if (wtrace && threadIdx.x == 0) LAB_TRACE(p)[523] = clock64();
}

// ------------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)f;
  });
  return fn;
}

// fp32 tensor with contiguous dims (innermost first) d0..d3; box b0..b3 (b0 = 32 floats = one 128-byte
// swizzle span).  mn_major selects the 32-byte-atom swizzle that MN-major tf32 operands require.
int make_map4(CUtensorMap* m, const void* ptr, const uint64_t d[4], const uint32_t b[4], bool mn_major) {
  EncodeTiledFn enc = get_encode();
  DDL_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
  DDL_REQUIRE(((uintptr_t)ptr & 15) == 0, "TMA base address must be 16-byte aligned");
  DDL_REQUIRE((d[0] * 4) % 16 == 0, "TMA row pitch (%llu floats) must be a multiple of 16 bytes",
              (unsigned long long)d[0]);
  cuuint64_t dims[4] = {d[0], d[1], d[2], d[3]};
  cuuint64_t strides[3] = {d[0] * 4, d[0] * d[1] * 4, d[0] * d[1] * d[2] * 4};
  cuuint32_t box[4] = {b[0], b[1], b[2], b[3]};
  cuuint32_t es[4] = {1, 1, 1, 1};
  for (int i = 0; i < 4; ++i) DDL_REQUIRE(box[i] >= 1 && box[i] <= 256, "TMA box dim %d = %u out of range", i, box[i]);
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DDL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Choose the pixel box (bw, bh, bn): rows = bw*bh*bn <= max_rows, rows % row_mult == 0, minimising
// the number of tiles (ties: fewer rows).  bh / bn may overshoot the tensor by up to row_mult-1:
// TMA zero-fills out-of-bounds rows, which contribute nothing to a reduction and are skipped by
// the epilogue of the act kernel.
// even_windows: 2x2 pooling windows never straddle two tiles (tile origins even, extents even or the whole axis).
void choose_box(int N, int H, int W, int max_rows, int row_mult, int* bw, int* bh, int* bn, bool even_windows = false) {
  int best_tiles = 1 << 30, best_rows = 0;
  *bw = 1; *bh = 1; *bn = row_mult;
  const int hmax = (H + row_mult - 1) / row_mult * row_mult;
  const int nmax = (N + row_mult - 1) / row_mult * row_mult;
  for (int w = 1; w <= W && w <= 256; ++w) {
    if (w != W && W % w != 0) continue;                 // full rows or exact divisors of the width
    if (even_windows && w != W && (w & 1)) continue;
    for (int h = 1; h <= hmax && h <= 256 && w * h <= max_rows; ++h) {
      if (even_windows && h < H && (h & 1)) continue;
      for (int n = 1; n <= nmax && n <= 256 && w * h * n <= max_rows; ++n) {
        int rows = w * h * n;
        if (rows % row_mult) continue;
        int tiles = ceil_div(W, w) * ceil_div(H, h) * ceil_div(N, n);
        if (tiles < best_tiles || (tiles == best_tiles && rows < best_rows) ||
            (even_windows && tiles == best_tiles && rows == best_rows && w > *bw)) {   // ties: longest contiguous run
          best_tiles = tiles; best_rows = rows; *bw = w; *bh = h; *bn = n;
        }
      }
    }
  }
}

}  // namespace

struct DdlTcPlan {
  int kind;                 // 0 act, 1 wgrad
  CUtensorMap map_a, map_b;
  CUtensorMap map_c;        // act plans with K4 pull: store map of the local weight copy (else = map_b)
  float* scratch = nullptr; // cluster split-K partial tiles (owned)
  ActParams act;
  WgradParams wg;
  dim3 grid;
  size_t smem;
};

// staged-epilogue options of an act plan (EPI_POOL for fprop, EPI_UNPOOL for dgrad)
struct PoolOpt {
  float* out2;            // pooled output / un-pooled gradient
  uint8_t* idx;           // argmax slots (written / read)
  float* dbias;           // EPI_UNPOOL: bias gradient (nullable)
  int UH, UW;             // EPI_UNPOOL: un-pooled plane
};

static DdlTcPlan* plan_act_impl(int dgrad, const float* act, const float* w, float* out, int N, int H, int W,
                                int cred, int nout, int ksize, int ntile, int ksplit, int flags, const float* bias,
                                const float* mask_act, float mask_scale, const DdlStepState* st,
                                uint32_t drop_stream_off, float keep_prob, float* colsum, float* w_store_dst,
                                const PoolOpt* po) {
  auto fail = [](const char* why) -> DdlTcPlan* { ddl::set_error("ddl_tc_plan_act: %s", why); return nullptr; };
  if (po) {
    if (flags || colsum || w_store_dst) return fail("staged pool epilogues take no other epilogue flags");
    if (ksplit != 1 && ksplit != 2 && ksplit != 4 && ksplit != 8) return fail("staged epilogue: cluster split must be 1, 2, 4 or 8");
    if (ntile % ksplit || (ntile / ksplit) % 4) return fail("staged epilogue: ntile / split must be a multiple of 4");
    if (!dgrad && bias == nullptr) return fail("pool epilogue needs the bias");
    if (dgrad && po->idx == nullptr) return fail("un-pool epilogue needs the argmax slots");
    flags = dgrad ? EPI_UNPOOL : EPI_POOL;
  }
  if (cred % 4 || nout % 4) return fail("channel counts must be multiples of 4 (16-byte TMA pitch)");
  if (ntile < 16 || ntile > 256 || ntile % 16) return fail("ntile must be a multiple of 16 in [16, 256]");
  if (!dgrad && ntile % 32) return fail("fprop ntile must be a multiple of 32 (MN-major boxes)");
  if (ksize != 1 && ksize != 5 && ksize != 3) return fail("ksize must be 1, 3 or 5");
  DdlTcPlan* pl = new (std::nothrow) DdlTcPlan();
  if (!pl) return fail("out of host memory");
  pl->kind = 0;
  ActParams& p = pl->act;
  p.N = N; p.H = H; p.W = W;
  choose_box(N, H, W, kTileM, 1, &p.bw, &p.bh, &p.bn, po && !dgrad);
  p.tiles_w = ceil_div(W, p.bw); p.tiles_h = ceil_div(H, p.bh); p.tiles_n = ceil_div(N, p.bn);
  p.ksize = ksize; p.flip = dgrad ? 1 : 0;
  p.cblocks = ceil_div(cred, kKB);
  p.kiters = ksize * ksize * p.cblocks;
  p.ksplit = ksplit < 1 ? 1 : (ksplit > p.kiters ? p.kiters : ksplit);
  p.nout = nout; p.ntile = ntile;
  p.b_mn_major = dgrad ? 0 : 1;
  // Split-K reduction, two ways (measured, profiles/launches_step_r1_d.csv):
  //  * plain outputs (conv fprop / dgrad: bias+ReLU live in the pool kernel) - red.add into a zeroed
  //    `out`; every CTA flushes its own tile in parallel.  17.8 us for conv4 fprop at ksplit 11;
  //  * outputs with a fused epilogue (FC layers: bias / ReLU / dropout / mask / column sums) - the ksplit
  //    (<= 8) CTAs of a tile form a thread-block cluster; ranks 1.. park partial tiles in plan-owned,
  //    L2-resident scratch, one cluster barrier, rank 0 adds them and runs the epilogue.  Keeps the
  //    fusion and needs no zeroing, but rank 0 alone reads (ksplit-1) tiles: 44.5 us for the same conv4
  //    case, so it is only used where the tile is narrow (N tile 32) and the epilogue must stay fused.
  const bool fused_epi = (flags & (EPI_BIAS | EPI_RELU | EPI_DROPOUT | EPI_MASKPOS)) != 0 || colsum != nullptr;
  const bool want_cluster = po != nullptr || (DDL_LAB_ENV("DDL_TC_NO_CLUSTER") == nullptr &&
                                             (fused_epi || DDL_LAB_ENV("DDL_TC_FORCE_CLUSTER") != nullptr));
  if (po && p.ksplit != (ksplit < 1 ? 1 : ksplit)) return (delete pl, fail("staged epilogue: split exceeds the k iterations"));
  if (want_cluster && p.ksplit > 8) p.ksplit = 8;
  p.cluster = (want_cluster && p.ksplit > 1) ? p.ksplit : 1;
  p.scratch = nullptr;
  p.flags = flags | ((p.ksplit > 1 && p.cluster == 1) ? EPI_ATOMIC : 0);
  // fused-epilogue cluster plans (the FC layers): reduce-scatter through distributed shared memory, every CTA finishes
  // ntile / cluster columns.  DDL_TC_LEADER_EPI=1 keeps the older leader-only reduction through global scratch; so do
  // K4 pull plans and cluster sizes that do not divide the tile into multiples of 4 columns.
  const bool staged_gen = !po && p.cluster > 1 && fused_epi && !(w_store_dst != nullptr && !dgrad && w_store_dst != w) &&
                          (p.cluster == 2 || p.cluster == 4 || p.cluster == 8) && ntile % p.cluster == 0 &&
                          (ntile / p.cluster) % 4 == 0 && DDL_LAB_ENV("DDL_TC_LEADER_EPI") == nullptr;
  if (staged_gen) p.flags |= EPI_STAGED;
  p.pool_out = po ? po->out2 : nullptr;
  p.pool_idx = po ? po->idx : nullptr;
  p.dbias = po ? po->dbias : nullptr;
  p.UH = po ? po->UH : 0; p.UW = po ? po->UW : 0;
  p.pull = (w_store_dst != nullptr && !dgrad && w_store_dst != w) ? 1 : 0;
  p.debug = 0;
  p.trace = nullptr;
#ifdef DDL_TC_LAB
  p.debug = ddl_lab_int("DDL_TC_DEBUG", 0);
  if (DDL_LAB_ENV("DDL_TC_TRACE")) {
    if (cudaMalloc((void**)&p.trace, 512 * 4 * sizeof(unsigned long long)) != cudaSuccess) { delete pl; return fail("cudaMalloc(trace) failed"); }
    cudaMemset(p.trace, 0, 512 * 4 * sizeof(unsigned long long));
  }
#endif
  p.out = out; p.bias = bias; p.mask_act = mask_act; p.mask_scale = mask_scale;
  p.colsum = colsum;
  if (colsum) p.flags |= EPI_COLSUM;
  if ((p.flags & EPI_COLSUM) && (p.flags & EPI_ATOMIC)) return (delete pl, fail("column sums need an unsplit reduction"));
  p.st = st; p.drop_stream_off = drop_stream_off;
  p.drop_thresh24 = keep_prob >= 1.f ? (1u << 24) : (uint32_t)lrintf(keep_prob * (float)(1 << 24));
  p.drop_inv_keep = keep_prob >= 1.f ? 1.f : 1.f / keep_prob;
  // K blocks per stage.  Two halve the barrier round trips, but under the ~110 KB budget that keeps two CTAs per SM
  // they leave only two stages, and the coarser refill granularity costs more than the saved waits: 356k vs 381k
  // samples/s for the whole step.  So one is the default and DDL_TC_KPS=2 the experiment (not with the K4
  // write-through warp, which walks single blocks).
  const int per_cta_blocks = ceil_div(p.kiters, p.ksplit);
  p.kps = 1;
  if (ddl_lab_int("DDL_TC_KPS", 1) >= 2 && !p.pull && per_cta_blocks >= 4) p.kps = 2;
  const int stage_bytes = p.kps * (kTileM * 128 + ntile * 128);
  // <= ~110 KB of stages: two CTAs per SM stay resident (grids of 150-300 CTAs run as one wave) and
  // 4 stages already cover the ~1000-cycle TMA latency at ~600 cycles per k-iteration
  int stages = (110 * 1024) / stage_bytes;
  if (stages < 2) stages = 2;
  if (stages > kMaxStages) stages = kMaxStages;
  const int fills = ceil_div(per_cta_blocks, p.kps);           // stage fills per CTA
  if (stages > fills) stages = fills < 2 ? 2 : fills;
  if (ddl_lab_int("DDL_TC_STAGES", 0) > 0) stages = ddl_lab_int("DDL_TC_STAGES", 0);
  // staged epilogues reuse the stage memory: D (cluster > 1: ntile x 128 floats) + S (128 x (cw + 1) floats)
  size_t epi_bytes = 0;
  if (po || (p.flags & EPI_STAGED)) {
    const int cw = ntile / p.cluster;
    epi_bytes = (p.cluster > 1 ? (size_t)ntile * 512 : 0) + (size_t)128 * (cw + 1) * sizeof(float);
    while ((size_t)stages * stage_bytes + stage_bytes <= epi_bytes && stages < kMaxStages && stages < fills) ++stages;
  }
  p.stages = stages;
  size_t bars_off = (size_t)stages * stage_bytes;
  if (epi_bytes > bars_off) bars_off = (epi_bytes + 15) & ~(size_t)15;
  p.bars_off = (int)bars_off;
  pl->smem = bars_off + sizeof(PipeBars) + 1024;
  if (pl->smem > 227 * 1024) return (delete pl, fail("tile does not fit shared memory"));
  pl->grid = dim3(p.tiles_w * p.tiles_h * p.tiles_n, ceil_div(nout, ntile), p.ksplit);
  if (p.cluster > 1 && !po && !(p.flags & EPI_STAGED)) {
    const size_t bytes = (size_t)pl->grid.x * pl->grid.y * (p.cluster - 1) * kTileM * ntile * sizeof(float);
    if (cudaMalloc((void**)&pl->scratch, bytes) != cudaSuccess) { delete pl; return fail("cudaMalloc(split-K scratch) failed"); }
    p.scratch = pl->scratch;
  }
  const uint64_t da[4] = {(uint64_t)cred, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint32_t ba[4] = {kKB, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
  if (make_map4(&pl->map_a, act, da, ba, false)) { ddl_tc_plan_destroy(pl); return nullptr; }
  const int taps = ksize * ksize;
  if (!dgrad) {           // W as [taps][cred][nout]: inner = nout
    const uint64_t dw[4] = {(uint64_t)nout, (uint64_t)cred, (uint64_t)taps, 1};
    const uint32_t bwb[4] = {32, kKB, 1, 1};
    if (make_map4(&pl->map_b, w, dw, bwb, true)) { delete pl; return nullptr; }
    pl->map_c = pl->map_b;
    if (p.pull && make_map4(&pl->map_c, w_store_dst, dw, bwb, true)) { delete pl; return nullptr; }
  } else {                // W as [taps][nout][cred]: inner = cred (the reduction)
    const uint64_t dw[4] = {(uint64_t)cred, (uint64_t)nout, (uint64_t)taps, 1};
    const uint32_t bwb[4] = {kKB, (uint32_t)ntile, 1, 1};
    if (make_map4(&pl->map_b, w, dw, bwb, false)) { delete pl; return nullptr; }
    pl->map_c = pl->map_b;
  }
  if (cudaFuncSetAttribute(tc_act_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) {
    delete pl;
    return fail("cudaFuncSetAttribute(max dynamic smem) failed");
  }
  return pl;
}

extern "C" {

DdlTcPlan* ddl_tc_plan_act(int dgrad, const float* act, const float* w, float* out, int N, int H, int W,
                           int cred, int nout, int ksize, int ntile, int ksplit, int flags, const float* bias,
                           const float* mask_act, float mask_scale, const DdlStepState* st,
                           uint32_t drop_stream_off, float keep_prob, float* colsum, float* w_store_dst) {
  return plan_act_impl(dgrad, act, w, out, N, H, W, cred, nout, ksize, ntile, ksplit, flags, bias, mask_act, mask_scale,
                       st, drop_stream_off, keep_prob, colsum, w_store_dst, nullptr);
}

/* conv fprop whose epilogue is bias + ReLU + 2x2 SAME max-pool: writes pooled [N, ceil(H/2), ceil(W/2), nout] and
 * (idx != NULL) the argmax slot of every window (0..3, 4 = ReLU-dead).  cluster in {1, 2, 4, 8}: the K loop is split
 * over a thread-block cluster whose CTAs reduce-scatter their partial tiles through distributed shared memory. */
DdlTcPlan* ddl_tc_plan_conv_pool(const float* act, const float* w, const float* bias, float* pooled, uint8_t* idx,
                                 int N, int H, int W, int cin, int cout, int ksize, int ntile, int cluster) {
  PoolOpt po = {pooled, idx, nullptr, 0, 0};
  return plan_act_impl(0, act, w, nullptr, N, H, W, cin, cout, ksize, ntile, cluster, 0, bias, nullptr, 1.f, nullptr, 0,
                       1.f, nullptr, nullptr, &po);
}

/* conv dgrad whose epilogue is the pool + ReLU backward of the layer below: dY [N, H, W, cout] -> gradient of the
 * un-pooled plane [N, UH, UW, cin] (value at the window's argmax slot, zeros elsewhere) and dbias += its column sums. */
DdlTcPlan* ddl_tc_plan_dgrad_unpool(const float* dy, const float* w, const uint8_t* idx, float* dfull, float* dbias,
                                    int N, int H, int W, int UH, int UW, int cout, int cin, int ksize, int ntile,
                                    int cluster) {
  PoolOpt po = {dfull, const_cast<uint8_t*>(idx), dbias, UH, UW};
  if ((UH + 1) / 2 != H || (UW + 1) / 2 != W) { ddl::set_error("ddl_tc_plan_dgrad_unpool: plane %dx%d does not pool to %dx%d", UH, UW, H, W); return nullptr; }
  return plan_act_impl(1, dy, w, nullptr, N, H, W, cout, cin, ksize, ntile, cluster, 0, nullptr, nullptr, 1.f, nullptr, 0,
                       1.f, nullptr, nullptr, &po);
}

/* conv dgrad whose epilogue is only the ReLU part of the pool backward: gp[N, H, W, cin] = dgrad value where the window
 * below is alive (idx < 4), 0 where it is ReLU-dead; dbias += column sums.  For the layer above conv1, whose un-pooled
 * gradient nobody needs: ddl_conv1_wgrad_pooled consumes gp + idx directly (4x fewer bytes than the un-pooled plane). */
DdlTcPlan* ddl_tc_plan_dgrad_poolgrad(const float* dy, const float* w, const uint8_t* idx, float* gp, float* dbias,
                                      int N, int H, int W, int cout, int cin, int ksize, int ntile, int cluster) {
  PoolOpt po = {gp, const_cast<uint8_t*>(idx), dbias, H, W};
  DdlTcPlan* pl = plan_act_impl(1, dy, w, nullptr, N, H, W, cout, cin, ksize, ntile, cluster, 0, nullptr, nullptr, 1.f, nullptr,
                                0, 1.f, nullptr, nullptr, &po);
  if (pl) pl->act.flags |= EPI_POOLGRAD;
  return pl;
}

DdlTcPlan* ddl_tc_plan_wgrad(const float* g, const float* x, float* out, int N, int H, int W, int ca, int cb,
                             int ksize, int ntile, int psplit) {
  auto fail = [](const char* why) -> DdlTcPlan* { ddl::set_error("ddl_tc_plan_wgrad: %s", why); return nullptr; };
  if (ca % 4 || cb % 4) return fail("channel counts must be multiples of 4 (16-byte TMA pitch)");
  if (ntile < 32 || ntile > 256 || ntile % 32) return fail("ntile must be a multiple of 32 in [32, 256]");
  DdlTcPlan* pl = new (std::nothrow) DdlTcPlan();
  if (!pl) return fail("out of host memory");
  pl->kind = 1;
  WgradParams& p = pl->wg;
  p.N = N; p.H = H; p.W = W;
  const int taps = ksize * ksize;
  const int b_boxes = ntile / 32;
  p.ksize = ksize; p.ca = ca; p.cb = cb;
  p.mtiles = ceil_div(ca, kTileM); p.ntile = ntile; p.ntiles = ceil_div(cb, ntile);
  // taps per CTA: up to 512 accumulator columns (two N <= 256 MMAs per K step), balanced over the groups;
  // the second MMA's B descriptor starts 8 atoms in, so a group wider than 256 columns needs ntile | 256
  int tpc_max = 512 / ntile;
  if (tpc_max > 1 && (256 % ntile) != 0) tpc_max = 256 / ntile > 0 ? 256 / ntile : 1;
  if (tpc_max < 1) tpc_max = 1;
  // Measured (tools/wgrad_perf.py, profiles/wgrad_policy.md): about five taps per CTA and one wave of CTAs is the
  // fastest split for all three conv layers (20-25 us alone vs 31-45 us with the widest tap groups): fewer shifted X
  // boxes per stage (deeper pipeline, larger pixel boxes) and a narrower red.add epilogue per CTA.
  // ... but inside the training step the wider grids take SMs from the critical dgrad chain running next to them
  // (336k vs 367k samples/s), so it is a lab-build switch: DDL_WGRAD_POLICY_WAVE=1
  static const bool old_policy = ddl_lab_int("DDL_WGRAD_POLICY_WAVE", 0) == 0;
  if (!old_policy && tpc_max > 5) tpc_max = 5;
  if (ddl_lab_int("DDL_WGRAD_TPC", 0) >= 1 && ddl_lab_int("DDL_WGRAD_TPC", 0) < tpc_max) tpc_max = ddl_lab_int("DDL_WGRAD_TPC", 0);
  if (tpc_max > taps) tpc_max = taps;
  p.tgroups = ceil_div(taps, tpc_max);
  p.tpc = ceil_div(taps, p.tgroups);
  const int a_live_max = ca >= kTileM ? kTileM / 32 : ceil_div(ca, 32);
  p.a_alloc = a_live_max;
  if (p.a_alloc + p.tpc * b_boxes < kTileM / 32) p.a_alloc = kTileM / 32;   // the M=128 descriptor spans 4 atoms
  const int boxes = p.a_alloc + p.tpc * b_boxes;
  // rows per stage: multiple of 8 (one UMMA K step = 8 pixel rows); two stages must fit in shared memory.
  // DDL_WGRAD_SMEM_KB (perf experiments): shared-memory budget of a wgrad CTA (default 216: one CTA per SM)
  static const int smem_kb = ddl_lab_int("DDL_WGRAD_SMEM_KB", 216);
  int max_rows = ((smem_kb * 1024 / 2) / (boxes * 128)) & ~7;
  if (max_rows > 128) max_rows = 128;
  if (max_rows < 8) { delete pl; return fail("tap group does not fit in shared memory"); }
  choose_box(N, H, W, max_rows, 8, &p.bw, &p.bh, &p.bn);
  const int R = p.bw * p.bh * p.bn;
  if (R % 8) { delete pl; return fail("no pixel box with a multiple of 8 rows"); }
  if (p.bw > 256 || p.bh > 256 || p.bn > 256 || R > 256) { delete pl; return fail("pixel box does not fit TMA limits"); }
  p.tiles_w = ceil_div(W, p.bw); p.tiles_h = ceil_div(H, p.bh); p.tiles_n = ceil_div(N, p.bn);
  const int ptiles = p.tiles_w * p.tiles_h * p.tiles_n;
  if (psplit <= 0) {
    // Every pixel split re-adds the whole dW tile with red.add, so splits are not free: aim for >= ~12
    // pixel tiles per CTA and never more than one wave of CTAs.
    const int groups = p.tgroups * p.mtiles * p.ntiles;
    if (old_policy) {
      const int wave = ceil_div(148, groups);
      psplit = ceil_div(ptiles, 12);
      if (psplit > wave) psplit = wave;
    } else {
      // one wave of CTAs, but at least ~4 pixel tiles each (every split re-adds the whole dW tile with red.add)
      psplit = 148 / groups;
      if (psplit > ceil_div(ptiles, 4)) psplit = ceil_div(ptiles, 4);
    }
  }
  p.psplit = psplit < 1 ? 1 : (psplit > ptiles ? ptiles : psplit);
  const int stage_bytes = boxes * R * 128;
  int stages = (smem_kb * 1024) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages > ceil_div(ptiles, p.psplit)) stages = ceil_div(ptiles, p.psplit) < 2 ? 2 : ceil_div(ptiles, p.psplit);
  if (stages < 1) { delete pl; return fail("stage does not fit in shared memory"); }
  p.stages = stages;
  p.out = out;
  p.trace = nullptr;
#ifdef DDL_TC_LAB
  if (DDL_LAB_ENV("DDL_TC_TRACE")) {
    if (cudaMalloc((void**)&p.trace, 512 * 4 * sizeof(unsigned long long)) != cudaSuccess) { delete pl; return fail("cudaMalloc(trace) failed"); }
    cudaMemset(p.trace, 0, 512 * 4 * sizeof(unsigned long long));
  }
#endif
  pl->smem = (size_t)stages * stage_bytes + sizeof(PipeBars) + 1024;
  pl->grid = dim3(p.tgroups, p.mtiles * p.ntiles, p.psplit);
  const uint64_t dg[4] = {(uint64_t)ca, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint64_t dx[4] = {(uint64_t)cb, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint32_t bx[4] = {32, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
  if (make_map4(&pl->map_a, g, dg, bx, true) || make_map4(&pl->map_b, x, dx, bx, true)) { delete pl; return nullptr; }
  if (cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) {
    delete pl;
    return fail("cudaFuncSetAttribute(max dynamic smem) failed");
  }
  return pl;
}

int ddl_tc_plan_run(const DdlTcPlan* pl, cudaStream_t s) {
  DDL_REQUIRE(pl != nullptr, "ddl_tc_plan_run: null plan");
  if (pl->kind == 0 && pl->act.cluster > 1) {
    DDL_CHECK_CUDA(ddl::launch(tc_act_gemm_kernel, pl->grid, dim3(kThreadsTc), pl->smem, s, (unsigned)pl->act.cluster,
                               pl->map_a, pl->map_b, pl->map_c, pl->act));
  } else if (pl->kind == 0)
    DDL_LAUNCH(tc_act_gemm_kernel, pl->grid, kThreadsTc, pl->smem, s, pl->map_a, pl->map_b, pl->map_c, pl->act);
  else
    DDL_LAUNCH(tc_wgrad_kernel, pl->grid, kThreadsTc, pl->smem, s, pl->map_a, pl->map_b, pl->wg);
  DDL_LAUNCH_CHECK();
  return 0;
}

void ddl_tc_plan_destroy(DdlTcPlan* pl) {
  if (pl && pl->scratch) cudaFree(pl->scratch);
  delete pl;
}

/* 1 when the plan accumulates with red.add and `out` must therefore be zeroed before every run */
int ddl_tc_plan_needs_zero(const DdlTcPlan* pl) { return pl && pl->kind == 0 && (pl->act.flags & EPI_ATOMIC) ? 1 : 0; }

/* perf experiments: copies the clock64 trace of an act plan created under DDL_TC_TRACE=1 */
int ddl_tc_plan_trace(const DdlTcPlan* pl, unsigned long long* host_out, int n) {
  const unsigned long long* src = pl ? (pl->kind == 0 ? pl->act.trace : pl->wg.trace) : nullptr;
  DDL_REQUIRE(src != nullptr, "no trace buffer");
  DDL_CHECK_CUDA(cudaMemcpy(host_out, src, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Allocation-free convolution block (no counterpart in the reference).
 *
 * Who this is for: a host program that must not call cudaMalloc.  Two kinds exist in practice.  Some
 * have handed the whole device to another allocator - a framework's caching pool told to own all free
 * memory, or an MPS / MIG slice with a hard quota - so one more cudaMalloc simply fails.  Others run a
 * latency-critical loop (serving, control) and cannot accept cudaMalloc's device-wide synchronisation
 * and unbounded latency in the middle of it.  Every other conv entry point of this library needs the
 * caller to provide (hence allocate) a buffer for the raw convolution sums; this one keeps that buffer
 * inside the library, in the file-scope array below, which every GPU that loads the library holds until
 * the process exits.  The caller passes only inputs and outputs.
 *
 * y[N, ceil(H/2), ceil(W/2), cout] = maxpool2x2_SAME(relu(conv5x5_SAME(x, w) + bias)),  idx optional.
 *
 * Rules:
 *  - the raw sums (N*H*W*cout floats) must fit DDL_STATIC_SCRATCH_BYTES, else the call is refused;
 *  - the first call on each device performs a one-time set-up (symbol address lookup, kernel attribute);
 *    make that first call before starting any stream capture;
 *  - all calls share the one array, which keeps the previous call's contents; this function clears the
 *    part it accumulates into, and two calls that overlap on the GPU (any streams) race on it - serialise
 *    them;
 *  - it calls no cudaMalloc / cudaFree and no allocating library routine.
 */
static __device__ __align__(256) float g_static_scratch[DDL_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;
float* g_static_ptr[kMaxDevices] = {};       // device address of g_static_scratch, per device

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: handing it to a
// kernel would fault (or, where the GPU can reach host memory, quietly use the wrong bytes).  The real
// address comes from cudaGetSymbolAddress, once per device, with that device current.
int static_scratch_setup(float** out) {
  int dev = -1;
  DDL_CHECK_CUDA(cudaGetDevice(&dev));
  DDL_REQUIRE(dev >= 0 && dev < kMaxDevices, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    DDL_CHECK_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    DDL_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    DDL_CHECK_CUDA(cudaFuncSetAttribute(tc_act_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    g_static_ptr[dev] = (float*)p;
  }
  *out = g_static_ptr[dev];
  return 0;
}
}  // namespace

int ddl_conv5x5_relu_pool_static(const float* x, const float* w, const float* bias, float* y, uint8_t* idx,
                                 int N, int H, int W, int cin, int cout, cudaStream_t s) {
  DDL_REQUIRE(N > 0 && H > 0 && W > 0 && cin > 0 && cout > 0, "ddl_conv5x5_relu_pool_static: empty shape");
  DDL_REQUIRE(cin % 4 == 0 && cout % 32 == 0, "ddl_conv5x5_relu_pool_static: cin %% 4 == 0 and cout %% 32 == 0 required");
  const size_t need = (size_t)N * H * W * cout * sizeof(float);
  DDL_REQUIRE(need <= (size_t)DDL_STATIC_SCRATCH_BYTES,
              "ddl_conv5x5_relu_pool_static: N*H*W*cout*4 = %zu bytes exceeds the built-in scratch of %d bytes "
              "(DDL_STATIC_SCRATCH_BYTES); use ddl_tc_plan_act with your own buffer instead",
              need, (int)DDL_STATIC_SCRATCH_BYTES);
  float* raw = nullptr;
  if (static_scratch_setup(&raw)) return -1;

  // a plan on the stack: same geometry rules as ddl_tc_plan_act(dgrad = 0), nothing heap-allocated
  DdlTcPlan pl;
  pl.kind = 0;
  ActParams& p = pl.act;
  p = ActParams();
  p.N = N; p.H = H; p.W = W;
  choose_box(N, H, W, kTileM, 1, &p.bw, &p.bh, &p.bn);
  p.tiles_w = ceil_div(W, p.bw); p.tiles_h = ceil_div(H, p.bh); p.tiles_n = ceil_div(N, p.bn);
  p.ksize = 5; p.flip = 0;
  p.cblocks = ceil_div(cin, kKB);
  p.kiters = 25 * p.cblocks;
  int ntile = 32;
  for (int c : {256, 128, 64, 32}) if (cout % c == 0) { ntile = c; break; }
  const int ctas = p.tiles_w * p.tiles_h * p.tiles_n * (cout / ntile);
  int ksplit = 1;
  if (ctas < 100) {
    ksplit = (140 + ctas / 2) / ctas;
    if (ksplit > p.kiters / 4) ksplit = p.kiters / 4 > 0 ? p.kiters / 4 : 1;
    if (ksplit > 16) ksplit = 16;
  }
  p.ksplit = ksplit;
  p.nout = cout; p.ntile = ntile;
  p.b_mn_major = 1;
  p.flags = ksplit > 1 ? EPI_ATOMIC : 0;
  p.pull = 0; p.debug = 0; p.trace = nullptr; p.cluster = 1; p.scratch = nullptr;
  p.out = raw; p.bias = nullptr; p.colsum = nullptr; p.mask_act = nullptr; p.mask_scale = 1.f;
  p.st = nullptr; p.drop_stream_off = 0; p.drop_thresh24 = 1u << 24; p.drop_inv_keep = 1.f;
  const int stage_bytes = kTileM * 128 + ntile * 128;
  int stages = (110 * 1024) / stage_bytes;
  if (stages < 2) stages = 2;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages > p.kiters) stages = p.kiters < 2 ? 2 : p.kiters;
  p.stages = stages;
  p.kps = 1;
  p.bars_off = stages * stage_bytes;           // the barriers sit right behind the pipeline stages
  pl.smem = (size_t)stages * stage_bytes + sizeof(PipeBars) + 1024;
  pl.grid = dim3(p.tiles_w * p.tiles_h * p.tiles_n, cout / ntile, ksplit);
  const uint64_t da[4] = {(uint64_t)cin, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint32_t ba[4] = {kKB, (uint32_t)p.bw, (uint32_t)p.bh, (uint32_t)p.bn};
  if (make_map4(&pl.map_a, x, da, ba, false)) return -1;
  const uint64_t dw[4] = {(uint64_t)cout, (uint64_t)cin, 25, 1};
  const uint32_t bwb[4] = {32, kKB, 1, 1};
  if (make_map4(&pl.map_b, w, dw, bwb, true)) return -1;

  // split-K partial sums are added with red.add, so the region must start at zero; the array still
  // holds whatever the previous call left there
  if (ksplit > 1) DDL_CHECK_CUDA(cudaMemsetAsync(raw, 0, need, s));
  // the kernel prefetches weight tiles before griddepcontrol.wait; launched without the programmatic-launch attribute it
  // starts only after everything earlier on the stream (e.g. an optimizer kernel of this library writing `w`) is done
  ddl::no_pdl_once();
  DDL_LAUNCH(tc_act_gemm_kernel, pl.grid, kThreadsTc, pl.smem, s, pl.map_a, pl.map_b, pl.map_b, pl.act);
  DDL_LAUNCH_CHECK();
  return ddl_maxpool_fwd(raw, bias, y, idx, N, H, W, cout, s);
}

/* host-only: the TMA pixel box (bw, bh, bn) chosen for an [N, H, W] activation; even_windows = the pooling variant */
int ddl_plan_choose_box(int N, int H, int W, int max_rows, int row_mult, int even_windows, int* box3) {
  DDL_REQUIRE(N > 0 && H > 0 && W > 0 && max_rows > 0 && row_mult > 0 && box3 != nullptr, "ddl_plan_choose_box: bad arguments");
  choose_box(N, H, W, max_rows, row_mult, &box3[0], &box3[1], &box3[2], even_windows != 0);
  return 0;
}

/* introspection for tests / profiling notes: fills {grid.x, grid.y, grid.z, stages, smem, bw, bh, bn} */
int ddl_tc_plan_info(const DdlTcPlan* pl, int* out8) {
  DDL_REQUIRE(pl != nullptr, "ddl_tc_plan_info: null plan");
  out8[0] = pl->grid.x; out8[1] = pl->grid.y; out8[2] = pl->grid.z;
  out8[3] = pl->kind == 0 ? pl->act.stages : pl->wg.stages;
  out8[4] = (int)pl->smem;
  out8[5] = pl->kind == 0 ? pl->act.bw : pl->wg.bw;
  out8[6] = pl->kind == 0 ? pl->act.bh : pl->wg.bh;
  out8[7] = pl->kind == 0 ? pl->act.bn : pl->wg.bn;
  return 0;
}

}  // extern "C"
