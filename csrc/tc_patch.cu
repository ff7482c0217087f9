// Patch-resident 5x5 convolution on tcgen05: one CTA per image, the zero-padded activation patch is loaded ONCE and
// every filter tap reads it through a shifted shared-memory descriptor.
//
// Why (profiles/timeline_step_r2.md, profiles/ncu_full_r1b_act_summary.csv): the tap-by-tap implicit GEMM of tc_gemm.cu
// re-fetches the 128-row activation box for each of the 25 taps, 16 KB per K block.  For the 14x14 layers that is
// 70 MB (conv2 forward) and 134 MB (conv2's dgrad, N = 32: 20 KB of operands per 131k MACs) of L2 -> shared-memory
// traffic for 2.5 MB / 5 MB of data; with every SM pulling, L2 feeds ~42 B/clk/SM and the tensor pipe sat at 7-14%.
// Here the activation crosses L2 once per image (42 KB per 32 channels) and only the weights stream.
//
// Reference ops: tf.nn.conv2d + bias + relu + max_pool2d of conv2 (mnist_sync/model/model.py:35-42) and its
// backprop-input (implied by compute_gradients, model.py:101).
//
// Geometry.  The patch of image n is a 4-D TMA box [32 ch, W+4, H+4, 1] starting at (w, h) = (-2, -2): out-of-bounds
// rows / columns are zero-filled, which IS SAME padding, and the box lands in shared memory as (H+4) * (W+4) rows of
// 128 bytes, row r = (h+2) * Wp + (w+2), Wp = W + 4, SWIZZLE_128B.  Output pixel (h, w) is M row r of the GEMM and tap
// (dh, dw) multiplies patch row r + dh*Wp + dw: the A operand of a tap is the SAME memory at a start address shifted by
// (dh*Wp + dw) * 128 bytes.  The swizzle XOR is a function of address bits [7,10), which the shift carries along, so a
// shifted descriptor reads exactly what TMA wrote.  M rows that fall on the pad columns (w < 0 or w >= W) are garbage
// and are dropped by the epilogue; an M tile of 128 rows holds rpt = 128 / Wp image rows, an image takes
// ceil(H / rpt) tiles = accumulators side by side in TMEM.
#include "common.cuh"
#include "../include/ddl_b200.h"

#include <mutex>
#include <new>
#include <stdlib.h>

namespace {

constexpr int kThreadsPatch = 192;     // warp 0: patch TMA, warp 1: MMA, warp 3: weight TMA + epilogue, warps 2, 4, 5: epilogue
constexpr int kMaxPatchStages = 8;
constexpr int kSlackBytes = 1024;      // 8 rows in front of the patch: the first tile's (-2, -2) tap starts 2 rows early

struct PatchParams {
  int N, H, W, Wp, Hp;
  int cblocks;                     // ceil(reduction channels / 32)
  int flip;                        // dgrad: tap offsets negated
  int nout;                        // output channels = UMMA N (one N tile)
  int b_mn_major;                  // 1: weights [tap][cred][nout] (fprop), 0: [tap][nout][cred] (dgrad)
  int rpt, tiles;                  // image rows per M tile, M tiles (accumulators) per image
  int tps;                         // filter taps per ring stage: 5 (a filter row, one barrier round trip per 5 taps) or 1
  int stages;                      // weight ring stages (tps taps x all channel blocks each)
  int patch_bytes;                 // shared memory per channel block of the patch (slack + rows + slack, multiple of 1024)
  int ring_off, bars_off;
  int mode;                        // 0: bias + ReLU + 2x2 max-pool -> out [N, PH, PW, nout] (+ idx);  1: ReLU mask of the
                                   // layer below (idx < 4) -> out [N, H, W, nout], dbias += column sums
  float* out;
  uint8_t* idx;
  const float* bias;
  float* dbias;
  unsigned long long* trace;       // lab build only (DDL_TC_LAB): globaltimer phase stamps of CTA 0
  int debug;                       // lab build only: 1 = load the weight ring once, never refill; 2 = epilogue warps sleep instead of polling
};
#ifdef DDL_TC_LAB
#define PDEBUG(p) ((p).debug)
#else
#define PDEBUG(p) 0
#endif

#ifdef DDL_TC_LAB
#define PSTAMP(i) do { if (p.trace && blockIdx.x == 0) p.trace[i] = ddl_globaltimer(); } while (0)
#else
#define PSTAMP(i) do { } while (0)
#endif

struct alignas(8) PatchBars {
  uint64_t full[kMaxPatchStages];
  uint64_t empty[kMaxPatchStages];
  uint64_t patch_full;
  uint64_t tmem_full;
  uint32_t tmem_addr;
  uint32_t pad;
};

__device__ __forceinline__ int spos(int c, int cw) { return (c & 3) * (cw >> 2) + (c >> 2); }
__device__ __forceinline__ uint32_t tmem_cols_pow2(int n) {
  return n <= 32 ? 32u : n <= 64 ? 64u : n <= 128 ? 128u : n <= 256 ? 256u : 512u;
}

// One ring stage = one filter row (5 taps) against every channel block and M tile, fully unrolled: every offset is an
// immediate, the only run-time inputs are the stage's base descriptor words.  Issue-side work per MMA is what the tensor
// pipe waits for (it accepts the next MMA only when the previous one is nearly done), so nothing is computed in between.
//   a_row0 : descriptor low word of patch row (2 + dh) * Wp - 2*sgn ... see the caller; step per tap = sgn * 8 (one row)
template <int CB, int T>
__device__ __forceinline__ void patch_issue_row(uint32_t a_lo_row, int a_tap_step, uint32_t a_cb, uint32_t a_tile,
                                                uint32_t b_lo_stage, uint32_t b_tap, uint32_t b_cb, uint32_t b_kstep,
                                                uint64_t a_tmpl, uint64_t b_tmpl, uint32_t tmem_base, uint32_t nout,
                                                uint32_t idesc, bool first_row, uint32_t issue) {
#pragma unroll
  for (int tt = 0; tt < 5; ++tt) {
#pragma unroll
    for (int cb = 0; cb < CB; ++cb) {
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const uint32_t a_lo = a_lo_row + (uint32_t)(tt * a_tap_step) + (uint32_t)cb * a_cb + (uint32_t)t * a_tile;
        const uint32_t b_lo = b_lo_stage + (uint32_t)tt * b_tap + (uint32_t)cb * b_cb;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t ad = a_tmpl | (uint64_t)((a_lo + 2u * k) & 0x3FFFu);
          const uint64_t bd = b_tmpl | (uint64_t)((b_lo + b_kstep * k) & 0x3FFFu);
          const uint32_t acc = (tt > 0 || cb > 0 || k > 0) ? 1u : (first_row ? 0u : 1u);
          umma_tf32_pred(tmem_base + (uint32_t)t * nout, ad, bd, idesc, acc, issue);
        }
      }
    }
  }
}

// grid = (N images); one CTA per SM is enough at batch 100 (100 CTAs, one wave)
__global__ void __launch_bounds__(kThreadsPatch, 1)
tc_patch_conv_kernel(const __grid_constant__ CUtensorMap map_act, const __grid_constant__ CUtensorMap map_w,
                     const PatchParams p) {
  pdl_launch_dependents();
  if (threadIdx.x == 0) PSTAMP(0);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  PatchBars* bars = (PatchBars*)(smem + (size_t)p.bars_off);
  uint8_t* ring = smem + (size_t)p.ring_off;
  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  const int n = blockIdx.x;
  const int b_bytes = p.nout * 128;                       // one (tap, channel block) weight tile
  const int stage_bytes = p.tps * p.cblocks * b_bytes;
  const int nstage = 25 / p.tps;                          // ring fills per image
  const uint32_t ncols = tmem_cols_pow2(p.tiles * p.nout);

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    mbar_init(&bars->patch_full, 1);
    mbar_init(&bars->tmem_full, 1);
    fence_barrier_init();
    tma_prefetch_desc(&map_act);
    tma_prefetch_desc(&map_w);
  }
  if (warp == 1) tmem_alloc(&bars->tmem_addr, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_addr;
  if (threadIdx.x == 0) PSTAMP(1);                      // prologue done

  // the weights are not produced by the preceding kernels of the stream (see ddl::no_pdl_once in the engine): warp 3
  // starts filling the ring before the wait; everybody else needs the previous kernel's activations
  if (warp != 3) pdl_wait();

  if (warp == 3) {
    // ================= weight producer: tps taps (all channel blocks) per stage, ONE box per (channel block, 32-column
    // atom) covering the stage's taps: shared-memory layout [cb][atom j][tap][32 rows][128 B] (fprop, MN-major B) or
    // [cb][tap][nout rows][128 B] (dgrad, K-major B)
    int s = 0;
    uint32_t ph = 0;
    for (int st = 0; st < nstage; ++st) {
      if ((PDEBUG(p) & 1) && st >= p.stages) break;
      mbar_wait(&bars->empty[s], ph ^ 1u);
      if (elect_one()) {
        mbar_expect_tx(&bars->full[s], (uint32_t)stage_bytes);
        uint8_t* sb = ring + (size_t)s * stage_bytes;
        for (int cb = 0; cb < p.cblocks; ++cb) {
          uint8_t* scb = sb + (size_t)cb * p.tps * b_bytes;
          if (p.b_mn_major) {
            for (int j = 0; j < p.nout / 32; ++j)
              tma_load_4d(scb + (size_t)j * p.tps * 4096, &map_w, &bars->full[s], 32 * j, cb * 32, st * p.tps, 0);
          } else {
            tma_load_4d(scb, &map_w, &bars->full[s], cb * 32, 0, st * p.tps, 0);
          }
        }
      }
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
    pdl_wait();
  } else if (warp == 0) {
    // ================= patch producer: the whole padded image, one box per 32 channels, once
    if (elect_one()) {
      mbar_expect_tx(&bars->patch_full, (uint32_t)(p.cblocks * p.Hp * p.Wp * 128));
      for (int cb = 0; cb < p.cblocks; ++cb)
        tma_load_4d(smem + (size_t)cb * p.patch_bytes + kSlackBytes, &map_act, &bars->patch_full, cb * 32, -2, -2, n);
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer: stage by stage (tps taps each), every tap against every M tile of the image.  The
    // loops are warp-uniform (all lanes compute the descriptors, from uniform values only), the tcgen05 instructions are
    // predicated on one lane.  The tensor pipe accepts the next MMA only when the previous one is nearly done
    // (tools/mma_probe.cu: issue time = completion time, 49 cycles per 128x64x8), so every cycle the issuing warp spends
    // on barriers and address arithmetic is an idle tensor core: one barrier round trip per filter ROW, descriptor low
    // words advanced by additions.
    const uint32_t idesc = umma_idesc_tf32(128, p.nout, 0, p.b_mn_major);
    const uint32_t patch0 = smem_u32(smem) + kSlackBytes;
    const uint32_t ring0 = smem_u32(ring);
    const uint32_t issue = elect_one() ? 1u : 0u;
    // descriptor high words are constant; the low word is (address >> 4) in its 14-bit field
    const uint64_t a_tmpl = umma_smem_desc(0, 16, 1024, DDL_UMMA_SW128);
    const uint64_t b_tmpl = p.b_mn_major ? umma_smem_desc(0, (uint32_t)(p.tps * 4096), 512, DDL_UMMA_SW128_BASE32B)
                                         : umma_smem_desc(0, 16, 1024, DDL_UMMA_SW128);
    const uint32_t b_kstep = p.b_mn_major ? (1024u >> 4) : (32u >> 4);
    const uint32_t b_tap = p.b_mn_major ? (4096u >> 4) : ((uint32_t)b_bytes >> 4);       // next tap inside a box
    const uint32_t b_cb = (uint32_t)(p.tps * b_bytes) >> 4;                               // next channel block
    const uint32_t a_cb = (uint32_t)p.patch_bytes >> 4;
    const uint32_t a_tile = (uint32_t)(p.rpt * p.Wp * 128) >> 4;
    const int sgn = p.flip ? -1 : 1;
    mbar_wait(&bars->patch_full, 0);
    if (lane == 0) PSTAMP(2);                           // patch landed
    int s = 0;
    uint32_t ph = 0;
    for (int st = 0; st < nstage; ++st) {
#ifdef DDL_TC_LAB
      const unsigned long long w0 = ddl_globaltimer();
#endif
      if (!((PDEBUG(p) & 1) && st >= p.stages)) mbar_wait(&bars->full[s], ph);
#ifdef DDL_TC_LAB
      if (p.trace && blockIdx.x == 0 && lane == 0) p.trace[9] = (st == 0 ? 0ull : p.trace[9]) + (ddl_globaltimer() - w0);   // time waiting for weights
#endif
      if (lane == 0 && st == 0) PSTAMP(3);              // first weight stage landed
#ifdef DDL_TC_LAB
      if (p.trace && blockIdx.x == 0 && lane == 0 && st == 0) p.trace[10] = (unsigned long long)clock64();
#endif
      if (lane == 0 && st == nstage / 2) PSTAMP(8);
      tc_fence_after();
      const uint32_t sb_lo = (ring0 + (uint32_t)(s * stage_bytes)) >> 4;
      bool fast = false;
      if (p.tps == 5) {
        // stage st = filter row dh = sgn * (st - 2); its taps dw = sgn * (tt - 2): one patch row apart (8 low-word units)
        const uint32_t a_row = (patch0 + (uint32_t)(((2 + sgn * (st - 2)) * p.Wp - 2 * sgn) * 128)) >> 4;
        const int a_step = 8 * sgn;
        fast = true;
        if (p.cblocks == 1 && p.tiles == 2)
          patch_issue_row<1, 2>(a_row, a_step, a_cb, a_tile, sb_lo, b_tap, b_cb, b_kstep, a_tmpl, b_tmpl, tmem_base, (uint32_t)p.nout, idesc, st == 0, issue);
        else if (p.cblocks == 2 && p.tiles == 2)
          patch_issue_row<2, 2>(a_row, a_step, a_cb, a_tile, sb_lo, b_tap, b_cb, b_kstep, a_tmpl, b_tmpl, tmem_base, (uint32_t)p.nout, idesc, st == 0, issue);
        else if (p.cblocks == 2 && p.tiles == 1)
          patch_issue_row<2, 1>(a_row, a_step, a_cb, a_tile, sb_lo, b_tap, b_cb, b_kstep, a_tmpl, b_tmpl, tmem_base, (uint32_t)p.nout, idesc, st == 0, issue);
        else
          fast = false;
      }
      for (int tt = 0; tt < (fast ? 0 : p.tps); ++tt) {
        const int tap = st * p.tps + tt;
        const int dh = sgn * (tap / 5 - 2), dw = sgn * (tap % 5 - 2);
        // M row i of tile t is patch row (2 + t*rpt) * Wp + i (pixel (t*rpt, -2) first); the tap shifts it
        const uint32_t a_lo0 = (patch0 + (uint32_t)(((2 + dh) * p.Wp + dw) * 128)) >> 4;
        const uint32_t b_lo0 = sb_lo + (uint32_t)tt * b_tap;
        for (int cb = 0; cb < p.cblocks; ++cb) {
          for (int t = 0; t < p.tiles; ++t) {
            const uint32_t a_lo = a_lo0 + (uint32_t)cb * a_cb + (uint32_t)t * a_tile;
            const uint32_t b_lo = b_lo0 + (uint32_t)cb * b_cb;
            const uint32_t td = tmem_base + (uint32_t)(t * p.nout);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              // base-offset field left 0: measured, the swizzle XOR follows the address bits (a non-zero field scrambles)
              const uint64_t ad = a_tmpl | (uint64_t)((a_lo + 2u * k) & 0x3FFFu);
              const uint64_t bd = b_tmpl | (uint64_t)((b_lo + b_kstep * k) & 0x3FFFu);
              umma_tf32_pred(td, ad, bd, idesc, (tap > 0 || cb > 0 || k > 0) ? 1u : 0u, issue);
            }
          }
        }
      }
      umma_commit_pred(&bars->empty[s], issue);
      if (st == nstage - 1) { umma_commit_pred(&bars->tmem_full, issue); if (lane == 0) PSTAMP(4); }   // last MMA issued
#ifdef DDL_TC_LAB
      if (p.trace && blockIdx.x == 0 && lane == 0 && st == nstage - 1) p.trace[11] = (unsigned long long)clock64();
#endif
      __syncwarp();
      if (++s == p.stages) { s = 0; ph ^= 1u; }
    }
  }

  // ================= epilogue (warps 2..5): accumulators -> shared memory S[padded pixel][nout + 1] -> fused output.
  // The patch and the weight ring are dead once tmem_full fires; S reuses them.  S row = h * Wp + (w + 2).
  const int sp = p.nout + 1;
  float* S = reinterpret_cast<float*>(smem);
  if (warp >= 2) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int live = p.rpt * p.Wp;                       // M rows of a tile that are image rows (the rest is overhang)
    // dgrad: this thread's argmax slots (written by the forward pass) are fetched while the MMAs run - the epilogue then
    // has no dependent global load left (it cost 9 us of the kernel when the loads were issued one trip at a time)
    constexpr int kPre = 16;
    uchar4 pre[kPre];
    const int te0 = (int)threadIdx.x - 64;
    const int cq4e = p.nout >> 2;
    const int items_d = p.H * p.W * cq4e;
    if (p.mode == 1) {
#pragma unroll
      for (int u = 0; u < kPre; ++u) {
        const int it = te0 + u * 128;
        pre[u] = make_uchar4(4, 4, 4, 4);
        if (it < items_d) {
          const int c4 = it % cq4e, pix = it / cq4e;
          pre[u] = *reinterpret_cast<const uchar4*>(p.idx + ((int64_t)n * p.H * p.W + pix) * p.nout + 4 * c4);
        }
      }
    }
    if (PDEBUG(p) & 2) __nanosleep(8000);
    if (lane == 0) mbar_wait(&bars->tmem_full, 0);
    __syncwarp();
    tc_fence_after();
    if (threadIdx.x == 64) PSTAMP(5);                   // accumulators complete
    for (int t = 0; t < p.tiles; ++t) {
      const int srow = t * live + row;
      const bool keep = row < live && srow < p.H * p.Wp;
      for (int c0 = 0; c0 < p.nout; c0 += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.nout + c0), v);
        if (keep) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float b = p.mode == 0 ? __ldg(p.bias + c0 + j) : 0.f;
            S[srow * sp + spos(c0 + j, p.nout)] = v[j] + b;
          }
        }
      }
    }
    named_bar_sync(1, 128);
    if (threadIdx.x == 64) PSTAMP(6);                   // staged
    const int te = (int)threadIdx.x - 64;
    const int cq4 = p.nout >> 2;
    if (p.mode == 0) {
      // bias (added above) + ReLU (running max starts at 0) + 2x2 stride-2 SAME max-pool, argmax slot 0..3 (4 = dead)
      const int PH = (p.H + 1) / 2, PW = (p.W + 1) / 2;
      const int items = PH * PW * cq4;
      for (int it = te; it < items; it += 128) {
        const int c4 = it % cq4, pp = it / cq4;
        const int pw = pp % PW, phh = pp / PW;
        float best[4] = {0.f, 0.f, 0.f, 0.f};
        int bi[4] = {4, 4, 4, 4};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int lh = 2 * phh + (k >> 1), lw = 2 * pw + (k & 1);
          if (lh < p.H && lw < p.W) {
            const float* sr = S + (lh * p.Wp + lw + 2) * sp + c4;
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const float x = sr[q4 * cq4];
              if (x > best[q4]) { best[q4] = x; bi[q4] = k; }
            }
          }
        }
        const int64_t o = (((int64_t)n * PH + phh) * PW + pw) * p.nout + 4 * c4;
        *reinterpret_cast<float4*>(p.out + o) = make_float4(best[0], best[1], best[2], best[3]);
        if (p.idx) *reinterpret_cast<uchar4*>(p.idx + o) = make_uchar4(bi[0], bi[1], bi[2], bi[3]);
      }
    } else {
      // ReLU mask of the layer below from its argmax slots (idx < 4: the window is alive), gradient kept pooled
      const int items = p.H * p.W * cq4;
      const bool fixed_col = (128 % cq4) == 0;
      float dbsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
      for (int u = 0; te + u * 128 < items; ++u) {
        const int it = te + u * 128;
        const int c4 = it % cq4, pix = it / cq4;
        const int w = pix % p.W, h = pix / p.W;
        const int64_t o = (((int64_t)n * p.H + h) * p.W + w) * p.nout + 4 * c4;
        uchar4 bi;
        if (u < kPre) {
          bi = pre[0];
#pragma unroll
          for (int j = 1; j < kPre; ++j) if (u == j) bi = pre[j];          // register array: selected, not indexed
        } else {
          bi = *reinterpret_cast<const uchar4*>(p.idx + o);
        }
        const float* sr = S + (h * p.Wp + w + 2) * sp + c4;
        const int b4[4] = {bi.x, bi.y, bi.z, bi.w};
        float g[4];
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) g[q4] = b4[q4] < 4 ? sr[q4 * cq4] : 0.f;
        *reinterpret_cast<float4*>(p.out + o) = make_float4(g[0], g[1], g[2], g[3]);
        if (p.dbias) {
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            if (fixed_col) dbsum[q4] += g[q4];
            else if (g[q4] != 0.f) atomicAdd(p.dbias + 4 * c4 + q4, g[q4]);
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
        const int col = 4 * (te % cq4);
        if (!pow2 || cq4 >= 32 || lane < cq4) {
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4)
            if (dbsum[q4] != 0.f) atomicAdd(p.dbias + col + q4, dbsum[q4]);
        }
      }
    }
  }
  if (threadIdx.x == 64) PSTAMP(7);                     // outputs written
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ncols);
}

// ------------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn patch_get_encode() {
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

// fp32 tensor, contiguous dims d0..d3 (innermost first), box b0..b3 with b0 = 32 floats (one 128-byte swizzle span)
int patch_make_map(CUtensorMap* m, const void* ptr, const uint64_t d[4], const uint32_t b[4], bool mn_major) {
  EncodeTiledFn enc = patch_get_encode();
  DDL_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
  DDL_REQUIRE(((uintptr_t)ptr & 15) == 0 && (d[0] * 4) % 16 == 0, "TMA base / pitch must be 16-byte aligned");
  cuuint64_t dims[4] = {d[0], d[1], d[2], d[3]};
  cuuint64_t strides[3] = {d[0] * 4, d[0] * d[1] * 4, d[0] * d[1] * d[2] * 4};
  cuuint32_t box[4] = {b[0], b[1], b[2], b[3]};
  cuuint32_t es[4] = {1, 1, 1, 1};
  for (int i = 0; i < 4; ++i) DDL_REQUIRE(box[i] >= 1 && box[i] <= 256, "TMA box dim %d = %u out of range", i, box[i]);
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DDL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

}  // namespace

struct DdlPatchPlan {
  CUtensorMap map_act, map_w;
  PatchParams p;
  dim3 grid;
  size_t smem;
};

extern "C" {

namespace {
// shared-memory plan: taps per stage (5 if two stages of a filter row fit, else 1), ring depth, offsets.  0 = does not fit.
int patch_layout(int H, int W, int cred, int nout, int* tps, int* stages, size_t* patch_bytes, size_t* body) {
  if (H < 1 || W < 1 || W + 4 > 128 || H + 4 > 256 || cred % 32 || nout % 32 || nout > 256) return 0;
  const int Wp = W + 4, rpt = 128 / Wp, tiles = (H + rpt - 1) / rpt;
  if (rpt < 1 || tiles * nout > 512) return 0;
  const int cblocks = cred / 32;
  const size_t patch = ((size_t)kSlackBytes + (size_t)(H + 4) * Wp * 128 + kSlackBytes + 1023) / 1024 * 1024;
  const size_t stg = (size_t)H * Wp * (nout + 1) * sizeof(float);
  const size_t budget = 224 * 1024 - sizeof(PatchBars) - 2048;
  for (int t : {5, 1}) {
    const size_t stage = (size_t)t * cblocks * nout * 128;
    if (cblocks * patch + 2 * stage > budget || stg > budget) continue;
    int st = (int)((budget - cblocks * patch) / stage);
    if (st > kMaxPatchStages) st = kMaxPatchStages;
    if (st > 25 / t) st = 25 / t;
    size_t bd = cblocks * patch + (size_t)st * stage;
    if (stg > bd) bd = (stg + 1023) / 1024 * 1024;
    *tps = t; *stages = st; *patch_bytes = patch; *body = bd;
    return 1;
  }
  return 0;
}
}  // namespace

/* 1 when the patch-resident kernel can run this layer: 5x5 SAME, channels in multiples of 32, one N tile (nout <= 256),
 * all M tiles of an image side by side in TMEM (tiles * nout <= 512), patch + two weight stages in shared memory */
int ddl_patch_conv_fits(int H, int W, int cred, int nout) {
  int tps, stages;
  size_t pb, body;
  return patch_layout(H, W, cred, nout, &tps, &stages, &pb, &body);
}

/* 1 when the patch-resident kernel is also the FASTER choice: its M rows include the pad columns, so it pays off where at
 * least 70% of the rows of its M tiles are real pixels (14x14: 77%; 7x7: 38% - there the tap-by-tap plans win) */
int ddl_patch_conv_preferred(int H, int W, int cred, int nout) {
  if (!ddl_patch_conv_fits(H, W, cred, nout)) return 0;
  const int Wp = W + 4, rpt = 128 / Wp, tiles = (H + rpt - 1) / rpt;
  return H * W * 10 >= tiles * 128 * 7 ? 1 : 0;
}

/* dgrad = 0: out [N, ceil(H/2), ceil(W/2), nout] = maxpool2x2_SAME(relu(conv5x5_SAME(act, w) + bias)), idx (nullable) the
 *            argmax slots; w [5,5,cred,nout].
 * dgrad = 1: out [N, H, W, nout] = conv-dgrad(act = dY [N,H,W,cred], w [5,5,nout,cred]) masked by idx < 4 (the ReLU of
 *            the layer below, pooled layout), dbias (nullable) += column sums.
 * Same results as ddl_tc_plan_conv_pool / ddl_tc_plan_dgrad_poolgrad (tf32 products, fp32 accumulation). */
DdlPatchPlan* ddl_patch_plan_create(int dgrad, const float* act, const float* w, const float* bias, float* out, uint8_t* idx,
                                    float* dbias, int N, int H, int W, int cred, int nout) {
  auto fail = [](const char* why) -> DdlPatchPlan* { ddl::set_error("ddl_patch_plan_create: %s", why); return nullptr; };
  if (!ddl_patch_conv_fits(H, W, cred, nout)) return fail("shape not supported by the patch-resident kernel");
  if (!dgrad && bias == nullptr) return fail("the forward variant needs the bias");
  if (dgrad && idx == nullptr) return fail("the dgrad variant needs the argmax slots");
  DdlPatchPlan* pl = new (std::nothrow) DdlPatchPlan();
  if (!pl) return fail("out of host memory");
  PatchParams& p = pl->p;
  p.N = N; p.H = H; p.W = W; p.Wp = W + 4; p.Hp = H + 4;
  p.cblocks = cred / 32; p.flip = dgrad ? 1 : 0; p.nout = nout; p.b_mn_major = dgrad ? 0 : 1;
  p.rpt = 128 / p.Wp; p.tiles = (H + p.rpt - 1) / p.rpt;
  int tps = 1, stages = 2;
  size_t pbytes = 0, body = 0;
  patch_layout(H, W, cred, nout, &tps, &stages, &pbytes, &body);
  p.patch_bytes = (int)pbytes;
  p.tps = tps; p.stages = stages;
  p.ring_off = (int)((size_t)p.cblocks * pbytes);
  p.bars_off = (int)body;
  pl->smem = body + sizeof(PatchBars) + 1024;
  p.mode = dgrad ? 1 : 0;
  p.out = out; p.idx = idx; p.bias = bias; p.dbias = dbias;
  p.trace = nullptr;
  p.debug = 0;
#ifdef DDL_TC_LAB
  p.debug = ddl_lab_int("DDL_PATCH_DEBUG", 0);
#endif
  pl->grid = dim3(N, 1, 1);
  const uint64_t da[4] = {(uint64_t)cred, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint32_t ba[4] = {32, (uint32_t)p.Wp, (uint32_t)p.Hp, 1};
  if (patch_make_map(&pl->map_act, act, da, ba, false)) { delete pl; return nullptr; }
  if (!dgrad) {
    const uint64_t dw[4] = {(uint64_t)nout, (uint64_t)cred, 25, 1};
    const uint32_t bw[4] = {32, 32, (uint32_t)p.tps, 1};
    if (patch_make_map(&pl->map_w, w, dw, bw, true)) { delete pl; return nullptr; }
  } else {
    const uint64_t dw[4] = {(uint64_t)cred, (uint64_t)nout, 25, 1};
    const uint32_t bw[4] = {32, (uint32_t)nout, (uint32_t)p.tps, 1};
    if (patch_make_map(&pl->map_w, w, dw, bw, false)) { delete pl; return nullptr; }
  }
  if (cudaFuncSetAttribute(tc_patch_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) {
    delete pl;
    return fail("cudaFuncSetAttribute(max dynamic smem) failed");
  }
  return pl;
}

int ddl_patch_plan_run(const DdlPatchPlan* pl, cudaStream_t s) {
  DDL_REQUIRE(pl != nullptr, "ddl_patch_plan_run: null plan");
  DDL_LAUNCH(tc_patch_conv_kernel, pl->grid, kThreadsPatch, pl->smem, s, pl->map_act, pl->map_w, pl->p);
  DDL_LAUNCH_CHECK();
  return 0;
}

void ddl_patch_plan_destroy(DdlPatchPlan* pl) { delete pl; }

/* lab build: device buffer (>= 16 x uint64) for the phase stamps of CTA 0; NULL switches them off */
int ddl_patch_plan_set_trace(DdlPatchPlan* pl, unsigned long long* trace) {
  DDL_REQUIRE(pl != nullptr, "ddl_patch_plan_set_trace: null plan");
  pl->p.trace = trace;
  return 0;
}

}  // extern "C"
