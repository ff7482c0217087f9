// SIMT (CUDA-core) kernels: optimizers, step state, conv1 (Cin = 1), pooling, fc3 + softmax
// cross-entropy, column sums.  Everything GEMM-shaped with >= 32 reduction channels lives in
// tc_gemm.cu on the tcgen05 tensor cores; what is left here is bandwidth- or latency-bound.
//
// Reference ops replaced: ApplyAdam (mnist_sync/parameter_server.py:21-27),
// conv1 + bias + ReLU + pool (mnist_sync/model/model.py:24-31), max_pool2d and its gradient
// (model.py:31,42,53,64), fc3 + softmax_cross_entropy_with_logits + reduce_mean (model.py:85-92),
// accuracy (model.py:104-105), dropout masks (model.py:74,82).
#include "common.cuh"
#include "../include/ddl_b200.h"

namespace {

constexpr int kThreads = 256;

inline int grid_for(int64_t n, int per_block, int cap = 148 * 8) {
  int64_t g = (n + per_block - 1) / per_block;
  if (g < 1) g = 1;
  if (g > cap) g = cap;
  return (int)g;
}

// ------------------------------------------------------------------------------------ step state
__global__ void step_advance_kernel(DdlStepState* st, float lr, float b1, float b2, int advance_opt) {
  pdl_enter();
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const uint32_t ws = st->wstep + 1;
    st->wstep = ws;
    st->stream = st->stream_base + (ws - 1) * st->stream_stride;
    st->loss_sum = 0.f;
    st->correct = 0;
    if (advance_opt) {
      const uint32_t t = st->t + 1;
      st->t = t;
      const double c1 = 1.0 - pow((double)b1, (double)t);
      const double c2 = 1.0 - pow((double)b2, (double)t);
      st->lr_t = (float)((double)lr * sqrt(c2) / c1);
    }
  }
}

// ------------------------------------------------------------------------------------ optimizers
// kind: 0 sgd, 1 momentum, 2 adam (TF1 form: eps outside the bias correction).
template <int KIND>
__global__ void optimizer_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                                 const float* __restrict__ g, int64_t n4, int64_t n, float step_size,
                                 float a, float b, float eps, float gscale, const DdlStepState* st) {
  pdl_enter();
  if (st != nullptr && KIND == 2) step_size = st->lr_t;
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 pv = reinterpret_cast<float4*>(p)[i];
    float gg[4] = {gv.x * gscale, gv.y * gscale, gv.z * gscale, gv.w * gscale};
    float pp[4] = {pv.x, pv.y, pv.z, pv.w};
    if (KIND == 0) {
#pragma unroll
      for (int k = 0; k < 4; ++k) pp[k] -= step_size * gg[k];
    } else if (KIND == 1) {
      float4 mv = reinterpret_cast<float4*>(m)[i];
      float mm[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) { mm[k] = a * mm[k] + gg[k]; pp[k] -= step_size * mm[k]; }
      reinterpret_cast<float4*>(m)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    } else {
      float4 mv = reinterpret_cast<float4*>(m)[i];
      float4 vv = reinterpret_cast<float4*>(v)[i];
      float mm[4] = {mv.x, mv.y, mv.z, mv.w};
      float vs[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        mm[k] = a * mm[k] + (1.f - a) * gg[k];
        vs[k] = b * vs[k] + (1.f - b) * gg[k] * gg[k];
        pp[k] -= step_size * mm[k] / (sqrtf(vs[k]) + eps);
      }
      reinterpret_cast<float4*>(m)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
      reinterpret_cast<float4*>(v)[i] = make_float4(vs[0], vs[1], vs[2], vs[3]);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
  }
  // scalar tail (n not a multiple of 4)
  for (int64_t j = n4 * 4 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n; j += stride) {
    float gg = g[j] * gscale;
    if (KIND == 0) {
      p[j] -= step_size * gg;
    } else if (KIND == 1) {
      float mm = a * m[j] + gg; m[j] = mm; p[j] -= step_size * mm;
    } else {
      float mm = a * m[j] + (1.f - a) * gg;
      float vs = b * v[j] + (1.f - b) * gg * gg;
      m[j] = mm; v[j] = vs;
      p[j] -= step_size * mm / (sqrtf(vs) + eps);
    }
  }
}

// ------------------------------------------------------------------------------------ conv1 (Cin = 1)
// One block = one image: the zero-padded image ((S+4)^2 floats) and the filter bank live in shared
// memory, so the inner loop has no bounds checks.  A work item = one pooled pixel x 4 channels: the 6x6
// input patch is read once (broadcast LDS) and reused for 4 channels x 4 conv outputs (2x2 pool window);
// weights come as float4.  Output: bias + ReLU + 2x2 max-pool, 2-bit arg-max (4 = "max is 0").
// x [B,S,S] (Cin = 1), w [25][C], y [B,S/2,S/2,C].  S even, C % 4 == 0.
__global__ void __launch_bounds__(256) conv1_relu_pool_kernel(const float* __restrict__ x,
                                                              const float* __restrict__ w,
                                                              const float* __restrict__ bias,
                                                              float* __restrict__ y, uint8_t* __restrict__ idx,
                                                              int B, int S, int C) {
  pdl_enter();
  extern __shared__ float sm1[];
  const int XP = S + 4;
  float* sw = sm1;                         // [25][C] + [C] bias
  float* sx = sm1 + 26 * C;                // [(S+4)][(S+4)]
  const int n = blockIdx.x;
  for (int i = threadIdx.x; i < 25 * C; i += blockDim.x) sw[i] = w[i];
  for (int i = threadIdx.x; i < C; i += blockDim.x) sw[25 * C + i] = bias[i];
  for (int i = threadIdx.x; i < XP * XP; i += blockDim.x) {
    const int r = i / XP - 2, cc = i % XP - 2;
    sx[i] = (r >= 0 && r < S && cc >= 0 && cc < S) ? x[(int64_t)n * S * S + r * S + cc] : 0.f;
  }
  __syncthreads();
  const int P = S / 2, C4 = C / 4;
  const int items = P * P * C4;
  for (int t = threadIdx.x; t < items; t += blockDim.x) {
    const int c4 = t % C4, pix = t / C4;
    const int pw = pix % P, ph = pix / P;
    float patch[6][6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j) patch[i][j] = sx[(2 * ph + i) * XP + 2 * pw + j];
    float acc[4][4];                       // [window position][channel]
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[k][q] = 0.f;
#pragma unroll
    for (int kh = 0; kh < 5; ++kh)
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) {
        const float4 wv = *reinterpret_cast<const float4*>(sw + (kh * 5 + kw) * C + c4 * 4);
        const float ww[4] = {wv.x, wv.y, wv.z, wv.w};
        const float p0 = patch[kh][kw], p1 = patch[kh][kw + 1], p2 = patch[kh + 1][kw], p3 = patch[kh + 1][kw + 1];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          acc[0][q] += p0 * ww[q]; acc[1][q] += p1 * ww[q]; acc[2][q] += p2 * ww[q]; acc[3][q] += p3 * ww[q];
        }
      }
    const float4 bv4 = *reinterpret_cast<const float4*>(sw + 25 * C + c4 * 4);
    const float bb[4] = {bv4.x, bv4.y, bv4.z, bv4.w};
    float best[4] = {0.f, 0.f, 0.f, 0.f};
    int bi[4] = {4, 4, 4, 4};
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float a = acc[k][q] + bb[q];
        if (a > best[q]) { best[q] = a; bi[q] = k; }
      }
    const int64_t o = ((int64_t)n * P * P + pix) * C + c4 * 4;
    *reinterpret_cast<float4*>(y + o) = make_float4(best[0], best[1], best[2], best[3]);
    if (idx) *reinterpret_cast<uchar4*>(idx + o) = make_uchar4(bi[0], bi[1], bi[2], bi[3]);
  }
}

// conv1 weight / bias gradient.  dY is the full-resolution gradient [B,S,S,C] (after pool-bwd).
// A thread owns one channel and one filter ROW (kh): it walks its strip of pixels left to right keeping
// the five x values under the filter row in a register window, so each pixel costs one dY load, one x
// load and five FMAs for five taps.  Block = 32 channels x 5 filter rows (+ warp 5: the bias);
// grid = (B * strips, C/32) with strips of S/4 image rows.
template <int S>
__global__ void __launch_bounds__(192) conv1_wgrad_kernel(const float* __restrict__ x,
                                                          const float* __restrict__ dy,
                                                          float* __restrict__ dw, float* __restrict__ db,
                                                          int B, int C, int strip) {
  pdl_enter();
  extern __shared__ float sx[];                  // [(strip+4)][S+4] zero-padded x rows of this strip
  const int XP = S + 4;
  const int strips = S / strip;
  const int n = blockIdx.x / strips, h0 = (blockIdx.x % strips) * strip;
  const int cg = blockIdx.y * 32;
  const int lane = threadIdx.x & 31, kh = threadIdx.x >> 5;      // kh = 5: bias warp
  for (int i = threadIdx.x; i < (strip + 4) * XP; i += blockDim.x) {
    const int r = i / XP + h0 - 2, cc = i % XP - 2;
    sx[i] = (r >= 0 && r < S && cc >= 0 && cc < S) ? x[(int64_t)n * S * S + r * S + cc] : 0.f;
  }
  __syncthreads();
  const float* dyi = dy + ((int64_t)n * S * S + (int64_t)h0 * S) * C + cg + lane;
  float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  if (kh < 5) {
    for (int py = 0; py < strip; ++py) {
      const float* xr = sx + (py + kh) * XP;     // x row under filter row kh for output row py
      float w0 = xr[0], w1 = xr[1], w2 = xr[2], w3 = xr[3];
      const float* gr = dyi + (int64_t)py * S * C;
#pragma unroll
      for (int px = 0; px < S; ++px) {       // fully unrolled: the register window rotates for free
        const float w4 = xr[px + 4];
        const float g = gr[px * C];
        acc[0] += w0 * g; acc[1] += w1 * g; acc[2] += w2 * g; acc[3] += w3 * g; acc[4] += w4 * g;
        w0 = w1; w1 = w2; w2 = w3; w3 = w4;
      }
    }
#pragma unroll
    for (int kw = 0; kw < 5; ++kw) atomicAdd(dw + (kh * 5 + kw) * C + cg + lane, acc[kw]);
  } else {
    float s2 = 0.f;
    for (int i = 0; i < strip * S; ++i) s2 += dyi[(int64_t)i * C];
    atomicAdd(db + cg + lane, s2);
  }
}

// ------------------------------------------------------------------------------------ pooling
// bias + ReLU + 2x2 stride-2 SAME max-pool over NHWC raw conv sums (bias may be NULL when the input
// is already post-ReLU).  Starting the running max at 0 IS the ReLU.
// idx: 0..3 position of the max inside the window, 4 = max is 0 (ReLU dead: no gradient).
__global__ void maxpool_fwd_kernel(const float* __restrict__ a, const float* __restrict__ bias,
                                   float* __restrict__ y, uint8_t* __restrict__ idx,
                                   int B, int H, int W, int C) {
  pdl_enter();
  const int PH = (H + 1) / 2, PW = (W + 1) / 2;
  const int C4 = C / 4;
  int64_t total = (int64_t)B * PH * PW * C4;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int c4 = (int)(t % C4);
    int64_t pix = t / C4;
    int pw = (int)(pix % PW), ph = (int)((pix / PW) % PH);
    int64_t n = pix / ((int64_t)PW * PH);
    float best[4] = {0.f, 0.f, 0.f, 0.f};
    int bi[4] = {4, 4, 4, 4};
    float4 bv = bias ? *reinterpret_cast<const float4*>(bias + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int h = 2 * ph + (k >> 1), w = 2 * pw + (k & 1);
      if (h < H && w < W) {
        float4 v = *reinterpret_cast<const float4*>(a + ((n * H + h) * W + w) * C + c4 * 4);
        float vv[4] = {v.x + bv.x, v.y + bv.y, v.z + bv.z, v.w + bv.w};
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (vv[q] > best[q]) { best[q] = vv[q]; bi[q] = k; }
      }
    }
    *reinterpret_cast<float4*>(y + pix * C + c4 * 4) = make_float4(best[0], best[1], best[2], best[3]);
    if (idx) {
      uchar4 o = make_uchar4(bi[0], bi[1], bi[2], bi[3]);
      *reinterpret_cast<uchar4*>(idx + pix * C + c4 * 4) = o;
    }
  }
}

// Pool + ReLU backward: routes the pooled gradient to the argmax position of each window, zero
// elsewhere; also accumulates the bias gradient (sum of dY over pixels) per channel.
// Block layout: threadIdx.x covers C4 = C/4 channel quads (fastest), threadIdx rows cover pixels.
__global__ void maxpool_bwd_kernel(const float* __restrict__ dp, const uint8_t* __restrict__ idx,
                                   float* __restrict__ dy, float* __restrict__ db,
                                   int B, int H, int W, int C) {
  pdl_enter();
  extern __shared__ float sb[];           // [C] bias-gradient partials
  const int PH = (H + 1) / 2, PW = (W + 1) / 2;
  const int C4 = C / 4;
  for (int i = threadIdx.x; i < C; i += blockDim.x) sb[i] = 0.f;
  __syncthreads();
  int64_t total = (int64_t)B * PH * PW * C4;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    int c4 = (int)(t % C4);
    int64_t pix = t / C4;
    int pw = (int)(pix % PW), ph = (int)((pix / PW) % PH);
    int64_t n = pix / ((int64_t)PW * PH);
    float4 g = *reinterpret_cast<const float4*>(dp + pix * C + c4 * 4);
    uchar4 bi = *reinterpret_cast<const uchar4*>(idx + pix * C + c4 * 4);
    float gg[4] = {g.x, g.y, g.z, g.w};
    int bb[4] = {bi.x, bi.y, bi.z, bi.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int h = 2 * ph + (k >> 1), w = 2 * pw + (k & 1);
      if (h < H && w < W) {
        float4 o = make_float4(bb[0] == k ? gg[0] : 0.f, bb[1] == k ? gg[1] : 0.f,
                               bb[2] == k ? gg[2] : 0.f, bb[3] == k ? gg[3] : 0.f);
        *reinterpret_cast<float4*>(dy + ((n * H + h) * W + w) * C + c4 * 4) = o;
      }
    }
    if (db) {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (bb[q] < 4 && gg[q] != 0.f) atomicAdd(&sb[c4 * 4 + q], gg[q]);
    }
  }
  if (db) {
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x)
      if (sb[i] != 0.f) atomicAdd(db + i, sb[i]);
  }
}

// ------------------------------------------------------------------------------------ column sums
// out[c] += sum_r x[r, c]   (bias gradients of the FC layers)
__global__ void colsum_kernel(const float* __restrict__ x, float* __restrict__ out, int R, int C) {
  pdl_enter();
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s = 0.f;
  for (int r = blockIdx.y; r < R; r += gridDim.y) s += x[(int64_t)r * C + c];
  atomicAdd(out + c, s);
}

// ------------------------------------------------------------------------------------ fc3 + loss
// One warp per sample row.  logits = h @ W + b (K x NC, NC <= 16), softmax cross-entropy against a
// class id, dlogits = (softmax - onehot) * scale, accuracy count, loss sum.
__global__ void fc3_xent_fwd_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                    const float* __restrict__ bias, const int* __restrict__ labels,
                                    float* __restrict__ logits, float* __restrict__ dlogits,
                                    DdlStepState* st, int B, int K, int NC, float scale, int use_smem) {
  pdl_enter();
  extern __shared__ float sw3[];            // [K][NC] when use_smem
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (use_smem) {
    for (int i = threadIdx.x; i < K * NC; i += blockDim.x) sw3[i] = w[i];
    __syncthreads();
  }
  if (row >= B) return;
  const float* wsrc = use_smem ? sw3 : w;
  float acc[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) acc[c] = 0.f;
  for (int kb = 0; kb < K; kb += 32 * 8) {                 // 8 independent h loads in flight per lane
    float hv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int k = kb + u * 32 + lane;
      hv[u] = k < K ? h[(int64_t)row * K + k] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int k = kb + u * 32 + lane;
      if (k < K) {
        const float* wr = wsrc + (int64_t)k * NC;
#pragma unroll
        for (int c = 0; c < 16; ++c)
          if (c < NC) acc[c] += hv[u] * wr[c];
      }
    }
  }
  float z = -INFINITY;
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    float v = acc[c];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == c && c < NC) z = v + bias[c];
  }
  if (lane < NC && logits) logits[(int64_t)row * NC + lane] = z;
  float mx = z;
  int am = lane;
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, mx, o);
    int oa = __shfl_xor_sync(0xffffffffu, am, o);
    if (om > mx || (om == mx && oa < am)) { mx = om; am = oa; }       // first max, like argmax
  }
  float e = (lane < NC) ? __expf(z - mx) : 0.f;
  float se = e;
  for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
  int lab = labels ? labels[row] : -1;
  if (dlogits && lane < NC) dlogits[(int64_t)row * NC + lane] = (e / se - (lane == lab ? 1.f : 0.f)) * scale;
  float zl = __shfl_sync(0xffffffffu, z, lab >= 0 ? lab : 0);
  if (lane == 0 && st && lab >= 0) {
    atomicAdd(&st->loss_sum, (logf(se) + mx - zl));
    if (am == lab) atomicAdd(&st->correct, 1u);
  }
}

// fc3 backward: dW[k, c] += sum_r h[r,k] dl[r,c];  db[c] += sum_r dl[r,c];
// dh[r,k] = (sum_c dl[r,c] W[k,c]) * dropout_mask(r*K+k) * inv_keep.
// grid.x = ceil(K / 32): a block owns 32 k-columns of h, staged in smem with dl.
__global__ void fc3_bwd_kernel(const float* __restrict__ h, const float* __restrict__ w,
                               const float* __restrict__ dl, float* __restrict__ dw, float* __restrict__ db,
                               float* __restrict__ dh, float* __restrict__ dbh, int B, int K, int NC,
                               uint32_t seed, uint32_t stream, uint32_t thresh24, float inv_keep,
                               const DdlStepState* st) {
  pdl_enter();
  extern __shared__ float sm[];
  __shared__ float colred[8][32];
  if (st) { seed = st->seed; stream += st->stream; }
  float* sdl = sm;                        // [B][NC]
  float* sh = sm + B * NC;                // [B][32]
  float* sw = sh + B * 32;                // [32][NC]
  const int k0 = blockIdx.x * 32;
  for (int i = threadIdx.x; i < B * NC; i += blockDim.x) sdl[i] = dl[i];
  for (int i = threadIdx.x; i < B * 32; i += blockDim.x) {
    int r = i >> 5, k = k0 + (i & 31);
    sh[i] = (k < K) ? h[(int64_t)r * K + k] : 0.f;
  }
  for (int i = threadIdx.x; i < 32 * NC; i += blockDim.x) {
    int k = k0 + i / NC;
    sw[i] = (k < K) ? w[(int64_t)k0 * NC + i] : 0.f;
  }
  __syncthreads();
  const int c = threadIdx.x & 31, kk = threadIdx.x >> 5;
  if (c < NC) {
    for (int ks = kk; ks < 32; ks += (blockDim.x >> 5)) {
      int k = k0 + ks;
      if (k >= K) break;
      float s2 = 0.f;
      for (int r = 0; r < B; ++r) s2 += sh[r * 32 + ks] * sdl[r * NC + c];
      dw[(int64_t)k * NC + c] += s2;
    }
    if (blockIdx.x == 0 && kk == 0) {
      float s2 = 0.f;
      for (int r = 0; r < B; ++r) s2 += sdl[r * NC + c];
      db[c] += s2;
    }
  }
  float colsum = 0.f;                      // blockDim.x is a multiple of 32: a thread keeps one column
  for (int t = threadIdx.x; t < B * 32; t += blockDim.x) {
    int r = t >> 5, ks = t & 31, k = k0 + ks;
    if (k >= K) continue;
    float s2 = 0.f;
    for (int cc = 0; cc < NC; ++cc) s2 += sdl[r * NC + cc] * sw[ks * NC + cc];
    uint32_t e = (uint32_t)(r * K + k);
    bool keep = thresh24 >= (1u << 24) || ddl_dropout_keep(seed, stream, e, thresh24);
    const float o = keep ? s2 * inv_keep : 0.f;
    dh[(int64_t)r * K + k] = o;
    colsum += o;
  }
  if (dbh) {                                // bias gradient of the layer that produced h: sum_r dh[r, k]
    colred[threadIdx.x >> 5][threadIdx.x & 31] = colsum;
    __syncthreads();
    if (threadIdx.x < 32 && k0 + threadIdx.x < K) {
      float t2 = 0.f;
      for (int l = 0; l < (int)(blockDim.x >> 5); ++l) t2 += colred[l][threadIdx.x];
      dbh[k0 + threadIdx.x] += t2;
    }
  }
}

// fc3 forward + softmax cross-entropy + fc3 backward in ONE kernel (training path).
// Every block recomputes the [B, NC] logits (B*K*NC MACs - tiny) so no grid-wide hand-off is needed
// between "dlogits ready" and the weight / input gradients; block b then owns the 32 k-columns
// [32b, 32b+32) exactly like fc3_bwd_kernel.  Block 0 alone accumulates the loss / accuracy.
__global__ void __launch_bounds__(256) fc3_fused_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                                        const float* __restrict__ bias, const int* __restrict__ labels,
                                                        float* __restrict__ dw, float* __restrict__ db,
                                                        float* __restrict__ dh, float* __restrict__ dbh,
                                                        DdlStepState* st, int B, int K, int NC, float scale,
                                                        uint32_t stream_off, uint32_t thresh24, float inv_keep) {
  pdl_enter();
  extern __shared__ float sm[];
  __shared__ float colred[8][32];
  float* sw3 = sm;                        // [K][NC]
  float* sdl = sw3 + K * NC;              // [B][NC]
  float* sh = sdl + B * NC;               // [B][32]
  const int k0 = blockIdx.x * 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const uint32_t seed = st->seed, stream = st->stream + stream_off;
  for (int i = threadIdx.x; i < K * NC; i += blockDim.x) sw3[i] = w[i];
  for (int i = threadIdx.x; i < B * 32; i += blockDim.x) {
    const int r = i >> 5, k = k0 + (i & 31);
    sh[i] = (k < K) ? h[(int64_t)r * K + k] : 0.f;
  }
  __syncthreads();
  // ---- phase 1: logits, softmax, dlogits for every row (warp per row)
  float loss_acc = 0.f;
  unsigned correct_acc = 0;
  for (int row = warp; row < B; row += nwarps) {
    float acc[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) acc[c] = 0.f;
    for (int kb = 0; kb < K; kb += 32 * 8) {
      float hv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int k = kb + u * 32 + lane;
        hv[u] = k < K ? h[(int64_t)row * K + k] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int k = kb + u * 32 + lane;
        if (k < K) {
          const float* wr = sw3 + k * NC;
#pragma unroll
          for (int c = 0; c < 16; ++c)
            if (c < NC) acc[c] += hv[u] * wr[c];
        }
      }
    }
    float z = -INFINITY;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      float v = acc[c];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == c && c < NC) z = v + bias[c];
    }
    float mx = z;
    int am = lane;
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (om > mx || (om == mx && oa < am)) { mx = om; am = oa; }
    }
    const float e = (lane < NC) ? __expf(z - mx) : 0.f;
    float se = e;
    for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
    const int lab = labels[row];
    if (lane < NC) sdl[row * NC + lane] = (e / se - (lane == lab ? 1.f : 0.f)) * scale;
    const float zl = __shfl_sync(0xffffffffu, z, lab);
    if (lane == 0) { loss_acc += logf(se) + mx - zl; correct_acc += (am == lab) ? 1u : 0u; }
  }
  if (blockIdx.x == 0 && lane == 0) {
    atomicAdd(&st->loss_sum, loss_acc);
    if (correct_acc) atomicAdd(&st->correct, correct_acc);
  }
  __syncthreads();
  // ---- phase 2: dW[k, c], db[c], dh[r, k] (+ column sums of dh = bias gradient of the previous layer)
  const int c = lane, kk = warp;
  if (c < NC) {
    for (int ks = kk; ks < 32; ks += nwarps) {
      const int k = k0 + ks;
      if (k >= K) break;
      float s2 = 0.f;
      for (int r = 0; r < B; ++r) s2 += sh[r * 32 + ks] * sdl[r * NC + c];
      dw[(int64_t)k * NC + c] += s2;
    }
    if (blockIdx.x == 0 && kk == 0) {
      float s2 = 0.f;
      for (int r = 0; r < B; ++r) s2 += sdl[r * NC + c];
      db[c] += s2;
    }
  }
  float colsum = 0.f;
  for (int t = threadIdx.x; t < B * 32; t += blockDim.x) {
    const int r = t >> 5, ks = t & 31, k = k0 + ks;
    if (k >= K) continue;
    float s2 = 0.f;
    for (int cc = 0; cc < NC; ++cc) s2 += sdl[r * NC + cc] * sw3[k * NC + cc];
    const uint32_t e = (uint32_t)(r * K + k);
    const bool keep = thresh24 >= (1u << 24) || ddl_dropout_keep(seed, stream, e, thresh24);
    const float o = keep ? s2 * inv_keep : 0.f;
    dh[(int64_t)r * K + k] = o;
    colsum += o;
  }
  if (dbh) {
    colred[warp][lane] = colsum;
    __syncthreads();
    if (threadIdx.x < 32 && k0 + threadIdx.x < K) {
      float t2 = 0.f;
      for (int l = 0; l < nwarps; ++l) t2 += colred[l][threadIdx.x];
      dbh[k0 + threadIdx.x] += t2;
    }
  }
}

__global__ void accuracy_kernel(const float* __restrict__ logits, const int* __restrict__ labels,
                                unsigned* __restrict__ correct, int B, int NC) {
  pdl_enter();
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= B) return;
  float mx = logits[(int64_t)r * NC];
  int am = 0;
  for (int c = 1; c < NC; ++c) {
    float z = logits[(int64_t)r * NC + c];
    if (z > mx) { mx = z; am = c; }
  }
  if (am == labels[r]) atomicAdd(correct, 1u);
}

}  // namespace

// ===================================================================================== C ABI
extern "C" {

int ddl_step_advance(DdlStepState* st, float lr, float b1, float b2, int advance_opt, cudaStream_t s) {
  DDL_LAUNCH(step_advance_kernel, 1, 32, 0, s, st, lr, b1, b2, advance_opt);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_optimizer_step(int kind, float* p, float* m, float* v, const float* g, int64_t n,
                       float step_size, float a, float b, float eps, float gscale,
                       const DdlStepState* st, cudaStream_t s) {
  DDL_REQUIRE(kind >= 0 && kind <= 2, "ddl_optimizer_step: kind %d", kind);
  DDL_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0,
              "ddl_optimizer_step: pointers must be 16-byte aligned");
  if (n <= 0) return 0;
  int64_t n4 = n / 4;
  int grid = grid_for(n4 > 0 ? n4 : 1, kThreads, 148 * 4);
  if (kind == 0) DDL_LAUNCH(optimizer_kernel<0>, grid, kThreads, 0, s, p, m, v, g, n4, n, step_size, a, b, eps, gscale, st);
  else if (kind == 1) DDL_LAUNCH(optimizer_kernel<1>, grid, kThreads, 0, s, p, m, v, g, n4, n, step_size, a, b, eps, gscale, st);
  else DDL_LAUNCH(optimizer_kernel<2>, grid, kThreads, 0, s, p, m, v, g, n4, n, step_size, a, b, eps, gscale, st);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_conv1_relu_pool(const float* x, const float* w, const float* bias, float* y, uint8_t* idx,
                        int B, int S, int C, cudaStream_t s) {
  DDL_REQUIRE(S % 2 == 0 && C % 32 == 0, "ddl_conv1_relu_pool: S even, C %% 32 == 0");
  const size_t smem = ((size_t)26 * C + (size_t)(S + 4) * (S + 4)) * sizeof(float);
  DDL_REQUIRE(smem <= 200 * 1024, "ddl_conv1_relu_pool: filter bank does not fit shared memory");
  if (smem > 48 * 1024) {
    static bool attr_done = false;
    if (!attr_done) {
      DDL_CHECK_CUDA(cudaFuncSetAttribute(conv1_relu_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_done = true;
    }
  }
  DDL_LAUNCH(conv1_relu_pool_kernel, B, 256, smem, s, x, w, bias, y, idx, B, S, C);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_conv1_wgrad(const float* x, const float* dy, float* dw, float* db, int B, int S, int C, cudaStream_t s) {
  DDL_REQUIRE(C % 32 == 0, "ddl_conv1_wgrad: C %% 32 == 0");
  const int strip = (S % 4 == 0) ? S / 4 : (S % 2 == 0 ? S / 2 : S);
  dim3 grid(B * (S / strip), C / 32);
  const size_t smem = (size_t)(strip + 4) * (S + 4) * sizeof(float);
  if (S == 28) DDL_LAUNCH(conv1_wgrad_kernel<28>, grid, 192, smem, s, x, dy, dw, db, B, C, strip);
  else if (S == 32) DDL_LAUNCH(conv1_wgrad_kernel<32>, grid, 192, smem, s, x, dy, dw, db, B, C, strip);
  else DDL_REQUIRE(false, "ddl_conv1_wgrad: image size %d not instantiated (28, 32)", S);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_maxpool_fwd(const float* a, const float* bias, float* y, uint8_t* idx, int B, int H, int W, int C,
                    cudaStream_t s) {
  DDL_REQUIRE(C % 4 == 0, "ddl_maxpool_fwd: C %% 4 == 0");
  int64_t total = (int64_t)B * ((H + 1) / 2) * ((W + 1) / 2) * (C / 4);
  DDL_LAUNCH(maxpool_fwd_kernel, grid_for(total, kThreads, 148 * 16), kThreads, 0, s, a, bias, y, idx, B, H, W, C);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_maxpool_bwd(const float* dp, const uint8_t* idx, float* dy, float* db, int B, int H, int W, int C,
                    cudaStream_t s) {
  DDL_REQUIRE(C % 4 == 0, "ddl_maxpool_bwd: C %% 4 == 0");
  int64_t total = (int64_t)B * ((H + 1) / 2) * ((W + 1) / 2) * (C / 4);
  DDL_LAUNCH(maxpool_bwd_kernel, grid_for(total, kThreads, 148 * 4), kThreads, C * sizeof(float), s, dp, idx, dy, db, B, H, W, C);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_colsum(const float* x, float* out, int R, int C, cudaStream_t s) {
  dim3 grid((C + 127) / 128, R >= 64 ? 8 : 1);
  DDL_LAUNCH(colsum_kernel, grid, 128, 0, s, x, out, R, C);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_fc3_xent_fwd(const float* h, const float* w, const float* bias, const int* labels, float* logits,
                     float* dlogits, DdlStepState* st, int B, int K, int NC, float scale, cudaStream_t s) {
  DDL_REQUIRE(NC <= 16, "ddl_fc3_xent_fwd: at most 16 classes");
  const size_t wbytes = (size_t)K * NC * sizeof(float);
  const int use_smem = wbytes <= 48 * 1024 ? 1 : 0;
  DDL_LAUNCH(fc3_xent_fwd_kernel, (B + 7) / 8, 256, use_smem ? wbytes : 0, s, h, w, bias, labels, logits, dlogits, st, B, K, NC, scale, use_smem);
  DDL_LAUNCH_CHECK();
  return 0;
}

static int fc3_bwd_launch(const float* h, const float* w, const float* dl, float* dw, float* db, float* dh,
                          float* dbh, int B, int K, int NC, uint32_t seed, uint32_t stream, const DdlStepState* st,
                          float keep_prob, cudaStream_t s) {
  size_t smem = ((size_t)B * (NC + 32) + 32 * NC) * sizeof(float);
  DDL_REQUIRE(NC <= 32 && smem <= 200 * 1024, "ddl_fc3_bwd: batch %d too large for smem staging", B);
  uint32_t thresh = keep_prob >= 1.f ? (1u << 24) : (uint32_t)lrintf(keep_prob * (float)(1 << 24));
  if (smem > 48 * 1024)
    DDL_CHECK_CUDA(cudaFuncSetAttribute(fc3_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  DDL_LAUNCH(fc3_bwd_kernel, (K + 31) / 32, 256, smem, s, h, w, dl, dw, db, dh, dbh, B, K, NC, seed, stream, thresh, keep_prob >= 1.f ? 1.f : 1.f / keep_prob, st);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_fc3_bwd(const float* h, const float* w, const float* dl, float* dw, float* db, float* dh, int B, int K,
                int NC, uint32_t seed, uint32_t stream, float keep_prob, cudaStream_t s) {
  return fc3_bwd_launch(h, w, dl, dw, db, dh, nullptr, B, K, NC, seed, stream, nullptr, keep_prob, s);
}

int ddl_fc3_bwd_st(const float* h, const float* w, const float* dl, float* dw, float* db, float* dh, float* dbh,
                   int B, int K, int NC, const DdlStepState* st, uint32_t stream_off, float keep_prob,
                   cudaStream_t s) {
  return fc3_bwd_launch(h, w, dl, dw, db, dh, dbh, B, K, NC, 0, stream_off, st, keep_prob, s);
}

int ddl_fc3_fused_st(const float* h, const float* w, const float* bias, const int* labels, float* dw, float* db,
                     float* dh, float* dbh, DdlStepState* st, int B, int K, int NC, float scale, uint32_t stream_off,
                     float keep_prob, cudaStream_t s) {
  DDL_REQUIRE(NC <= 16 && st != nullptr && labels != nullptr, "ddl_fc3_fused_st: NC <= 16, st and labels required");
  const size_t smem = ((size_t)K * NC + (size_t)B * NC + (size_t)B * 32) * sizeof(float);
  DDL_REQUIRE(smem <= 200 * 1024, "ddl_fc3_fused_st: K=%d, B=%d do not fit shared memory", K, B);
  if (smem > 48 * 1024) {
    static bool attr_done = false;
    if (!attr_done) {
      DDL_CHECK_CUDA(cudaFuncSetAttribute(fc3_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_done = true;
    }
  }
  const uint32_t thresh = keep_prob >= 1.f ? (1u << 24) : (uint32_t)lrintf(keep_prob * (float)(1 << 24));
  DDL_LAUNCH(fc3_fused_kernel, (K + 31) / 32, 256, smem, s, h, w, bias, labels, dw, db, dh, dbh, st, B, K, NC, scale,
             stream_off, thresh, keep_prob >= 1.f ? 1.f : 1.f / keep_prob);
  return 0;
}

int ddl_accuracy(const float* logits, const int* labels, unsigned* correct, int B, int NC, cudaStream_t s) {
  DDL_LAUNCH(accuracy_kernel, (B + 127) / 128, 128, 0, s, logits, labels, correct, B, NC);
  DDL_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
