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
    st->loss_mean = 0.f;
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

// conv1 weight / bias gradient from the POOLED layout: after pool + ReLU backward only the argmax pixel of every
// 2x2 window carries gradient, so the sum runs over B * (S/2)^2 windows instead of B * S^2 pixels (4x less work, and
// the 75% zeros of dY are never read).  idx [B, S/2, S/2, C] holds the window slot (0..3, 4 = dead).
// Block = one image x one group of 32 channels, 160 threads = 32 channels x 5 filter rows.
//   phase 1 (all threads, batched loads): slot + gradient of every window -> shared memory;
//   phase 2: thread (c, kh) walks the windows: 5 LDS of the x row under filter row kh + 5 FMAs each.
template <int S>
__global__ void __launch_bounds__(160) conv1_wgrad_sparse_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                 const uint8_t* __restrict__ idx, float* __restrict__ dw,
                                                                 float* __restrict__ db, int B, int C) {
  pdl_launch_dependents();
  constexpr int PS = (S + 1) / 2, XP = S + 4, NW = PS * PS;
  extern __shared__ __align__(16) float ssm[];
  float* sx = ssm;                               // [(S + 4)][XP] zero-padded image
  float* sg = sx + (S + 4) * XP;                 // [NW][32] gradient at the argmax pixel (0 when dead)
  int* so = reinterpret_cast<int*>(sg + NW * 32); // [NW][32] offset of that pixel's top-left tap in sx
  const int n = blockIdx.x, cg = blockIdx.y * 32;
  const int t = threadIdx.x, lane = t & 31, kh = t >> 5;
  // the image is an input of the step: fetch it before waiting for the kernel that produces dY
  for (int i = t; i < (S + 4) * XP; i += 160) {
    const int r = i / XP - 2, cc = i % XP - 2;
    sx[i] = (r >= 0 && r < S && cc >= 0 && cc < S) ? __ldg(x + (int64_t)n * S * S + r * S + cc) : 0.f;
  }
  pdl_wait();
  // phase 1: 8 windows per trip per thread, slot loads first, then the dependent gradient loads
  for (int base = 0; base < NW * 32; base += 160 * 8) {
    int b[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = base + u * 160 + t;
      b[u] = 4;
      if (i < NW * 32) b[u] = idx[((int64_t)n * NW + (i >> 5)) * C + cg + (i & 31)];
    }
    float g[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = base + u * 160 + t;
      g[u] = 0.f;
      if (b[u] < 4) {
        const int w = i >> 5, ph = w / PS, pw = w % PS;
        const int uh = 2 * ph + (b[u] >> 1), uw = 2 * pw + (b[u] & 1);
        g[u] = dy[(((int64_t)n * S + uh) * S + uw) * C + cg + (i & 31)];
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = base + u * 160 + t;
      if (i < NW * 32) {
        const int w = i >> 5, ph = w / PS, pw = w % PS;
        const int bb = b[u] < 4 ? b[u] : 0;
        sg[i] = g[u];
        so[i] = (2 * ph + (bb >> 1)) * XP + 2 * pw + (bb & 1);
      }
    }
  }
  __syncthreads();
  // phase 2
  float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  float bsum = 0.f;
  const float* xk = sx + kh * XP;
#pragma unroll 4
  for (int w = 0; w < NW; ++w) {
    const float g = sg[w * 32 + lane];
    const float* xr = xk + so[w * 32 + lane];
    acc[0] += g * xr[0]; acc[1] += g * xr[1]; acc[2] += g * xr[2]; acc[3] += g * xr[3]; acc[4] += g * xr[4];
    bsum += g;
  }
#pragma unroll
  for (int kw = 0; kw < 5; ++kw) atomicAdd(dw + (kh * 5 + kw) * C + cg + lane, acc[kw]);
  if (kh == 0 && db) atomicAdd(db + cg + lane, bsum);
}

// conv1 weight gradient from the POOLED gradient gp [B, S/2, S/2, C] (the compact output of conv2's dgrad,
// ddl_tc_plan_dgrad_poolgrad: ReLU-dead windows are already 0) and the argmax slots idx: the un-pooled plane is never
// built.  Block = one image x 32 channels, 320 threads = 2 window halves x 5 filter rows x 32 channels.
//   before griddepcontrol.wait: the zero-padded image -> shared memory (an input of the step);
//   phase 1: gp (float4) and idx (uchar4) of the whole image, one level of coalesced loads, every load issued before
//            the first shared store -> sg[w][c] (value), so[w][c] (offset of the argmax pixel's top-left tap);
//   phase 2: thread (half, kh, c) walks its 98 windows: 2 LDS + 5 LDS of x under filter row kh + 5 FMAs each.  The row
//            pitch XP = S + 6 puts the four possible argmax pixels of a window in four different banks.
template <int S, int SPLIT>
__global__ void __launch_bounds__(320) conv1_wgrad_pooled_kernel(const float* __restrict__ x, const float* __restrict__ gp,
                                                                 const uint8_t* __restrict__ idx, float* __restrict__ dw,
                                                                 float* __restrict__ db, int B, int C) {
  pdl_launch_dependents();
  constexpr int PS = S / 2, XP = S + 6, NW = PS * PS / SPLIT, NV = NW * 8;   // NW windows, NV float4 / uchar4 items per block
  constexpr int kT = 320, kTrips = (NV + kT - 1) / kT;
  static_assert((PS * PS) % (2 * SPLIT) == 0, "windows must split evenly");
  extern __shared__ __align__(16) float ssm[];
  float* sx = ssm;                                       // [(S + 4)][XP]
  float* sg = sx + (S + 4) * XP;                         // [NW][32]
  unsigned short* so = reinterpret_cast<unsigned short*>(sg + NW * 32);   // [NW][32]
  float* sred = reinterpret_cast<float*>(so + NW * 32);  // [26][32] partial sums of the second half
  const int n = blockIdx.x / SPLIT, w0 = (blockIdx.x % SPLIT) * NW, cg = blockIdx.y * 32;
  const int t = threadIdx.x, lane = t & 31, kh = (t >> 5) % 5, half = t / 160;
  for (int i = t; i < (S + 4) * XP; i += kT) {
    const int r = i / XP - 2, cc = i % XP - 2;
    sx[i] = (r >= 0 && r < S && cc >= 0 && cc < S) ? __ldg(x + (int64_t)n * S * S + r * S + cc) : 0.f;
  }
  pdl_wait();
  {
    float4 g4[kTrips];
    uchar4 b4[kTrips];
#pragma unroll
    for (int u = 0; u < kTrips; ++u) {
      const int i = u * kT + t;                          // local window = i / 8, channel quad = i % 8
      g4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      b4[u] = make_uchar4(0, 0, 0, 0);
      if (i < NV) {
        const int64_t o = ((int64_t)n * PS * PS + w0 + (i >> 3)) * C + cg + 4 * (i & 7);
        g4[u] = __ldcg(reinterpret_cast<const float4*>(gp + o));
        b4[u] = *reinterpret_cast<const uchar4*>(idx + o);
      }
    }
#pragma unroll
    for (int u = 0; u < kTrips; ++u) {
      const int i = u * kT + t;
      if (i < NV) {
        const int wl = i >> 3, w = w0 + wl, ph = w / PS, pw = w % PS;
        const int base = 2 * ph * XP + 2 * pw;
        const int bb[4] = {b4[u].x & 3, b4[u].y & 3, b4[u].z & 3, b4[u].w & 3};    // dead windows carry g = 0
        *reinterpret_cast<float4*>(sg + wl * 32 + 4 * (i & 7)) = g4[u];
        ushort4 ov;
        ov.x = (unsigned short)(base + (bb[0] >> 1) * XP + (bb[0] & 1));
        ov.y = (unsigned short)(base + (bb[1] >> 1) * XP + (bb[1] & 1));
        ov.z = (unsigned short)(base + (bb[2] >> 1) * XP + (bb[2] & 1));
        ov.w = (unsigned short)(base + (bb[3] >> 1) * XP + (bb[3] & 1));
        *reinterpret_cast<ushort4*>(so + wl * 32 + 4 * (i & 7)) = ov;
      }
    }
  }
  __syncthreads();
  float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  float bsum = 0.f;
  const float* xk = sx + kh * XP;
  constexpr int HW = NW / 2;
#pragma unroll 7
  for (int w = half * HW; w < (half + 1) * HW; ++w) {
    const float g = sg[w * 32 + lane];
    const float* xr = xk + so[w * 32 + lane];
    acc[0] += g * xr[0]; acc[1] += g * xr[1]; acc[2] += g * xr[2]; acc[3] += g * xr[3]; acc[4] += g * xr[4];
    bsum += g;
  }
  if (half == 1) {
#pragma unroll
    for (int kw = 0; kw < 5; ++kw) sred[(kh * 5 + kw) * 32 + lane] = acc[kw];
    if (kh == 0) sred[25 * 32 + lane] = bsum;
  }
  __syncthreads();
  if (half == 0) {
#pragma unroll
    for (int kw = 0; kw < 5; ++kw) {
      const float v = acc[kw] + sred[(kh * 5 + kw) * 32 + lane];
      if (v != 0.f) atomicAdd(dw + (kh * 5 + kw) * C + cg + lane, v);
    }
    if (kh == 0 && db) atomicAdd(db + cg + lane, bsum + sred[25 * 32 + lane]);
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
    atomicAdd(&st->loss_mean, (logf(se) + mx - zl) * scale);
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

// fc3 forward + softmax cross-entropy + fc3 backward in ONE kernel (training path): a thread-block cluster of
// kFc3Cluster CTAs x 1024 threads.  CTA r computes the logits / softmax / dlogits of rows r, r + CS, ... (warp per
// row) and stores each dlogits value into EVERY CTA's shared memory (st.shared::cluster); after one cluster barrier
// CTA r owns the k-columns [r * KS, (r + 1) * KS) of the backward: dW3, dh2 (+ dropout mask) and the column sums of
// dh2 (= fc2's bias gradient).  What shaped it (tools/fc3_phase.py, tools/timeline.py): the two-kernel version took
// 16 + 15 us for ~3 MFLOP because 256-thread blocks walked long dependent LDS -> FMA chains with runtime div / mod
// per output.  Here: the class count is a compile-time padded width (NCT = 12 or 16, zero columns), every inner loop
// is vector LDS + independent FMAs, 32 warps per SM hide the latencies, everything that does not depend on the
// previous kernel (W3 -> shared memory, bias, labels) is loaded BEFORE griddepcontrol.wait.
constexpr int kFc3Cluster = 8;
constexpr int kFc3Threads = 1024;
// perf experiments: globaltimer stamps of CTA 0 / thread 0 (tools/fc3_phase.py); off unless ddl_debug_stamps(1)
static int g_stamps_host = 0;                 // passed as a kernel argument: no device-side flag load on the hot path
__device__ unsigned long long g_stamps[32];
#define DDL_STAMP(i) do { if (stamps && threadIdx.x == 0) g_stamps[i] = ddl_globaltimer(); } while (0)

template <int NCT>
__global__ void __launch_bounds__(kFc3Threads) fc3_cluster_kernel(const float* __restrict__ h, const float* __restrict__ w,
                                                                  const float* __restrict__ bias, const int* __restrict__ labels,
                                                                  float* __restrict__ dw, float* __restrict__ db,
                                                                  float* __restrict__ dh, float* __restrict__ dbh,
                                                                  DdlStepState* st, int B, int K, int NC, int KS, float scale,
                                                                  uint32_t stream_off, uint32_t thresh24, float inv_keep,
                                                                  int stamps_on) {
  pdl_launch_dependents();
  cluster_arrive_release();                // "I am running": peers may store into my shared memory after the matching wait
  extern __shared__ __align__(16) float sm[];
  float* swT = sm;                         // [NCT][K]    W3 transposed (rows >= NC are zero): lanes along k, LDS.128
  float* sdl = swT + NCT * K;              // [B][NCT]    dlogits of every row (filled by all CTAs of the cluster)
  float* shs = sdl + B * NCT;              // [B][KS + 1] this CTA's k-columns of h
  float* scol = shs + B * (KS + 1);        // [KS]        column sums of dh
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int crank = (int)cluster_ctarank();
  const bool stamps = stamps_on && crank == 0;
  DDL_STAMP(0);
  const int k0 = crank * KS;
  // ---- phase 0: loads that do not depend on the previous kernel
  {
    const int n4 = (K * NC) >> 2;
    for (int base = 0; base < n4; base += kFc3Threads * 2) {
      float4 buf[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i4 = base + u * kFc3Threads + t;
        if (i4 < n4) buf[u] = __ldg(reinterpret_cast<const float4*>(w) + i4);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i4 = base + u * kFc3Threads + t;
        if (i4 < n4) {
          const float v4[4] = {buf[u].x, buf[u].y, buf[u].z, buf[u].w};
          int kk = (4 * i4) / NC, cc = (4 * i4) % NC;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            swT[cc * K + kk] = v4[j];
            if (++cc == NC) { cc = 0; ++kk; }
          }
        }
      }
    }
    for (int i = t; i < (NCT - NC) * K; i += kFc3Threads) swT[NC * K + i] = 0.f;
    for (int i = t; i < KS; i += kFc3Threads) scol[i] = 0.f;
  }
  const float bz = lane < NC ? __ldg(bias + lane) : 0.f;
  const int my_row = crank + kFc3Cluster * warp;        // one row per warp: B <= 32 * kFc3Cluster
  const int lab = my_row < B ? __ldg(labels + my_row) : 0;
  DDL_STAMP(1);
  pdl_wait();                                // h (and the step state) come from the previous kernels
  DDL_STAMP(2);
  const uint32_t seed = st->seed, stream = st->stream + stream_off;
  // this CTA's k-columns of h for the backward: issued now, parked in shared memory after the logits
  constexpr int kHs = 2;
  float4 hs[kHs];
  const int ks4 = KS >> 2, nslice4 = B * ks4;
#pragma unroll
  for (int u = 0; u < kHs; ++u) {
    const int o4 = u * kFc3Threads + t;
    hs[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (o4 < nslice4) {
      const int r = o4 / ks4, k = k0 + 4 * (o4 % ks4);
      if (k < K) hs[u] = *reinterpret_cast<const float4*>(h + (int64_t)r * K + k);
    }
  }
  // my row of h: lanes own 4 consecutive k per 128-column chunk
  float4 hv[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int k = 4 * lane + 128 * u;
    hv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (my_row < B && k < K) hv[u] = *reinterpret_cast<const float4*>(h + (int64_t)my_row * K + k);
  }
  __syncthreads();                           // swT complete
  DDL_STAMP(3);
  cluster_wait_acquire();                    // every CTA of the cluster has started
  DDL_STAMP(4);
  // ---- phase 1: logits, softmax, dlogits of my row, pushed into every CTA's sdl
  if (my_row < B) {
    float acc[NCT];
#pragma unroll
    for (int c = 0; c < NCT; ++c) acc[c] = 0.f;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = 4 * lane + 128 * u;
      if (k < K) {
#pragma unroll
        for (int c = 0; c < NCT; ++c) {
          const float4 w4 = *reinterpret_cast<const float4*>(swT + c * K + k);
          acc[c] += hv[u].x * w4.x + hv[u].y * w4.y + hv[u].z * w4.z + hv[u].w * w4.w;
        }
      }
    }
    for (int kb = 512; kb < K; kb += 128) {            // K > 512: the rest of the row, one chunk at a time
      const int k = kb + 4 * lane;
      if (k < K) {
        const float4 h4 = *reinterpret_cast<const float4*>(h + (int64_t)my_row * K + k);
#pragma unroll
        for (int c = 0; c < NCT; ++c) {
          const float4 w4 = *reinterpret_cast<const float4*>(swT + c * K + k);
          acc[c] += h4.x * w4.x + h4.y * w4.y + h4.z * w4.z + h4.w * w4.w;
        }
      }
    }
    float z = -INFINITY;
#pragma unroll
    for (int c = 0; c < NCT; ++c) {
      float v = acc[c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == c && c < NC) z = v + bz;
    }
    float mx = z;
    int am = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, am, o);
      if (om > mx || (om == mx && oa < am)) { mx = om; am = oa; }
    }
    const float e = (lane < NC) ? __expf(z - mx) : 0.f;
    float se = e;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
    if (lane < NCT) {
      const float d = lane < NC ? (e / se - (lane == lab ? 1.f : 0.f)) * scale : 0.f;
      const uint32_t a = smem_u32(sdl) + (uint32_t)((my_row * NCT + lane) * 4);
#pragma unroll
      for (int pr = 0; pr < kFc3Cluster; ++pr) st_dsmem_f32(cluster_map_shared(a, (uint32_t)pr), d);
    }
    const float zl = __shfl_sync(0xffffffffu, z, lab);
    if (lane == 0) {
      atomicAdd(&st->loss_sum, logf(se) + mx - zl);
      atomicAdd(&st->loss_mean, (logf(se) + mx - zl) * scale);
      if (am == lab) atomicAdd(&st->correct, 1u);
    }
  }
  DDL_STAMP(5);
#pragma unroll
  for (int u = 0; u < kHs; ++u) {
    const int o4 = u * kFc3Threads + t;
    if (o4 < nslice4) {
      float* d = shs + (o4 / ks4) * (KS + 1) + 4 * (o4 % ks4);
      d[0] = hs[u].x; d[1] = hs[u].y; d[2] = hs[u].z; d[3] = hs[u].w;
    }
  }
  for (int o4 = kHs * kFc3Threads + t; o4 < nslice4; o4 += kFc3Threads) {     // beyond the register-staged batches
    const int r = o4 / ks4, k = k0 + 4 * (o4 % ks4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K) v = *reinterpret_cast<const float4*>(h + (int64_t)r * K + k);
    float* d = shs + r * (KS + 1) + 4 * (o4 % ks4);
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
  }
  DDL_STAMP(6);
  cluster_arrive_release();                  // every CTA's dlogits have landed in every sdl; shs / scol complete
  cluster_wait_acquire();
  DDL_STAMP(7);
  // ---- phase 2: backward of my k-columns.  KS is a power of two in [32, 256]: t -> (ks, group) without divisions.
  const int ks = t & (KS - 1), grp = t / KS, ngrp = kFc3Threads / KS;
  const int k = k0 + ks;
  {  // dW3[k, c] += sum_r h[r, k] dl[r, c]: group = (column quarter, row partition)
    constexpr int CQ = NCT / 4;
    const int cq = grp & 3, rp = grp >> 2, nrp = ngrp >> 2;
    float a[CQ];
#pragma unroll
    for (int j = 0; j < CQ; ++j) a[j] = 0.f;
    for (int r = rp; r < B; r += nrp) {
      const float hval = shs[r * (KS + 1) + ks];
#pragma unroll
      for (int j = 0; j < CQ; ++j) a[j] += hval * sdl[r * NCT + cq * CQ + j];
    }
    if (k < K) {
#pragma unroll
      for (int j = 0; j < CQ; ++j)
        if (cq * CQ + j < NC) atomicAdd(dw + (int64_t)k * NC + cq * CQ + j, a[j]);
    }
  }
  DDL_STAMP(8);
  if (crank == 0 && warp < NC) {             // db3[c] += sum_r dl[r, c]: warp c, lanes over rows
    float s2 = 0.f;
    for (int r = lane; r < B; r += 32) s2 += sdl[r * NCT + warp];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    if (lane == 0) atomicAdd(db + warp, s2);
  }
  DDL_STAMP(9);
  {  // dh2[r, k] = (sum_c dl[r, c] W3[k, c]) * mask / keep: a thread keeps W3[k, :] in registers and walks rows
    float wk[NCT];
#pragma unroll
    for (int c = 0; c < NCT; ++c) wk[c] = k < K ? swT[c * K + k] : 0.f;
    float colsum = 0.f;
    for (int r = grp; r < B; r += ngrp) {
      float s2 = 0.f;
#pragma unroll
      for (int c4 = 0; c4 < NCT; c4 += 4) {
        const float4 d4 = *reinterpret_cast<const float4*>(sdl + r * NCT + c4);       // broadcast
        s2 += d4.x * wk[c4] + d4.y * wk[c4 + 1] + d4.z * wk[c4 + 2] + d4.w * wk[c4 + 3];
      }
      if (k < K) {
        const uint32_t e = (uint32_t)(r * K + k);
        const bool keep = thresh24 >= (1u << 24) || ddl_dropout_keep(seed, stream, e, thresh24);
        const float v = keep ? s2 * inv_keep : 0.f;
        dh[(int64_t)r * K + k] = v;
        colsum += v;
      }
    }
    if (dbh && colsum != 0.f) atomicAdd(scol + ks, colsum);
  }
  DDL_STAMP(10);
  if (dbh) {
    __syncthreads();
    for (int i = t; i < KS; i += kFc3Threads)
      if (k0 + i < K && scol[i] != 0.f) atomicAdd(dbh + k0 + i, scol[i]);
  }
  DDL_STAMP(11);
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

/* conv1 weight / bias gradient from the pooled layout (see conv1_wgrad_sparse_kernel): dy is the un-pooled gradient
 * [B,S,S,C] (only its argmax pixels are read), idx [B,S/2,S/2,C] the argmax slots.  x is read before
 * griddepcontrol.wait: it must not be produced by the kernel launched just before. */
int ddl_conv1_wgrad_sparse(const float* x, const float* dy, const uint8_t* idx, float* dw, float* db, int B, int S, int C,
                           cudaStream_t s) {
  DDL_REQUIRE(C % 32 == 0 && S == 28, "ddl_conv1_wgrad_sparse: C %% 32 == 0 and S == 28 (got C=%d S=%d)", C, S);
  constexpr int S28 = 28, PS = 14;
  const size_t smem = ((size_t)(S28 + 4) * (S28 + 4) + (size_t)PS * PS * 32 * 2) * sizeof(float);
  static bool attr_done = false;
  if (!attr_done) {
    DDL_CHECK_CUDA(cudaFuncSetAttribute(conv1_wgrad_sparse_kernel<28>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  DDL_LAUNCH(conv1_wgrad_sparse_kernel<28>, dim3(B, C / 32), 160, smem, s, x, dy, idx, dw, db, B, C);
  DDL_LAUNCH_CHECK();
  return 0;
}

int ddl_conv1_wgrad_pooled(const float* x, const float* gp, const uint8_t* idx, float* dw, float* db, int B, int S, int C,
                           cudaStream_t s) {
  DDL_REQUIRE(C % 32 == 0 && S == 28, "ddl_conv1_wgrad_pooled: C %% 32 == 0 and S == 28 (got C=%d S=%d)", C, S);
  // two blocks per image (98 windows each): 200 blocks, two resident per SM - twice the warps per SM of one block per
  // image, and every SM gets work (100 blocks left 48 SMs idle)
  constexpr int S28 = 28, SPLIT = 2, NW = 14 * 14 / SPLIT;
  const size_t smem = (size_t)(S28 + 4) * (S28 + 6) * sizeof(float) + (size_t)NW * 32 * (sizeof(float) + sizeof(unsigned short)) +
                      26 * 32 * sizeof(float);
  DDL_LAUNCH((conv1_wgrad_pooled_kernel<28, SPLIT>), dim3(B * SPLIT, C / 32), 320, smem, s, x, gp, idx, dw, db, B, C);
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

static int fc3_fused_ks(int K) {
  int ks = 32;
  while (ks * kFc3Cluster < K) ks <<= 1;
  return ks;
}

static size_t fc3_fused_smem(int B, int K, int NC) {
  const int nct = NC <= 12 ? 12 : 16, KS = fc3_fused_ks(K);
  return ((size_t)K * nct + (size_t)B * nct + (size_t)B * (KS + 1) + KS) * sizeof(float);
}

/* 1 when ddl_fc3_fused_st can run this shape (shared memory, one row per warp, k-slice a power of two <= 256) */
int ddl_fc3_fused_fits(int B, int K, int NC) {
  if (NC > 16 || NC < 1 || K % 4 || B > 32 * kFc3Cluster || K > 256 * kFc3Cluster) return 0;
  return fc3_fused_smem(B, K, NC) <= 200 * 1024 ? 1 : 0;
}

int ddl_fc3_fused_st(const float* h, const float* w, const float* bias, const int* labels, float* dw, float* db,
                     float* dh, float* dbh, DdlStepState* st, int B, int K, int NC, float scale, uint32_t stream_off,
                     float keep_prob, cudaStream_t s) {
  DDL_REQUIRE(st != nullptr && labels != nullptr, "ddl_fc3_fused_st: st and labels required");
  DDL_REQUIRE(ddl_fc3_fused_fits(B, K, NC), "ddl_fc3_fused_st: shape B=%d K=%d NC=%d does not fit (see ddl_fc3_fused_fits)", B, K, NC);
  const int KS = fc3_fused_ks(K);
  const size_t smem = fc3_fused_smem(B, K, NC);
  static bool attr_done = false;
  if (!attr_done) {
    DDL_CHECK_CUDA(cudaFuncSetAttribute(fc3_cluster_kernel<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    DDL_CHECK_CUDA(cudaFuncSetAttribute(fc3_cluster_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_done = true;
  }
  const uint32_t thresh = keep_prob >= 1.f ? (1u << 24) : (uint32_t)lrintf(keep_prob * (float)(1 << 24));
  const float inv_keep = keep_prob >= 1.f ? 1.f : 1.f / keep_prob;
  if (NC <= 12)
    DDL_CHECK_CUDA(ddl::launch(fc3_cluster_kernel<12>, dim3(1, 1, kFc3Cluster), dim3(kFc3Threads), smem, s, (unsigned)kFc3Cluster,
                               h, w, bias, labels, dw, db, dh, dbh, st, B, K, NC, KS, scale, stream_off, thresh, inv_keep, g_stamps_host));
  else
    DDL_CHECK_CUDA(ddl::launch(fc3_cluster_kernel<16>, dim3(1, 1, kFc3Cluster), dim3(kFc3Threads), smem, s, (unsigned)kFc3Cluster,
                               h, w, bias, labels, dw, db, dh, dbh, st, B, K, NC, KS, scale, stream_off, thresh, inv_keep, g_stamps_host));
  DDL_LAUNCH_CHECK();
  return 0;
}

/* perf experiments: enable / read the globaltimer stamps of fc3_cluster_kernel */
int ddl_debug_stamps(int on, unsigned long long* out, int n) {
  g_stamps_host = on;
  if (out && n > 0) DDL_CHECK_CUDA(cudaMemcpyFromSymbol(out, g_stamps, sizeof(unsigned long long) * (n > 32 ? 32 : n)));
  return 0;
}

int ddl_accuracy(const float* logits, const int* labels, unsigned* correct, int B, int NC, cudaStream_t s) {
  DDL_LAUNCH(accuracy_kernel, (B + 127) / 128, 128, 0, s, logits, labels, correct, B, NC);
  DDL_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
