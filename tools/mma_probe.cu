// tcgen05.mma kind::tf32 issue / execution rate probe: ONE CTA issues `iters` x 4 MMAs (128 x N x 8, operands in shared
// memory, SWIZZLE_128B K-major A, K-major B) from a warp-uniform loop and times them with clock64 until the commit lands.
// Varied: N, the A start address (1024-aligned or shifted by whole 128-byte rows), cta occupancy (1 or 2 CTAs per SM).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc -o tools/mma_probe.bin tools/mma_probe.cu
#include <cstdio>
#include "common.cuh"
namespace ddl { void set_error(const char*, ...) {} const char* get_error() { return ""; } bool pdl_enabled() { return false; } void no_pdl_once() {} void note_launch() {} }

__global__ void __launch_bounds__(128, 1) mma_rate(int N, int shift_rows, int iters, int rotate, int alt, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t taddr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 1.0f;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc(&taddr, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    const uint32_t issue = elect_one() ? 1u : 0u;
    const uint32_t idesc = umma_idesc_tf32(128, N, 0, 0);
    const uint64_t tmpl = umma_smem_desc(0, 16, 1024, DDL_UMMA_SW128);
    const uint32_t a0 = (smem_u32(smem) + 1024 + shift_rows * 128) >> 4;        // A: 128+ rows of 128 B
    const uint32_t b0 = (smem_u32(smem) + 48 * 1024) >> 4;                        // B: N rows of 128 B
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t ar = rotate ? (uint32_t)((i & 7) * (128 >> 4)) : 0u;         // rotate: a different row shift every iteration
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = tmpl | (uint64_t)((a0 + ar + 2u * k) & 0x3FFFu);
        const uint64_t bd = tmpl | (uint64_t)((b0 + 2u * k) & 0x3FFFu);
        umma_tf32_pred(taddr + (alt ? (uint32_t)((i & 1) * N) : 0u), ad, bd, idesc, 1u, issue);
      }
    }
    const long long t1 = clock64();
    umma_commit_pred(&bar, issue);
    mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (lane == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(taddr, 512);
}

int main() {
  long long* out; cudaMalloc(&out, 16);
  cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const int iters = 500;
  printf("%-6s %-10s %-8s %-8s | cycles per MMA (issue loop) | cycles per MMA (until all complete)\n", "N", "A shift", "rotate", "CTAs");
  for (int ctas : {1, 148})
    for (int N : {32, 64, 128, 256})
      for (int alt : {0, 1})
        for (int rot : {0, 1}) {
          const int shift = 0;
          mma_rate<<<ctas, 128, 98 * 1024>>>(N, shift, iters, rot, alt, out);
          long long h[2];
          cudaError_t e = cudaDeviceSynchronize();
          cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          printf("%-6d alt-acc %-3d %-8d %-8d | %8.1f | %8.1f\n", N, alt, rot, ctas, h[0] / (4.0 * iters), h[1] / (4.0 * iters));
        }
  return 0;
}
