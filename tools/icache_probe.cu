// Instruction-fetch probe: how fast does ONE warp walk straight-line code it has never executed on this SM, with the code
// (a) resident in L2, (b) evicted from L2 as well (256 MiB written in between, as bench.py does between timed steps)?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/icache_probe.bin tools/icache_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
template <int N, int SALT>
__global__ void walk(unsigned* out, unsigned long long* ns, unsigned seed) {
  unsigned a = seed, b = seed * 3 + SALT, c = seed ^ 5, d = seed + 7;
  const unsigned long long t0 = gt();
  a += (unsigned)(t0 >> 63);              // the chain starts after the first timer read ...
#pragma unroll
  for (int i = 0; i < N; ++i) {          // a dependent chain of 4 * N IMADs, fully unrolled: no loop, 16 B per instruction
    asm volatile("mad.lo.u32 %0, %0, 3, %1;" : "+r"(a) : "r"(b));
    asm volatile("mad.lo.u32 %0, %0, 5, %1;" : "+r"(b) : "r"(c));
    asm volatile("mad.lo.u32 %0, %0, 7, %1;" : "+r"(c) : "r"(d));
    asm volatile("mad.lo.u32 %0, %0, 9, %1;" : "+r"(d) : "r"(a));
  }
  unsigned long long t1;                  // ... and the second read consumes its result
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1) : "r"(a ^ b ^ c ^ d));
  if (threadIdx.x == 0) { out[blockIdx.x] = a ^ b ^ c ^ d; if (blockIdx.x == 0) ns[0] = t1 - t0; }
}
__global__ void fill(float4* p, size_t n) { for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = make_float4(1, 2, 3, 4); }
int main() {
  unsigned* out; unsigned long long* ns; float4* big;
  cudaMalloc(&out, 4096); cudaMalloc(&ns, 8); cudaMalloc(&big, 256u << 20);
  unsigned long long h = 0;
  auto run = [&](const char* name, auto k, int blocks) {
    k<<<blocks, 32>>>(out, ns, 1u);
    cudaError_t e1 = cudaGetLastError(), e2 = cudaDeviceSynchronize(), e3 = cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost);
    if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) printf("error: %s / %s / %s\n", cudaGetErrorString(e1), cudaGetErrorString(e2), cudaGetErrorString(e3));
    printf("%-62s %8.2f us\n", name, h / 1e3);
  };
  constexpr int N = 2048;                // 8192 IMADs = 128 KiB of SASS
  for (int rep = 0; rep < 2; ++rep) {
    run("first launch (everything cold)                        1 block", walk<N, 1>, 1);
    run("second launch, same code (warm)                       1 block", walk<N, 1>, 1);
    walk<N, 2><<<148 * 4, 32>>>(out, ns, 2u); cudaDeviceSynchronize();               // other code on every SM: evicts the SM i-caches
    run("after another 128 KiB kernel ran on all SMs (L2-warm) 1 block", walk<N, 1>, 1);
    walk<N, 2><<<148 * 4, 32>>>(out, ns, 2u); fill<<<148 * 8, 256>>>(big, (256u << 20) / 16); cudaDeviceSynchronize();
    run("after that AND a 256 MiB fill (L2-cold)               1 block", walk<N, 1>, 1);
    walk<N, 2><<<148 * 4, 32>>>(out, ns, 2u); fill<<<148 * 8, 256>>>(big, (256u << 20) / 16); cudaDeviceSynchronize();
    run("same, 148 blocks walk the code at once (L2-cold)             ", walk<N, 1>, 148);
  }
  printf("code walked: %d instructions = %d KiB\n", N * 4, N * 4 * 16 / 1024);
  return 0;
}
