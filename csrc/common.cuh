// Shared device/host helpers for the ddl_b200 sm_100a kernels.
//
// PTX wrappers for the Blackwell primitives used by csrc/tc_gemm.cu (mbarrier, TMA tiled loads,
// tcgen05 alloc / mma / commit / ld) plus small utilities shared by every translation unit.
// The reference (epikjjh/DIstributed-Deep-Learning) has no native code at all; everything here
// replaces what TensorFlow / MPI did implicitly.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

// ------------------------------------------------------------------------------------ errors
namespace ddl {
void set_error(const char* fmt, ...);
const char* get_error();
}  // namespace ddl

#define DDL_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ddl::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

#define DDL_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      ddl::set_error(__VA_ARGS__);        \
      return -2;                          \
    }                                     \
  } while (0)

#define DDL_LAUNCH_CHECK() DDL_CHECK_CUDA(cudaGetLastError())

// Planning / scheduling experiment switches (tile shapes, ring depths, split policies, ...) are read from the environment
// only in the lab build (python -m ddl_b200.ops.build --lab, used by the sweeps under tools/); the product library always
// takes the measured defaults.  Product switches are the documented ones: DDL_PDL, DDL_WAIT_TIMEOUT_NS (compile time).
#ifdef DDL_TC_LAB
#define DDL_LAB_ENV(name) getenv(name)
inline int ddl_lab_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
#else
#define DDL_LAB_ENV(name) (static_cast<const char*>(nullptr))
inline int ddl_lab_int(const char*, int dflt) { return dflt; }
#endif

namespace ddl {
bool pdl_enabled();        // DDL_PDL=0 disables programmatic dependent launch
// The next launch of this thread carries no programmatic-serialization attribute: it starts only after everything
// before it in its stream has completed.  Used for the first kernel of a forward pass: later kernels prefetch their
// weight operands BEFORE griddepcontrol.wait, which is only sound if no earlier kernel that writes weights (the
// optimizer / pull of the previous step) can still be running - this cuts the chain of early launches there.
void no_pdl_once();
void note_launch();   // bumps the counter behind ddl_launch_count()
// one launch path for every kernel: PDL attribute (+ optional cluster along z)
template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                          unsigned cluster_z, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 1; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = cluster_z;
    ++n;
  }
  cfg.attrs = attr; cfg.numAttrs = n;
  note_launch();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
}  // namespace ddl
#define DDL_LAUNCH(kernel, grid, block, smem, stream, ...) \
  DDL_CHECK_CUDA(ddl::launch(kernel, dim3(grid), dim3(block), smem, stream, 1u, __VA_ARGS__))

// ------------------------------------------------------------------------------------ misc device
#define DDL_DEVICE __device__ __forceinline__

// Time bound for every wait loop in this library (mbarrier waits and cross-GPU flag waits).  A
// wait that exceeds it traps, turning a would-be hang (a dead peer rank, a mis-programmed
// pipeline) into a launch error the host sees - the reference simply blocks forever in MPI
// (SURVEY 5.3).
#ifndef DDL_WAIT_TIMEOUT_NS
#define DDL_WAIT_TIMEOUT_NS 4000000000ull
#endif
#define DDL_SPIN_LIMIT 0x7fffffffu

__device__ __forceinline__ uint64_t ddl_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// call once per failed poll; traps after DDL_WAIT_TIMEOUT_NS of wall time
__device__ __forceinline__ void ddl_spin_guard(uint32_t& spins, uint64_t& t0) {
  if ((++spins & 0x3ff) == 0) {
    const uint64_t now = ddl_globaltimer();
    if (t0 == 0) t0 = now;
    else if (now - t0 > DDL_WAIT_TIMEOUT_NS) __trap();
  }
}

// One elected lane of a fully converged warp (elect.sync).  Keeping the surrounding loop warp-uniform
// and predicating only the single-thread instructions (TMA issue, tcgen05.mma, commit) lets ptxas
// keep addresses and descriptors in uniform registers; an `if (lane == 0)` loop instead forces an
// R2UR per operand and costs ~90 cycles per MMA issue (measured, profiles/tc_issue_latency.md).
DDL_DEVICE bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
      "elect.sync rx|px, 0xffffffff;\n\t"
      "@px mov.s32 %0, 1;\n\t}"
      : "+r"(pred));
  return pred != 0;
}
DDL_DEVICE int warp_id_uniform() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }

DDL_DEVICE uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Counter-based dropout hash.  Must stay bit-identical to ddl_b200/models/oracle.py:dropout_keep_mask.
DDL_DEVICE uint32_t ddl_mix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
DDL_DEVICE bool ddl_dropout_keep(uint32_t seed, uint32_t stream, uint32_t idx, uint32_t thresh24) {
  uint32_t h = idx * 0x9E3779B9u + seed * 0x85EBCA6Bu + stream * 0xC2B2AE35u;
  return (ddl_mix32(h) >> 8) < thresh24;
}

// ------------------------------------------------------------------------------------ mbarrier
DDL_DEVICE void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
DDL_DEVICE void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
DDL_DEVICE void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
DDL_DEVICE void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
DDL_DEVICE void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
DDL_DEVICE bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
DDL_DEVICE bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
DDL_DEVICE void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (!mbar_try_wait(bar, parity)) ddl_spin_guard(spins, t0);
}
// pure polling wait (no hardware suspend)
DDL_DEVICE void mbar_wait_poll(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (!mbar_test_wait(bar, parity)) ddl_spin_guard(spins, t0);
}

// ------------------------------------------------------------------------------------ TMA
DDL_DEVICE void tma_prefetch_desc(const void* desc) {
  asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)desc) : "memory");
}
DDL_DEVICE void tma_load_4d(void* smem_dst, const void* desc, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"((uint64_t)desc), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
DDL_DEVICE void tma_load_3d(void* smem_dst, const void* desc, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"((uint64_t)desc), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// smem -> global tiled store (used by the pull-through GEMM to persist peer-fetched weight tiles)
DDL_DEVICE void tma_store_3d(const void* desc, const void* smem_src, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
      ::"l"((uint64_t)desc), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
DDL_DEVICE void tma_store_4d(const void* desc, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
      ::"l"((uint64_t)desc), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
DDL_DEVICE void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
DDL_DEVICE void tma_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
DDL_DEVICE void tma_store_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ------------------------------------------------------------------------------------ tcgen05
DDL_DEVICE void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DDL_DEVICE void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// whole-warp (.sync.aligned): allocate `ncols` (power of two >= 32) TMEM columns, address -> *smem_slot
DDL_DEVICE void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
DDL_DEVICE void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32 (fp32 storage, tf32 multiply, fp32 accumulate)
DDL_DEVICE void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Predicated forms for WARP-UNIFORM issue loops: every lane runs the loop and computes the (uniform) descriptors, only the
// instruction itself is predicated on the elected lane.  Inside an `if (elect_one()) { loop }` body the operands live in
// per-thread registers and every tcgen05.mma pays a chain of R2UR moves (~100 cycles per MMA measured in the
// patch-resident conv kernel, tools/patch_phase.py); with the loop outside the predicate ptxas keeps them in uniform
// registers.
DDL_DEVICE void umma_tf32_pred(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate, uint32_t issue) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(issue)
      : "memory");
}
DDL_DEVICE void umma_commit_pred(uint64_t* bar, uint32_t issue) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "setp.ne.b32 q, %1, 0;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      ::"r"(smem_u32(bar)), "r"(issue) : "memory");
}
// arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed
DDL_DEVICE void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// TMEM -> registers: this warp's 32 lanes x 16 consecutive fp32 columns
DDL_DEVICE void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor (sm_100 "version 1").
//   bits [0,14)  start address >> 4        bits [16,30) leading-dim byte offset >> 4
//   bits [32,46) stride-dim byte offset >> 4   bits [46,48) version = 1   bits [61,64) layout type
// layout 2 = SWIZZLE_128B (16-byte atoms; K-major operands: 8 rows x 128 B per swizzle atom)
// layout 1 = SWIZZLE_128B_BASE32B (32-byte atoms) - the ONLY legal layout for MN-major 32-bit (tf32)
//            operands: 4 K-rows x 128 B per atom, filled by TMA with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.
#define DDL_UMMA_SW128 2u
#define DDL_UMMA_SW128_BASE32B 1u
DDL_DEVICE uint64_t umma_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// Instruction descriptor for kind::tf32, fp32 accumulate, M=128.
//   [4,6) c_format=1 (F32)  [7,10) a_format=2 (TF32)  [10,13) b_format=2  [15] a_major  [16] b_major
//   [17,23) N>>3  [24,29) M>>4
__host__ __device__ inline uint32_t umma_idesc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4;
  d |= 2u << 7;
  d |= 2u << 10;
  d |= (uint32_t)(a_mn_major & 1) << 15;
  d |= (uint32_t)(b_mn_major & 1) << 16;
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

// ------------------------------------------------------------------------------------ programmatic dependent launch
// Every kernel of this library is launched with programmatic stream serialization (PDL): it may begin
// while its predecessor in the stream is still running, so launch latency, block scheduling and the
// smem / mbarrier / TMEM prologue overlap the predecessor's tail.  Contract for every kernel:
//   pdl_launch_dependents() first (lets the successor start early), then pdl_wait() before the first
//   access to global memory - it returns once the predecessor grid has completed and flushed.
// EVERY thread of EVERY kernel executes pdl_wait(), also on early-exit paths: a grid that finishes
// without waiting would let its own successor overtake the grid it skipped.
DDL_DEVICE void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
DDL_DEVICE void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
DDL_DEVICE void pdl_enter() { pdl_launch_dependents(); pdl_wait(); }

// ------------------------------------------------------------------------------------ thread-block clusters
DDL_DEVICE uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
DDL_DEVICE void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
DDL_DEVICE void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// distributed shared memory: address of `local_smem_addr` in the CTA with cluster rank `rank`
DDL_DEVICE uint32_t cluster_map_shared(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
DDL_DEVICE float4 ld_dsmem_v4(uint32_t cluster_addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(cluster_addr) : "memory");
  return v;
}
DDL_DEVICE void st_dsmem_f32(uint32_t cluster_addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(cluster_addr), "f"(v) : "memory");
}
// named barrier over `nthreads` threads (id 1..15; 0 is __syncthreads)
DDL_DEVICE void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// ------------------------------------------------------------------------------------ system-scope sync
DDL_DEVICE void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DDL_DEVICE uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
DDL_DEVICE uint32_t atom_add_release_sys(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.add.release.sys.global.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
// NVSwitch multicast (NVLS): `p` must be a multicast address
DDL_DEVICE float4 multimem_ld_reduce_add_v4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
DDL_DEVICE void multimem_st_v4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
DDL_DEVICE void red_add_v4_f32(float* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
