// Thread-local error string behind ddl_last_error().
#include "common.cuh"
#include "../include/ddl_b200.h"
#include <stdarg.h>
#include <stdlib.h>
#include <atomic>

namespace ddl {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }
static thread_local bool g_no_pdl_once = false;
void no_pdl_once() { g_no_pdl_once = true; }
bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("DDL_PDL"); return !(e && e[0] == '0'); }();
  if (g_no_pdl_once) { g_no_pdl_once = false; return false; }
  return on;
}
static std::atomic<unsigned long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
}  // namespace ddl

extern "C" const char* ddl_last_error(void) { return ddl::get_error(); }
extern "C" int ddl_version(void) { return 200; }
/* kernels launched by this library in this process since it was loaded (every launch goes through ddl::launch; a launch
 * recorded during stream capture counts once, however often the graph is replayed) */
extern "C" unsigned long long ddl_launch_count(void) { return ddl::g_launches.load(std::memory_order_relaxed); }

extern "C" int ddl_memcpy_async(void* dst, const void* src, size_t bytes, cudaStream_t s) {
  DDL_CHECK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, s));
  return 0;
}
