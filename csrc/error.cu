// Error reporting and launch accounting shared by every entry point.
#include <atomic>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "jllama_b200.h"

namespace jl {
static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
unsigned long long launches() { return g_launches.load(std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return JL_ERR_CUDA;
}
}  // namespace jl

extern "C" JL_API const char* jl_last_error(void) { return jl::g_err; }
extern "C" JL_API int jl_version(void) { return 100; }
extern "C" JL_API unsigned long long jl_launch_count(void) { return jl::launches(); }

namespace jl {
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("JLLAMA_NO_PDL");
    return !(e && e[0] && e[0] != '0');
  }();
  return on;
}
}  // namespace jl
