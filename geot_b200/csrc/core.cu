// geot_b200/csrc/core.cu — error slot, launch counter, ABI version.
#include <atomic>
#include <cstdarg>
#include "common.cuh"

namespace geot {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

}  // namespace geot

GEOT_API int geot_abi_version(void) { return 1; }
GEOT_API const char* geot_last_error(void) { return geot::g_err; }
GEOT_API uint64_t geot_launch_count(void) { return geot::g_launches.load(std::memory_order_relaxed); }
