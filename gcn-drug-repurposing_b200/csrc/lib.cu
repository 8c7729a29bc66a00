// lib.cu — library-level entry points: version, error string, launch counter.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"

namespace gdr {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace gdr

extern "C" {

int gdr_version(void) { return 100; }  // 0.1.0

const char *gdr_last_error_string(void) { return gdr::g_err; }

int64_t gdr_launch_count(void) { return gdr::g_launches.load(); }

void gdr_launch_count_reset(void) { gdr::g_launches.store(0); }

}  // extern "C"
