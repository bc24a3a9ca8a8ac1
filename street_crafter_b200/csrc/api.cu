// Error reporting and launch accounting for libscb200.so (see include/scb200.h).
#include <atomic>
#include <cstdarg>
#include "common.cuh"

namespace {
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
}  // namespace

void sc_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void sc_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" const char* sc_last_error(void) { return g_err; }
extern "C" int sc_version(void) { return 100; }
extern "C" int64_t sc_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
