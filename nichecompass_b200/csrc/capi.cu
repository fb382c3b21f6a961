// Library-level C ABI plumbing: error reporting, version, launch counter.
#include <atomic>
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void nc_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void nc_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

extern "C" const char* nc_last_error(void) { return g_err; }
extern "C" int nc_version(void) { return 100; }
extern "C" int64_t nc_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
