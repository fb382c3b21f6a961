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
// launches replayed from a captured CUDA graph (counted by the caller at capture time)
extern "C" int64_t nc_launch_count_add(int64_t n) { return g_launches.fetch_add(n, std::memory_order_relaxed) + n; }

// ---------------------------------------------------------------------------- pinned host blocks
// Host-resident count store (data/store.py, location="host"): page-locked, device-mapped memory owned by the
// caller for the lifetime of the store.  Blocks come straight from cudaHostAlloc and go back with cudaFreeHost
// (which waits for the device), never through a caching allocator: a block that a zero-copy kernel may still be
// reading -- or whose lines a previous tenant left in a GPU cache -- is never handed to a new tenant.
extern "C" int nc_host_alloc(int64_t bytes, void** out_host) {
    NC_REQUIRE(bytes >= 0 && out_host != nullptr, NC_EINVAL, "nc_host_alloc: bad arguments");
    *out_host = nullptr;
    if (bytes == 0) return NC_OK;
    cudaError_t e = cudaHostAlloc(out_host, (size_t)bytes, cudaHostAllocPortable | cudaHostAllocMapped);
    if (e != cudaSuccess) {
        cudaGetLastError();
        nc_set_error("nc_host_alloc: cudaHostAlloc(%lld B) failed: %s", (long long)bytes, cudaGetErrorString(e));
        return NC_ECUDA;
    }
    return NC_OK;
}

extern "C" int nc_host_free(void* host) {
    if (host == nullptr) return NC_OK;
    cudaError_t e = cudaFreeHost(host);             // implicit device synchronisation: no kernel still reads it
    if (e != cudaSuccess) {
        cudaGetLastError();
        nc_set_error("nc_host_free: cudaFreeHost failed: %s", cudaGetErrorString(e));
        return NC_ECUDA;
    }
    return NC_OK;
}

// Raises a kernel's dynamic shared-memory limit once per (kernel, device); the result of
// cudaFuncSetAttribute is checked (a failure used to surface only as "invalid argument" at launch).
int nc_set_max_smem(const void* func, int bytes, bool* done_per_device /* [64] */, const char* name) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess && dev >= 0 && dev < 64 && done_per_device[dev]) return NC_OK;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        nc_set_error("%s: cudaFuncSetAttribute(max dynamic smem = %d) failed: %s", name, bytes, cudaGetErrorString(e));
        return NC_ECUDA;
    }
    if (dev >= 0 && dev < 64) done_per_device[dev] = true;
    return NC_OK;
}
