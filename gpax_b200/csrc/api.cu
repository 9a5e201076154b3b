// C-ABI plumbing shared by all entry points: version, thread-local diagnostics.
#include <cstdarg>
#include <cstdio>
#include <atomic>
#include <mutex>
#include <map>
#include <utility>
#include "common.cuh"

namespace gpx {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static std::atomic<unsigned long long> g_launches{0};
void count_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
unsigned long long launches() { return g_launches.load(std::memory_order_relaxed); }

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: raise it once per
// (device, kernel) under a lock, so the entry points are thread-safe and work on every GPU of
// a process.  Returns 0 or the cudaError_t.
int ensure_dyn_smem(const void *func, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void *>, size_t> done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_pair(dev, func);
    auto it = done.find(key);
    if (it != done.end() && it->second >= bytes) return 0;
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) {
        set_error("cudaFuncSetAttribute(MaxDynamicSharedMemorySize = %zu) failed: %d (%s)", bytes, (int)e,
                  cudaGetErrorString(e));
        return (int)e;
    }
    done[key] = bytes;
    return 0;
}
}  // namespace gpx

extern "C" int gpx_abi_version(void) { return GPX_ABI_VERSION; }
extern "C" const char *gpx_last_error_string(void) { return gpx::g_err; }
extern "C" uint64_t gpx_launch_count(void) { return gpx::launches(); }
