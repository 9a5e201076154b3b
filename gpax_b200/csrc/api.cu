// C-ABI plumbing shared by all entry points: version, thread-local diagnostics.
#include <cstdarg>
#include <cstdio>
#include <atomic>
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
}  // namespace gpx

extern "C" int gpx_abi_version(void) { return GPX_ABI_VERSION; }
extern "C" const char *gpx_last_error_string(void) { return gpx::g_err; }
extern "C" uint64_t gpx_launch_count(void) { return gpx::launches(); }
