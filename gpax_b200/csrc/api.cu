// C-ABI plumbing shared by all entry points: version, thread-local diagnostics.
#include <cstdarg>
#include <cstdio>
#include "common.cuh"

namespace gpx {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace gpx

extern "C" int gpx_abi_version(void) { return GPX_ABI_VERSION; }
extern "C" const char *gpx_last_error_string(void) { return gpx::g_err; }
