#include "kx_common.cuh"
#include <stdarg.h>

namespace kx {
static thread_local char t_err[512] = "";
std::atomic<int64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
}
}  // namespace kx

extern "C" const char *kx_last_error(void) { return kx::t_err; }
extern "C" int kx_version(void) { return 100; }
extern "C" int64_t kx_launch_count(void) { return kx::g_launches.load(); }
