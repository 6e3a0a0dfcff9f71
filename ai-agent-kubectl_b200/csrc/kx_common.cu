#include "kx_common.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <string>
#include <map>
#include <mutex>
#include <utility>
#include <tuple>

namespace kx {
static thread_local char t_err[512] = "";
std::atomic<int64_t> g_launches{0};
thread_local TraceCtx *t_trace = nullptr;

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
}

bool debug_flag(const char *name) {
    const char *v = getenv("KX_DEBUG");              // read when an engine is created, not on the request path
    const std::string flags = std::string(",") + (v ? v : "") + ",";
    return flags.find(std::string(",") + name + ",") != std::string::npos;
}

namespace {
struct Scratch { void *ptr = nullptr; size_t cap = 0; };
std::mutex g_scratch_mu;
std::map<std::tuple<int, cudaStream_t, int>, Scratch> g_scratch;
}  // namespace

int stream_scratch(cudaStream_t stream, int tag, size_t bytes, void **out, bool zero_on_alloc) {
    int dev = 0;
    KX_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    Scratch &s = g_scratch[std::make_tuple(dev, stream, tag)];
    if (s.cap < bytes) {
        if (s.ptr) {
            KX_CUDA(cudaStreamSynchronize(stream));   // earlier work on this stream may still use it
            KX_CUDA(cudaFree(s.ptr));
            s.ptr = nullptr; s.cap = 0;
        }
        size_t cap = bytes + bytes / 4 + 256;
        KX_CUDA(cudaMalloc(&s.ptr, cap));
        s.cap = cap;
        if (zero_on_alloc) KX_CUDA(cudaMemsetAsync(s.ptr, 0, cap, stream));
    }
    *out = s.ptr;
    return KX_OK;
}

void release_scratch() {
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    int prev = 0;
    cudaGetDevice(&prev);
    for (auto &kv : g_scratch) {
        cudaSetDevice(std::get<0>(kv.first));
        cudaFree(kv.second.ptr);
    }
    g_scratch.clear();
    cudaSetDevice(prev);
}
}  // namespace kx

extern "C" const char *kx_last_error(void) { return kx::t_err; }
extern "C" int kx_version(void) { return 100; }
extern "C" void kx_release_scratch(void) { kx::release_scratch(); }
extern "C" int64_t kx_launch_count(void) { return kx::g_launches.load(); }
