// Shared host-side helpers for libkx_b200.so (error reporting, launch counting).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/kx.h"

namespace kx {

void set_error(const char *fmt, ...);
extern std::atomic<int64_t> g_launches;

// Library-owned device scratch, cached per (current device, stream, purpose) and grown on demand
// (cudaMalloc only when it has to grow).  Valid until the next call on the same stream.
enum ScratchTag { SCRATCH_K9_QUEUE = 0, SCRATCH_GEMM = 1, SCRATCH_ATTN = 2, SCRATCH_TAGS = 3 };
int stream_scratch(cudaStream_t stream, int tag, size_t bytes, void **out, bool zero_on_alloc = false);
void release_scratch();

// Development switches, all in ONE environment variable read when an engine is created: KX_DEBUG is a comma-separated list of
//   trace            per-launch timeline of the decode step (kx_engine_trace)
//   no-fused-norm    decode keeps separate RMSNorm launches (A/B against the deferred-norm epilogues)
//   no-attn-tc       prefill attention on the mma.sync kernel
//   no-shared-prefix every sequence re-reads a cached prefix in decode attention
//   no-fused-argmax  lm_head writes logits, a second kernel picks the token
// Nothing in the product path depends on it being set.
bool debug_flag(const char *name);

inline void count_launch(int k = 1) { g_launches.fetch_add(k, std::memory_order_relaxed); }

// Function attributes (dynamic shared memory limit, cluster opt-in) are per device: a process that drives several
// GPUs (one engine per device behind a ReplicaRouter) has to set them on each.  `first()` is true once per
// (call site, current device); `done()` records it after the attributes were set.
struct PerDeviceOnce {
    std::atomic<uint64_t> mask{0};
    static int device() { int d = 0; cudaGetDevice(&d); return d & 63; }
    bool first() const { return !(mask.load(std::memory_order_acquire) & (1ull << device())); }
    void done() { mask.fetch_or(1ull << device(), std::memory_order_release); }
};

// ---- optional kernel timeline (diagnostic, KX_TRACE=1 at engine creation) -----------------------------
// A slot is eight u64 globaltimer values of one launch: [0] earliest CTA start, [1] earliest return from
// griddepcontrol.wait (the predecessor's results are visible), [2] latest "main loop done" mark, [3] latest
// CTA end, [4] latest "split-K slices exchanged" mark, [5..7] spare.  The host hands every instrumented launch its
// slot (or NULL when tracing is off).
struct TraceCtx {
    unsigned long long *buf = nullptr;     // device, cap slots of 8 u64
    int cap = 0, n = 0;
    const char *names[2048];
};
extern thread_local TraceCtx *t_trace;
inline unsigned long long *trace_slot(const char *name) {
    TraceCtx *t = t_trace;
    if (!t || t->n >= t->cap || t->n >= 2048) return nullptr;
    t->names[t->n] = name;
    return t->buf + 8 * (size_t)(t->n++);
}
#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long trace_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_min(unsigned long long *slot, int i) { if (slot) atomicMin(slot + i, trace_now()); }
__device__ __forceinline__ void trace_max(unsigned long long *slot, int i) { if (slot) atomicMax(slot + i, trace_now()); }
#endif

#define KX_CUDA(call)                                                                        \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            kx::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__,               \
                          cudaGetErrorString(e__));                                          \
            return KX_ECUDA;                                                                 \
        }                                                                                    \
    } while (0)

#define KX_REQUIRE(cond, code, ...)                                                          \
    do {                                                                                     \
        if (!(cond)) {                                                                       \
            kx::set_error(__VA_ARGS__);                                                      \
            return (code);                                                                   \
        }                                                                                    \
    } while (0)

}  // namespace kx
