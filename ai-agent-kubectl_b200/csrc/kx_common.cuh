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

inline void count_launch(int k = 1) { g_launches.fetch_add(k, std::memory_order_relaxed); }

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
