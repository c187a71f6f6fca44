// Host-side helpers shared by all translation units of libvpg_b200.so: error reporting across the C ABI,
// launch counting, device-property cache.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/vpg_b200.h"

namespace vpg {

int set_error(int code, const char* fmt, ...);
void count_launch(int n = 1);
int sm_count();  // SMs of the current device (cached per device)

#define VPG_CHECK_ARG(cond, ...)                                   \
    do {                                                           \
        if (!(cond)) return ::vpg::set_error(VPG_ERR_INVALID, __VA_ARGS__); \
    } while (0)

#define VPG_CUDA(call)                                                                                   \
    do {                                                                                                 \
        cudaError_t _e = (call);                                                                         \
        if (_e != cudaSuccess)                                                                           \
            return ::vpg::set_error(VPG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), \
                                    __FILE__, __LINE__);                                                 \
    } while (0)

// call after a <<<>>> launch (cudaGetLastError also clears a non-sticky launch error so that it is reported once)
#define VPG_LAUNCH_CHECK()                  \
    do {                                    \
        ::vpg::count_launch();              \
        VPG_CUDA(cudaGetLastError());       \
    } while (0)

// tuning knobs: read from the environment ONCE (callers keep the value in a function-local static)
static inline int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

}  // namespace vpg
