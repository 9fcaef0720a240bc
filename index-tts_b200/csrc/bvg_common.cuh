// Shared host-side helpers of libbvg.so (error reporting, launch accounting) and device constants.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>

#include "../../include/bvg.h"

namespace bvg {

int set_error(int code, const char* fmt, ...);
void count_launch(int n = 1);

#define BVG_REQUIRE(cond, code, ...)                       \
    do {                                                   \
        if (!(cond)) return ::bvg::set_error((code), __VA_ARGS__); \
    } while (0)

#define BVG_CUDA(call)                                                                                     \
    do {                                                                                                   \
        cudaError_t e__ = (call);                                                                          \
        if (e__ != cudaSuccess)                                                                            \
            return ::bvg::set_error(BVG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                                    __FILE__, __LINE__);                                                   \
    } while (0)

// Launch check: cudaGetLastError only (no synchronisation).
#define BVG_LAUNCH_CHECK(name)                                                                                       \
    do {                                                                                                             \
        cudaError_t e__ = cudaGetLastError();                                                                        \
        if (e__ != cudaSuccess)                                                                                      \
            return ::bvg::set_error(BVG_ERR_CUDA, "launch of %s failed: %s", (name), cudaGetErrorString(e__));       \
        ::bvg::count_launch();                                                                                       \
    } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

// 12-tap Kaiser-windowed sinc, cutoff 0.25, half-width 0.3: the fp32 values the reference registers as
// `upsample.filter` / `downsample.lowpass.filter` (alias_free_activation/torch/filter.py:30-62); symmetric, so six
// distinct taps: F0 = f[0] = f[11], ..., F5 = f[5] = f[6].  Exact bit patterns pinned by tests/golden/filter.npz.
#define BVG_F0 2.0289646927e-03f
#define BVG_F1 9.3894656748e-03f
#define BVG_F2 -2.5543458760e-02f
#define BVG_F3 -5.7657383382e-02f
#define BVG_F4 1.2857258320e-01f
#define BVG_F5 4.4320979714e-01f

}  // namespace bvg
