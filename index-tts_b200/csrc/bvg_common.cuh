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

// ---- bf16x3 split operands (BVG_F32_SPLIT_TC, the fp32 parity mode on the tensor cores) --------------------------------
// An fp32 value is the exact sum of three bf16 pieces v = h + m + l (8 + 8 + 8 significant bits; both subtractions below
// are exact in fp32).  A product x*w is then the sum of nine bf16 x bf16 products, each exact in the fp32 accumulator of
// the tensor core; the six largest (down to 2^-16 of the product) are kept, what is dropped is below 2^-24.  The six
// terms ride on the K axis of the ordinary bf16 implicit-GEMM kernel: activations are stored as
// [.., 6*C] = [l | m | m | h | h | h], weights as [.., 6*Cin] = [h | m | h | l | m | h] (pairs xl*wh, xm*wm, xm*wh, xh*wl,
// xh*wm, xh*wh).  SMALL TERMS FIRST: the tensor core's fp32 accumulation TRUNCATES (round toward zero) at every K = 16
// step, a bias of ~2^-24 of the running sum per step; the kernel walks K-blocks in slot order, so the five correction
// terms are summed while the accumulator is still 2^-8 of its final size and only the Cin*taps/16 steps of the leading
// xh*wh term truncate at full magnitude (measured: 6x smaller error than leading-term-first, profiles/r02_fp32_split.md).
#define BVG_SPLIT_TERMS 6
struct Bf16x3 {
    __nv_bfloat16 h, m, l;
};
__device__ __forceinline__ Bf16x3 split_bf16x3(float v) {
    Bf16x3 s;
    s.h = __float2bfloat16_rn(v);
    float r = v - __bfloat162float(s.h);
    s.m = __float2bfloat16_rn(r);
    r -= __bfloat162float(s.m);
    s.l = __float2bfloat16_rn(r);
    return s;
}
// piece of activation term slot t (0..5) / weight term slot t
__device__ __forceinline__ __nv_bfloat16 split_act_term(const Bf16x3& s, int t) { return t == 0 ? s.l : (t < 3 ? s.m : s.h); }
__device__ __forceinline__ __nv_bfloat16 split_w_term(const Bf16x3& s, int t) {
    return (t == 0 || t == 2 || t == 5) ? s.h : ((t == 1 || t == 4) ? s.m : s.l);
}

}  // namespace bvg
