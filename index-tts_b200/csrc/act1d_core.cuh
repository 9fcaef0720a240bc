// Device-side core of Activation1d in the channels-last layout: packed-FP32 helpers, SnakeBeta, the per-thread sliding
// window march (polyphase up-sampler, snake, transposed-form low-pass).  Used by act1d.cu; kept separate so that a kernel
// that produces or consumes activated rows on chip can share it (the tail experiment of round 2).  Reference: alias_free_activation/torch/act.py:25-30,
// resample.py:33-58, activations.py:99-120.
#pragma once
#include "bvg_common.cuh"

namespace bvg {

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 splat(float a) { return make_float2(a, a); }

// Per-channel-pair snake constants.
struct SnakeCoef {
    float2 a;     // exp(alpha)               (precise)   | 2*exp(alpha)            (fast)
    float2 invb;  // 1/(exp(beta)+1e-9)       (precise)   | 0.5/(exp(beta)+1e-9)    (fast)
};

template <bool PRECISE>
__device__ __forceinline__ SnakeCoef make_coef(float al0, float al1, float be0, float be1) {
    SnakeCoef c;
    float a0 = expf(al0), a1 = expf(al1);
    float b0 = 1.0f / (expf(be0) + 1e-9f), b1 = 1.0f / (expf(be1) + 1e-9f);
    if (PRECISE) {
        c.a = make_float2(a0, a1);
        c.invb = make_float2(b0, b1);
    } else {
        c.a = make_float2(2.0f * a0, 2.0f * a1);
        c.invb = make_float2(0.5f * b0, 0.5f * b1);
    }
    return c;
}

// sin(t)^2 for two values at once, fp32-grade, every operation a packed FFMA2 / FMUL2 / FADD2 (13 of them): the libdevice
// sinf it replaces is ~20 scalar instructions per value, which made the precise mode 2.7-3.6x slower than the MUFU mode.
//   k = rint(t / pi) by the 1.5 * 2^23 trick, r = t - k * pi with a three-term Cody-Waite split of pi (k * PI_A is exact:
//   PI_A has 8 significant bits), |r| <= pi/2, sin(r) = r + r^3 * P(r^2) with the degree-11 Taylor polynomial (truncation
//   error 5.7e-8 at pi/2); the sign of sin(r) is irrelevant for the square.  |t| >= 2^21 (never reached by audio-scale
//   activations: the argument's own fp32 rounding is already +-0.25 rad there) falls back to libdevice.
__device__ __forceinline__ float2 sin_sq2(float2 t) {
    if (fmaxf(fabsf(t.x), fabsf(t.y)) >= 2097152.0f) {
        const float s0 = sinf(t.x), s1 = sinf(t.y);
        return make_float2(s0 * s0, s1 * s1);
    }
    const float2 magic = splat(12582912.0f);
    float2 k = ffma2(t, splat(0.318309886183790672f), magic);
    k = fadd2(k, splat(-12582912.0f));
    float2 r = ffma2(k, splat(-3.140625f), t);
    r = ffma2(k, splat(-9.67502593994140625e-4f), r);
    r = ffma2(k, splat(-1.509957990978376e-7f), r);
    const float2 z = fmul2(r, r);
    float2 q = ffma2(z, splat(-2.50521083854417188e-8f), splat(2.75573192239858907e-6f));
    q = ffma2(z, q, splat(-1.98412698412698413e-4f));
    q = ffma2(z, q, splat(8.33333333333333333e-3f));
    q = ffma2(z, q, splat(-1.66666666666666667e-1f));
    const float2 sn = ffma2(fmul2(r, z), q, r);
    return fmul2(sn, sn);
}

// PRECISE: v = u + invb * sin(a u)^2 with the range-reduced polynomial sine above.
// fast   : sin^2(t) = (1 - cos 2t)/2  ->  v = u - c*cos(2a u) + c with c = invb/2; the constant "+c" is NOT added
//          here: the down filter has unit DC gain and replicate padding, so it is added once per output instead.
template <bool PRECISE>
__device__ __forceinline__ float2 snake2(float2 u, const SnakeCoef& c) {
    if (PRECISE) {
        return ffma2(c.invb, sin_sq2(fmul2(u, c.a)), u);
    } else {
        float2 t = fmul2(u, c.a);
        float2 cs = make_float2(__cosf(t.x), __cosf(t.y));
        return ffma2(make_float2(-c.invb.x, -c.invb.y), cs, u);
    }
}

template <typename T>
struct PairIO;
template <>
struct PairIO<float> {
    static __device__ __forceinline__ float2 load(const float* p) { return __ldg(reinterpret_cast<const float2*>(p)); }
    static __device__ __forceinline__ void store(float* p, float2 v, int64_t) { *reinterpret_cast<float2*>(p) = v; }
};
template <>
struct PairIO<__nv_bfloat16> {
    static __device__ __forceinline__ float2 load(const __nv_bfloat16* p) {
        unsigned int raw = __ldg(reinterpret_cast<const unsigned int*>(p));
        return make_float2(__uint_as_float(raw << 16), __uint_as_float(raw & 0xffff0000u));
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, float2 v, int64_t) {
        *reinterpret_cast<__nv_bfloat162*>(p) = __float22bfloat162_rn(v);
    }
};
// fp32 rows of a shared-memory tile: generic loads / stores, rows need not be 8-byte aligned
struct TileF32 {
    float v;
};
template <>
struct PairIO<TileF32> {
    static __device__ __forceinline__ float2 load(const TileF32* p) {
        const float* q = reinterpret_cast<const float*>(p);
        return make_float2(q[0], q[1]);
    }
    static __device__ __forceinline__ void store(TileF32* p, float2 v, int64_t) {
        float* q = reinterpret_cast<float*>(p);
        q[0] = v.x;
        q[1] = v.y;
    }
};
// bf16x3 split output (fp32 parity mode on the tensor cores, bvg_common.cuh): an output row holds the six term slots
// [l | m | m | h | h | h] of C channels each, so a row is 12*C bytes and slot t of this channel pair is t * row_bytes / 6
// bytes further.  Output-only type.
struct Split6 {
    uint8_t bytes[12];
};
static_assert(sizeof(Split6) == 12, "Split6 must be 12 bytes: row pitch = C * sizeof(TOut)");
template <>
struct PairIO<Split6> {
    static __device__ __forceinline__ void store(Split6* p, float2 v, int64_t row_bytes) {
        const Bf16x3 a = split_bf16x3(v.x), b = split_bf16x3(v.y);
        const __nv_bfloat162 h = __halves2bfloat162(a.h, b.h), m = __halves2bfloat162(a.m, b.m),
                             l = __halves2bfloat162(a.l, b.l);
        const int64_t slot = row_bytes / BVG_SPLIT_TERMS;  // C * 2 bytes
        uint8_t* q = reinterpret_cast<uint8_t*>(p);
        *reinterpret_cast<__nv_bfloat162*>(q) = l;
        *reinterpret_cast<__nv_bfloat162*>(q + slot) = m;
        *reinterpret_cast<__nv_bfloat162*>(q + 2 * slot) = m;
        *reinterpret_cast<__nv_bfloat162*>(q + 3 * slot) = h;
        *reinterpret_cast<__nv_bfloat162*>(q + 4 * slot) = h;
        *reinterpret_cast<__nv_bfloat162*>(q + 5 * slot) = h;
    }
};
// byte offset of channel `ch` inside an output row
template <typename TOut>
__device__ __forceinline__ int64_t out_ch_bytes(int ch) { return static_cast<int64_t>(ch) * static_cast<int64_t>(sizeof(TOut)); }
template <>
__device__ __forceinline__ int64_t out_ch_bytes<Split6>(int ch) { return static_cast<int64_t>(ch) * 2; }

// ------------------------------------------------------------------------------------------------------------
// channels-last kernel
// ------------------------------------------------------------------------------------------------------------
// Thread (pair, seg) produces y[b, seg*L .. seg*L+L-1, 2*pair .. 2*pair+1].  L + 6 must be a multiple of 7.
// EDGE = false: the segment and its +-6-row halo lie strictly inside [0,T): no clamping, no padding logic.
// EDGE = true: first / last segments of a sequence (replicate padding of x and of the activated 2x signal).
struct ActEdge {
    int rel_lo, rel_hi;  // row 0 and row T-1 relative to row t0-6 (capped)
    int m_valid;         // steps g < m_valid have pair index m < T
    int n_out;           // outputs of this segment
    bool first_seg;      // t0 == 0
};

// One 7-step body: consumes cur[0..6] (rows rel 6+g0 .. 12+g0), produces outputs rel rows g0-6 .. g0.
// The low-pass runs in TRANSPOSED form: the pair (v[2m], v[2m+1]) produced at a step is scattered into the seven
// pending outputs y[m-3 .. m+3] (y[i] = sum_k g[k] v[2i+k-5]; v[2m] carries tap 2(m-i)+5, v[2m+1] tap 2(m-i)+6), so
// the twelve FFMA2 of a step are mutually independent (no 12-deep dependent chain) and only seven accumulators live
// in registers instead of a 14-entry history of the activated signal.  acc[(s+j+3)%7] holds y[m+j] at step s.
// NH > 1 ("heads"): the three AMP blocks of a stage apply their first Activation1d to the SAME tensor with different
// alpha / beta (reference bigvgan.py:369-374 calls resblocks[i*3+j](x) on one x).  The up-sampler does not depend on the
// snake parameters, so one pass computes it once and runs NH snake + low-pass pipelines on it: 12 + 16 NH packed FP32
// operations per row instead of 28 NH, and x is read once.  Head h stores at yp + ydelta[h] (ydelta[0] = 0).
template <typename TOut, bool PRECISE, bool EDGE, bool FIRST, int NH>
__device__ __forceinline__ void act1d_nlc_body(float2 (&xw)[7], float2 (&acc)[NH][7], const float2 (&cur)[7],
                                               float2 (&vlast)[NH], const SnakeCoef (&coef)[NH], const ActEdge& e, int g0,
                                               uint8_t*& yp, int64_t row_bytes, const int64_t (&ydelta)[NH]) {
    const float2 E0 = splat(2.f * BVG_F0), E1 = splat(2.f * BVG_F1), E2 = splat(2.f * BVG_F2),
                 E3 = splat(2.f * BVG_F3), E4 = splat(2.f * BVG_F4), E5 = splat(2.f * BVG_F5);
    const float2 G0 = splat(BVG_F0), G1 = splat(BVG_F1), G2 = splat(BVG_F2), G3 = splat(BVG_F3), G4 = splat(BVG_F4),
                 G5 = splat(BVG_F5);
#pragma unroll
    for (int s = 0; s < 7; ++s) {
        const int g = g0 + s;
        xw[(s + 6) % 7] = cur[s];
        const float2 xm3 = xw[(s + 0) % 7], xm2 = xw[(s + 1) % 7], xm1 = xw[(s + 2) % 7], x0 = xw[(s + 3) % 7],
                     xp1 = xw[(s + 4) % 7], xp2 = xw[(s + 5) % 7], xp3 = xw[(s + 6) % 7];
        float2 ue = fmul2(E0, xm3);
        float2 uo = fmul2(E1, xm2);
        ue = ffma2(E2, xm2, ue);
        uo = ffma2(E3, xm1, uo);
        ue = ffma2(E4, xm1, ue);
        uo = ffma2(E5, x0, uo);
        ue = ffma2(E5, x0, ue);
        uo = ffma2(E4, xp1, uo);
        ue = ffma2(E3, xp1, ue);
        uo = ffma2(E2, xp2, uo);
        ue = ffma2(E1, xp2, ue);
        uo = ffma2(E0, xp3, uo);
#pragma unroll
        for (int h = 0; h < NH; ++h) {
            // fast mode: the constant term of sin^2 = (1 - cos)/2 passes the unit-DC-gain low-pass unchanged, so it
            // seeds every accumulator instead of being added per sample
            const float2 hoisted_c = PRECISE ? splat(0.f) : coef[h].invb;
            float2 a = snake2<PRECISE>(ue, coef[h]);
            float2 o = snake2<PRECISE>(uo, coef[h]);
            if (EDGE) {
                if (g >= e.m_valid) {  // m >= T: replicate padding of the 2x signal on the right: v[2T-1]
                    a = vlast[h];
                    o = vlast[h];
                } else {
                    vlast[h] = o;
                }
            }
            float2& yp3 = acc[h][(s + 6) % 7];
            float2& yp2 = acc[h][(s + 5) % 7];
            float2& yp1 = acc[h][(s + 4) % 7];
            float2& y00 = acc[h][(s + 3) % 7];
            float2& ym1 = acc[h][(s + 2) % 7];
            float2& ym2 = acc[h][(s + 1) % 7];
            float2& ym3 = acc[h][(s + 0) % 7];
            if (EDGE && FIRST && s == 3 && e.first_seg) {
                // m == 0: replicate padding on the left, v[n < 0] = v[0]: the virtual pairs m = -3, -2, -1 are (a, a)
                const float2 g0a = ffma2(G0, a, hoisted_c);
                const float2 g12 = splat(BVG_F1 + BVG_F2), g34 = splat(BVG_F3 + BVG_F4);
                yp2 = g0a;                                     // y[2]: pair -1 tap 0
                yp1 = ffma2(g12, a, g0a);                      // y[1]: pair -2 tap 0, pair -1 taps 1,2
                y00 = ffma2(g34, a, ffma2(g12, a, g0a));       // y[0]: pair -3 tap 0, pair -2 taps 1,2, pair -1 taps 3,4
                // (whatever steps 0..2 scattered into these three from the clamped-x pairs is discarded here)
            }
            yp3 = ffma2(G0, o, hoisted_c);
            yp2 = ffma2(G1, a, yp2);
            yp1 = ffma2(G3, a, yp1);
            y00 = ffma2(G5, a, y00);
            ym1 = ffma2(G4, a, ym1);
            ym2 = ffma2(G2, a, ym2);
            ym3 = ffma2(G0, a, ym3);
            yp2 = ffma2(G2, o, yp2);
            yp1 = ffma2(G4, o, yp1);
            y00 = ffma2(G5, o, y00);
            ym1 = ffma2(G3, o, ym1);
            ym2 = ffma2(G1, o, ym2);
            if (!FIRST || s == 6) {
                const int i = g - 6;  // output row relative to t0
                if (!EDGE || i < e.n_out) PairIO<TOut>::store(reinterpret_cast<TOut*>(yp + ydelta[h]), ym3, row_bytes);
            }
        }
        yp += row_bytes;  // running byte pointer (row g - 6 of the segment): one 64-bit add per row, no multiply per store
    }
}

// Cy: row pitch of the output in TOut elements when it differs from the input's (shared-memory tiles);
// n_lim: store only the first n_lim outputs of the segment (EDGE instantiation only: a tile that ends inside the segment).
template <typename TIn, typename TOut, bool PRECISE, bool EDGE>
__device__ __forceinline__ void act1d_nlc_segment(const TIn* __restrict__ xb, TOut* __restrict__ yb,
                                                  const SnakeCoef& coef, int64_t T, int C, int L, int64_t t0, int Cy = -1,
                                                  int n_lim = -1) {
    if (Cy < 0) Cy = C;
    // rows are addressed relative to row (t0 - 6): rel = row - (t0 - 6)
    const TIn* xrel = xb + (t0 - 6) * C;
    TOut* yrel = yb + t0 * Cy;
    ActEdge e;
    e.rel_lo = EDGE ? (int)(6 - t0 > 0 ? 6 - t0 : 0) : 0;
    e.rel_hi = EDGE ? (int)min((int64_t)(L + 64), T - 1 - (t0 - 6)) : 0;
    e.m_valid = EDGE ? (int)min((int64_t)(L + 64), T - (t0 - 3)) : 0;
    e.n_out = (int)min((int64_t)(n_lim >= 0 && n_lim < L ? n_lim : L), T - t0);
    e.first_seg = (t0 == 0);
    auto ldx = [&](int rel) -> float2 {
        if (EDGE) rel = rel < e.rel_lo ? e.rel_lo : (rel > e.rel_hi ? e.rel_hi : rel);
        return PairIO<TIn>::load(xrel + (int64_t)rel * C);
    };
    // x window: x[m+j], j=-3..3, lives in xw[(s+j+3)%7] at step s of a 7-step body; pair m in ve/vo[s].
    // Loads run two 7-row bodies (~14 rows, > 1 us of work per warp) ahead of their use: global latency is hidden by
    // the register prefetch queue cur <- nxt <- nxt2 rather than by occupancy (2 CTAs/SM at ~110 registers).
    float2 xw[7], acc[1][7], cur[7], nxt[7], nxt2[7];
    float2 vlast[1] = {splat(0.f)};
    const SnakeCoef coefs[1] = {coef};
    const int64_t ydelta[1] = {0};
#pragma unroll
    for (int j = 0; j < 6; ++j) xw[j] = ldx(j);
#pragma unroll
    for (int s = 0; s < 7; ++s) {
        cur[s] = ldx(6 + s);
        acc[0][s] = splat(0.f);
    }
    const int nbody = (L + 6) / 7;
#pragma unroll
    for (int s = 0; s < 7; ++s) nxt[s] = nbody > 1 ? ldx(13 + s) : splat(0.f);
#pragma unroll
    for (int s = 0; s < 7; ++s) nxt2[s] = nbody > 2 ? ldx(20 + s) : splat(0.f);
    const int64_t row_bytes = (int64_t)Cy * (int64_t)sizeof(TOut);
    uint8_t* yp = reinterpret_cast<uint8_t*>(yrel) - 6 * row_bytes;  // row -6: never dereferenced before row 0
    act1d_nlc_body<TOut, PRECISE, EDGE, true, 1>(xw, acc, cur, vlast, coefs, e, 0, yp, row_bytes, ydelta);
#pragma unroll 1
    for (int body = 1; body < nbody; ++body) {
        const int g0 = 7 * body;
#pragma unroll
        for (int s = 0; s < 7; ++s) { cur[s] = nxt[s]; nxt[s] = nxt2[s]; }
        if (body + 2 < nbody) {
#pragma unroll
            for (int s = 0; s < 7; ++s) nxt2[s] = ldx(g0 + 20 + s);
        }
        act1d_nlc_body<TOut, PRECISE, EDGE, false, 1>(xw, acc, cur, vlast, coefs, e, g0, yp, row_bytes, ydelta);
    }
}


}  // namespace bvg
