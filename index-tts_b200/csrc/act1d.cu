// Fused anti-aliased SnakeBeta activation (Activation1d): y = lowpass_down2( snake_beta( up2(x) ) ).
//
// Replaces the reference's fused extension kernel (alias_free_activation/cuda/anti_alias_activation_cuda.cu:43-179)
// and its torch path (alias_free_activation/torch/act.py:25-30, resample.py:30-38, filter.py:94-101,
// activations.py:107-120).  Written from the closed forms, cl() = clamp to [0,T-1]:
//     u[2m]   = sum_j E_even[j] x[cl(m+j)]   j=-3..2     (E = 2*f, polyphase up-sampler, gain 2 folded in)
//     u[2m+1] = sum_j E_odd[j]  x[cl(m+j)]   j=-2..3
//     v[n]    = u[n] + 1/(exp(beta)+1e-9) * sin^2(exp(alpha) * u[n])
//     y[i]    = sum_{k<12} f[k] v[clamp(2i+k-5, 0, 2T-1)]
// No 2x-rate tensor is ever written: the 2x signal lives in registers.
//
// Two kernels:
//  * act1d_nlc_kernel  -- internal channels-last layout [B,T,C].  One thread owns a channel PAIR and marches along
//    time with register sliding windows, so every FIR/snake op is a packed f32x2 instruction (FFMA2) over the two
//    channels and each input element is loaded exactly once (plus a 6-row warm-up per time segment).  Loads and
//    stores are coalesced across the channel axis.
//  * act1d_ncl_kernel  -- public [B,C,T] layout.  A CTA stages a halo'd time tile of one (b,c) row in shared
//    memory with one bulk asynchronous copy (TMA engine, cp.async.bulk + mbarrier), each thread computes 24
//    consecutive outputs from registers and writes them with 128-bit stores.
#include "bvg_common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

#include <atomic>
#include <cstdlib>

#ifndef BVG_ACT_SMEM_KB_DEFAULT
#define BVG_ACT_SMEM_KB_DEFAULT 100
#endif

#include "act1d_core.cuh"

namespace bvg {

// act1d_umma.cu: both FIRs on the tensor cores (fast-sin mode only); positive return = shape not taken
int activation1d_nlc_umma(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha, const float* beta, int64_t B,
                          int64_t T, int64_t C, cudaStream_t st);
bool act_umma_requested(int64_t C);
int activation1d_umma_init();

// grid = (ceil(n_seg / segs_per_block), C/2 / S, B); block = S * segs_per_block threads.
template <typename TIn, typename TOut, bool PRECISE>
__global__ void __launch_bounds__(256)
act1d_nlc_kernel(const TIn* __restrict__ x, TOut* __restrict__ y, const float* __restrict__ alpha,
                 const float* __restrict__ beta, int64_t T, int C, int L, int S, int64_t n_seg) {
    const int pair = blockIdx.y * S + (threadIdx.x % S);
    const int64_t seg = (int64_t)blockIdx.x * (blockDim.x / S) + (threadIdx.x / S);
    if (seg >= n_seg) return;
    const int64_t b = blockIdx.z;
    const int64_t t0 = seg * L;
    const TIn* xb = x + (b * T) * C + 2 * pair;
    TOut* yb = reinterpret_cast<TOut*>(reinterpret_cast<uint8_t*>(y + (b * T) * C) + out_ch_bytes<TOut>(2 * pair));
    const SnakeCoef coef = make_coef<PRECISE>(alpha[2 * pair], alpha[2 * pair + 1], beta[2 * pair], beta[2 * pair + 1]);
    // the segment's loads touch rows t0-6 .. t0+L+5 (the last 7-row prefetch is skipped)
    const bool interior = (t0 >= 6) && (t0 + L + 6 <= T);
    if (interior)
        act1d_nlc_segment<TIn, TOut, PRECISE, false>(xb, yb, coef, T, C, L, t0);
    else
        act1d_nlc_segment<TIn, TOut, PRECISE, true>(xb, yb, coef, T, C, L, t0);
}

// ------------------------------------------------------------------------------------------------------------
// channels-last kernel, TMA-pipelined (the one bvg_forward uses)
// ------------------------------------------------------------------------------------------------------------
// A work unit is (batch item, chunk of CC channels, group of `nsegs` consecutive time segments of L rows).  The 256
// compute threads of a CTA map to (channel pair, segment); each marches along its segment with the same register
// sliding windows as above, but its input rows arrive through a ring of shared-memory "slabs": slab q of a unit holds
// rows t0_s - 7 + 7q .. +6 of EVERY segment s of the group (one TMA box [CC x 7 rows] per segment, all completing on
// one mbarrier).  A producer warp keeps the ring full, so the compute warps never wait on global memory: they read 7
// rows into registers, hand the slab straight back, and spend the rest of the body on FFMA2/MUFU work.  TMA
// zero-fills rows outside [0,T); the sequence ends (replicate padding of x and of the activated 2x signal) are fixed
// up in registers on the EDGE path.
constexpr int ACT_TMA_COMPUTE = 256;
constexpr int ACT_TMA_THREADS = ACT_TMA_COMPUTE + 32;
constexpr int ACT_TMA_MAX_STAGES = 8;
#ifndef ACT_TMA_MIN_CTAS
#define ACT_TMA_MIN_CTAS 2
#endif
#ifndef ACT_TMA_PREFETCH
#define ACT_TMA_PREFETCH 0  // register prefetch of the next slab: measured neutral (gpurun r01s), costs 14 MOV per body
#endif

struct ActTmaParams {
    int64_t T;
    int C, CC, P, nsegs, L, nbody, n_cchunks;
    uint32_t groups_per_b, n_units;
    int stages;
    int rb;                // 7-row blocks per slab: one TMA box / mbarrier hand-shake feeds rb bodies
    uint32_t seg_stride;   // bytes between the segments of one slab (multiple of 128)
    uint32_t row_bytes;    // CC * sizeof(TIn)
    uint32_t stage_bytes;  // nsegs * seg_stride
    uint32_t tx_bytes;     // nsegs * 7 * row_bytes
};

// Slab reads use explicit shared-space loads on a 32-bit address (the aligned dynamic-smem pointer is otherwise demoted
// to generic LD by the compiler).
template <typename T>
struct SmemPair;
template <>
struct SmemPair<float> {
    static __device__ __forceinline__ float2 load(uint32_t a) {
        float2 v;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
        return v;
    }
};
template <>
struct SmemPair<__nv_bfloat16> {
    static __device__ __forceinline__ float2 load(uint32_t a) {
        unsigned int raw;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(raw) : "r"(a));
        return make_float2(__uint_as_float(raw << 16), __uint_as_float(raw & 0xffff0000u));
    }
};

// Outputs and snake parameters of the NH heads of one launch (NH = 1: the plain Activation1d).
template <int NH>
struct ActHeads {
    void* y[NH];
    const float* alpha[NH];
    const float* beta[NH];
};
// CTAs per SM: the three-head body keeps 21 accumulators live (~150 registers), one CTA per SM
constexpr int act_tma_min_ctas(int nh) { return nh == 1 ? ACT_TMA_MIN_CTAS : 1; }

template <typename TIn, typename TOut, bool PRECISE, int NH>
__global__ void __launch_bounds__(ACT_TMA_THREADS, act_tma_min_ctas(NH))
act1d_nlc_tma_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ ActHeads<NH> hd,
                     const __grid_constant__ ActTmaParams p) {
    extern __shared__ uint8_t act_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(act_smem_raw) + 127) & ~uintptr_t(127));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.stages * p.stage_bytes);
    uint64_t* empty = full + ACT_TMA_MAX_STAGES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    griddep_launch_dependents();  // PDL: see sm100_ptx.cuh; only the producer warp touches produced data before its wait
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_x);
        for (int i = 0; i < p.stages; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], ACT_TMA_COMPUTE / 32);
        }
        fence_barrier_init();
    }
    __syncthreads();
    const int n_slabs = (p.nbody + 1 + p.rb - 1) / p.rb;  // blocks: 1 window block + nbody bodies, rb per slab
    const uint32_t smem_base = smem_u32(smem);
    int st = 0;
    uint32_t ph = 0;

    if (warp == ACT_TMA_COMPUTE / 32) {
        // ------------------------------ TMA producer ------------------------------
        griddep_wait();  // x is produced by the previous kernel of the stream
        for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
            const uint32_t cch = unit % p.n_cchunks;
            const uint32_t rest = unit / p.n_cchunks;
            const uint32_t b = rest / p.groups_per_b;
            const uint32_t g = rest - b * p.groups_per_b;
            const int row_base = static_cast<int>(g) * p.nsegs * p.L - 7;
            for (int q = 0; q < n_slabs; ++q) {
                mbar_wait(&empty[st], ph ^ 1);
                if (lane == 0) mbar_arrive_expect_tx(&full[st], p.tx_bytes);
                __syncwarp();
                uint8_t* dst = smem + st * p.stage_bytes;
                for (int s = lane; s < p.nsegs; s += 32)
                    tma_load_3d(dst + s * p.seg_stride, &map_x, &full[st], static_cast<int>(cch) * p.CC,
                                row_base + s * p.L + 7 * p.rb * q, static_cast<int>(b));
                if (++st == p.stages) { st = 0; ph ^= 1; }
            }
        }
        return;
    }

    // ------------------------------ compute warps ------------------------------
    const int tid = threadIdx.x;
    const int seg_raw = tid / p.P;
    const bool lane_used = seg_raw < p.nsegs;
    const int seg = lane_used ? seg_raw : 0;
    const int pair = tid - seg_raw * p.P;
    const uint32_t my_off = seg * p.seg_stride + pair * 2u * static_cast<uint32_t>(sizeof(TIn));
    const int L = p.L, C = p.C;
    const int64_t T = p.T;

    for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
        const uint32_t cch = unit % p.n_cchunks;
        const uint32_t rest = unit / p.n_cchunks;
        const uint32_t b = rest / p.groups_per_b;
        const uint32_t g = rest - b * p.groups_per_b;
        const int64_t t0 = (static_cast<int64_t>(g) * p.nsegs + seg) * L;
        const bool active = lane_used && t0 < T;
        const bool interior = (t0 > 0) && (t0 + L + 6 <= T);
        const bool edge_path = __any_sync(0xffffffffu, active && !interior);
        const int ch = static_cast<int>(cch) * p.CC + 2 * pair;
        SnakeCoef coef[NH];
        int64_t ydelta[NH];
#pragma unroll
        for (int h = 0; h < NH; ++h) {
            coef[h] = make_coef<PRECISE>(hd.alpha[h][ch], hd.alpha[h][ch + 1], hd.beta[h][ch], hd.beta[h][ch + 1]);
            ydelta[h] = static_cast<const uint8_t*>(hd.y[h]) - static_cast<const uint8_t*>(hd.y[0]);
        }
        TOut* yrel = reinterpret_cast<TOut*>(
            reinterpret_cast<uint8_t*>(static_cast<TOut*>(hd.y[0]) + (static_cast<int64_t>(b) * T + t0) * C) +
            out_ch_bytes<TOut>(ch));
        const int64_t row_bytes = static_cast<int64_t>(C) * static_cast<int64_t>(sizeof(TOut));
        uint8_t* yp = reinterpret_cast<uint8_t*>(yrel) - 6 * row_bytes;  // row -6: never dereferenced before row 0

        ActEdge e;
        e.rel_lo = e.rel_hi = 0;
        e.m_valid = static_cast<int>(min(static_cast<int64_t>(L + 64), T - (t0 - 3)));
        e.n_out = static_cast<int>(min(static_cast<int64_t>(L), T - t0));
        e.first_seg = (t0 == 0);

        float2 xw[7], acc[NH][7], cur[7];
#if ACT_TMA_PREFETCH
        float2 nxt[7];
#endif
        float2 vlast[NH], xlast = splat(0.f);
#pragma unroll
        for (int h = 0; h < NH; ++h) vlast[h] = splat(0.f);
        // The unit's rows arrive as 7-row blocks: block 0 = rows t0-7 .. t0-1 (the window starts with rows t0-6 .. t0-1),
        // block b+1 = the rows of body b.  A slab holds rb consecutive blocks of every segment; it is waited for at its
        // first block and handed back to the producer after its last one.
        const int last_blk = p.nbody;
        int blk = 0, sub = 0;
        auto release = [&]() {
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[st]);
            if (++st == p.stages) { st = 0; ph ^= 1; }
            sub = 0;
        };
        mbar_wait(&full[st], ph);
        {
            const uint32_t sp = smem_base + st * p.stage_bytes + my_off;
#pragma unroll
            for (int j = 0; j < 6; ++j) xw[j] = SmemPair<TIn>::load(sp + (j + 1) * p.row_bytes);
        }
        if (p.rb == 1 || last_blk == 0) release(); else sub = 1;
        blk = 1;
#pragma unroll
        for (int h = 0; h < NH; ++h)
#pragma unroll
            for (int s = 0; s < 7; ++s) acc[h][s] = splat(0.f);
        auto fetch = [&](float2 (&dst)[7]) {
            if (sub == 0) mbar_wait(&full[st], ph);
            const uint32_t sp = smem_base + st * p.stage_bytes + my_off + sub * 7u * p.row_bytes;
#pragma unroll
            for (int s = 0; s < 7; ++s) dst[s] = SmemPair<TIn>::load(sp + s * p.row_bytes);
            if (sub + 1 == p.rb || blk == last_blk) release(); else ++sub;
            ++blk;
        };
#if ACT_TMA_PREFETCH
        fetch(nxt);
#endif

#pragma unroll 1
        for (int body = 0; body < p.nbody; ++body) {
#if ACT_TMA_PREFETCH
#pragma unroll
            for (int s = 0; s < 7; ++s) cur[s] = nxt[s];
            if (body + 1 < p.nbody) fetch(nxt);
#else
            fetch(cur);
#endif
            if (!active) continue;
            const int g0 = 7 * body;
            if (edge_path) {
                // replicate padding of x: rows >= T repeat x[T-1]; rows < 0 (first segment only) repeat x[0]
#pragma unroll
                for (int s = 0; s < 7; ++s) {
                    if (t0 + g0 + s >= T) cur[s] = xlast; else xlast = cur[s];
                }
                if (body == 0) {
                    if (e.first_seg) {
#pragma unroll
                        for (int j = 0; j < 6; ++j) xw[j] = cur[0];
                    }
                    act1d_nlc_body<TOut, PRECISE, true, true, NH>(xw, acc, cur, vlast, coef, e, 0, yp, row_bytes, ydelta);
                } else {
                    act1d_nlc_body<TOut, PRECISE, true, false, NH>(xw, acc, cur, vlast, coef, e, g0, yp, row_bytes, ydelta);
                }
            } else {
                if (body == 0)
                    act1d_nlc_body<TOut, PRECISE, false, true, NH>(xw, acc, cur, vlast, coef, e, 0, yp, row_bytes, ydelta);
                else
                    act1d_nlc_body<TOut, PRECISE, false, false, NH>(xw, acc, cur, vlast, coef, e, g0, yp, row_bytes, ydelta);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// [B,C,T] kernel with a TMA-staged halo'd time tile
// ------------------------------------------------------------------------------------------------------------
constexpr int NCL_THREADS = 128;
// 24 outputs per thread: 25 % halo work in the up-sampler instead of 37.5 % at 16, and a 96-byte lane stride makes the
// 128-bit shared-memory reads 2-way instead of 4-way bank conflicted; 32 (128-byte stride: 8-way conflicts, 130 registers)
// is 35 % slower.  cfg3, C = 1536, T = 2^21, fp32: 16 -> 0.75, 24 -> 0.81, 32 -> 0.51 of the HBM peak
// (profiles/r02_ncl_outputs_per_thread.log).
#ifndef BVG_NCL_PT
#define BVG_NCL_PT 24
#endif
constexpr int NCL_PER_THREAD = BVG_NCL_PT;              // outputs per thread (multiple of 8)
constexpr int NCL_XR = NCL_PER_THREAD + 16;             // per-thread input window x[i0-8 .. i0+PT+8)
constexpr int NCL_TILE = NCL_THREADS * NCL_PER_THREAD;  // outputs per CTA
constexpr int NCL_HALO = 8;                             // >= 6; multiple of 8 keeps interior tiles 16-byte aligned
constexpr int NCL_SMEM_ELEMS = NCL_TILE + 2 * NCL_HALO + 16;

__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

template <typename TIO>
struct NclIO;
template <>
struct NclIO<float> {
    // xr[j] = tile[base + j], j < 32, base % 4 == 0
    static __device__ __forceinline__ void load32(const float* tile, int base, float (&xr)[NCL_XR]) {
        const float4* p = reinterpret_cast<const float4*>(tile + base);
#pragma unroll
        for (int j = 0; j < NCL_XR / 4; ++j) {
            float4 v = p[j];
            xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
        }
    }
    static __device__ __forceinline__ float get(const float* tile, int i) { return tile[i]; }
    static __device__ __forceinline__ void store16(float* y, const float (&o)[NCL_PER_THREAD]) {
        float4* p = reinterpret_cast<float4*>(y);
#pragma unroll
        for (int j = 0; j < NCL_PER_THREAD / 4; ++j) p[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
    }
    static __device__ __forceinline__ void store1(float* y, float v) { *y = v; }
};
template <>
struct NclIO<__nv_bfloat16> {
    static __device__ __forceinline__ void load32(const __nv_bfloat16* tile, int base, float (&xr)[NCL_XR]) {
        const uint4* p = reinterpret_cast<const uint4*>(tile + base);  // base % 8 == 0
#pragma unroll
        for (int j = 0; j < NCL_XR / 8; ++j) {
            uint4 v = p[j];
            const unsigned int w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                xr[8 * j + 2 * q] = __uint_as_float(w[q] << 16);
                xr[8 * j + 2 * q + 1] = __uint_as_float(w[q] & 0xffff0000u);
            }
        }
    }
    static __device__ __forceinline__ float get(const __nv_bfloat16* tile, int i) { return __bfloat162float(tile[i]); }
    static __device__ __forceinline__ void store16(__nv_bfloat16* y, const float (&o)[NCL_PER_THREAD]) {
        uint4* p = reinterpret_cast<uint4*>(y);
#pragma unroll
        for (int j = 0; j < NCL_PER_THREAD / 8; ++j) {
            __nv_bfloat162 h0 = __floats2bfloat162_rn(o[8 * j], o[8 * j + 1]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(o[8 * j + 2], o[8 * j + 3]);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(o[8 * j + 4], o[8 * j + 5]);
            __nv_bfloat162 h3 = __floats2bfloat162_rn(o[8 * j + 6], o[8 * j + 7]);
            p[j] = make_uint4(*reinterpret_cast<unsigned int*>(&h0), *reinterpret_cast<unsigned int*>(&h1),
                              *reinterpret_cast<unsigned int*>(&h2), *reinterpret_cast<unsigned int*>(&h3));
        }
    }
    static __device__ __forceinline__ void store1(__nv_bfloat16* y, float v) { *y = __float2bfloat16_rn(v); }
};

// xr[j] = tile[base4 + D + j], j < 32, base4 % 4 == 0, 0 < D < 4 (rows that start D floats into a 16-byte vector)
template <int D>
__device__ __forceinline__ void ncl_load32_shifted(const float* tile, int base4, float (&xr)[NCL_XR]) {
    const float4* p = reinterpret_cast<const float4*>(tile + base4);
    float buf[NCL_XR + 4];
#pragma unroll
    for (int j = 0; j < NCL_XR / 4 + 1; ++j) {
        const float4 v = p[j];
        buf[4 * j] = v.x; buf[4 * j + 1] = v.y; buf[4 * j + 2] = v.z; buf[4 * j + 3] = v.w;
    }
#pragma unroll
    for (int j = 0; j < NCL_XR; ++j) xr[j] = buf[j + D];
}
// 16 fp32 outputs to a row address that is H floats short of a 16-byte boundary (0 < H < 4): H scalars, three aligned
// vectors, 4 - H scalars.
template <int H>
__device__ __forceinline__ void ncl_store16_shifted(float* y, const float (&o)[NCL_PER_THREAD]) {
#pragma unroll
    for (int j = 0; j < H; ++j) y[j] = o[j];
#pragma unroll
    for (int v = 0; v < NCL_PER_THREAD / 4 - 1; ++v)
        *reinterpret_cast<float4*>(y + H + 4 * v) = make_float4(o[H + 4 * v], o[H + 4 * v + 1], o[H + 4 * v + 2], o[H + 4 * v + 3]);
#pragma unroll
    for (int j = H + NCL_PER_THREAD - 4; j < NCL_PER_THREAD; ++j) y[j] = o[j];
}

// out[q] = y[i0 + q], q < 16, from xr[j] = x[cl(i0 - 8 + j)].  EDGE: the thread touches an end of the row (replicate
// padding of the activated 2x signal).
template <bool PRECISE, bool EDGE>
__device__ __forceinline__ void ncl_compute16(const float (&xr)[NCL_XR], const SnakeCoef& coef, int64_t i0, int64_t T,
                                              float (&out)[NCL_PER_THREAD]) {
    const float e0 = 2.f * BVG_F0, e1 = 2.f * BVG_F1, e2 = 2.f * BVG_F2, e3 = 2.f * BVG_F3, e4 = 2.f * BVG_F4,
                e5 = 2.f * BVG_F5;
    // The packed up-sampler reads (x[a], x[a+1]) for every a: an FFMA2 operand is an even-aligned 64-bit register pair,
    // so the odd-start pairs are materialised ONCE (30 MOVs) instead of being re-assembled at each of their six uses
    // (ncu r02y: ~250 MOVs of 880 instructions per thread, issue slots 81 % busy against 67 % for the FMA pipe).
    float2 xe[NCL_XR / 2], xo[NCL_XR / 2 - 1];
#pragma unroll
    for (int i = 0; i < NCL_XR / 2; ++i) xe[i] = make_float2(xr[2 * i], xr[2 * i + 1]);
#pragma unroll
    for (int i = 0; i < NCL_XR / 2 - 1; ++i) xo[i] = make_float2(xr[2 * i + 1], xr[2 * i + 2]);
#define NCL_X2(a) (((a) & 1) ? xo[(a) >> 1] : xe[(a) >> 1])
    // v2[p] = (v[2m], v[2m+1]) for m = i0 - 3 + p
    float2 v2[NCL_PER_THREAD + 6];
    float vlast = 0.f;
    // pairs with m >= T repeat v[2T-1]: in tile-relative 32-bit arithmetic (the row end is at most 2^31 away or irrelevant)
    const int p_end = EDGE ? static_cast<int>(min(T - (i0 - 3), (int64_t)(NCL_PER_THREAD + 6))) : NCL_PER_THREAD + 6;
#pragma unroll
    for (int p = 0; p < NCL_PER_THREAD + 6; ++p) {
        // x[m+j] = xr[p + 5 + j]
        float2 u = fmul2(make_float2(e0, e1), NCL_X2(p + 2));
        u = ffma2(make_float2(e2, e3), NCL_X2(p + 3), u);
        u = ffma2(make_float2(e4, e5), NCL_X2(p + 4), u);
        u = ffma2(make_float2(e5, e4), NCL_X2(p + 5), u);
        u = ffma2(make_float2(e3, e2), NCL_X2(p + 6), u);
        u = ffma2(make_float2(e1, e0), NCL_X2(p + 7), u);
        float2 v = snake2<PRECISE>(u, coef);
        if (EDGE) {
            if (p >= p_end) {
                v = splat(vlast);
            } else {
                vlast = v.y;
            }
        }
        v2[p] = v;
    }
#undef NCL_X2
    if (EDGE && i0 == 0) {  // m = -3,-2,-1 -> v[0]
        v2[0] = v2[1] = v2[2] = splat(v2[3].x);
    }
    const float hoisted_c = PRECISE ? 0.f : coef.invb.x;
#pragma unroll
    for (int q = 0; q < NCL_PER_THREAD; ++q) {
        float2 acc = fmul2(make_float2(BVG_F1, BVG_F2), v2[q + 1]);
        acc = ffma2(make_float2(BVG_F3, BVG_F4), v2[q + 2], acc);
        acc = ffma2(make_float2(BVG_F5, BVG_F5), v2[q + 3], acc);
        acc = ffma2(make_float2(BVG_F4, BVG_F3), v2[q + 4], acc);
        acc = ffma2(make_float2(BVG_F2, BVG_F1), v2[q + 5], acc);
        float r = fmaf(BVG_F0, v2[q].y, acc.x);
        r = fmaf(BVG_F0, v2[q + 6].x, r);
        out[q] = r + acc.y + hoisted_c;
    }
}

// grid = (ceil(T / NCL_TILE), C, B)
template <typename TIO, bool PRECISE>
__global__ void __launch_bounds__(NCL_THREADS)
act1d_ncl_kernel(const TIO* __restrict__ x, TIO* __restrict__ y, const float* __restrict__ alpha,
                 const float* __restrict__ beta, int64_t C, int64_t T, int64_t stride_b, int64_t stride_c,
                 int vec_store_ok) {
    // also reused as the fp32 output staging tile (NCL_TILE + NCL_TILE/32 words) on the unaligned-row path
    constexpr int kTileElems = (NCL_TILE + NCL_TILE / 32) * 4 / (int)sizeof(TIO) > NCL_SMEM_ELEMS
                                   ? (NCL_TILE + NCL_TILE / 32) * 4 / (int)sizeof(TIO) : NCL_SMEM_ELEMS;
    __shared__ __align__(128) TIO tile[kTileElems];
    __shared__ __align__(8) uint64_t bar;
    const int64_t c = blockIdx.y, b = blockIdx.z;
    const int64_t t0 = (int64_t)blockIdx.x * NCL_TILE;
    const TIO* row = x + b * stride_b + c * stride_c;
    const int64_t lo = max(t0 - (int64_t)NCL_HALO, (int64_t)0);
    const int64_t hi = min(t0 + (int64_t)(NCL_TILE + NCL_HALO), T);
    const uintptr_t a_lo = reinterpret_cast<uintptr_t>(row + lo), a_hi = reinterpret_cast<uintptr_t>(row + hi);
    const uintptr_t c_lo = a_lo & ~uintptr_t(15), c_hi = (a_hi + 15) & ~uintptr_t(15);
    const int delta = static_cast<int>((a_lo - c_lo) / sizeof(TIO));

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t bytes = static_cast<uint32_t>(c_hi - c_lo);
        mbar_arrive_expect_tx(&bar, bytes);
        bulk_copy_g2s(tile, reinterpret_cast<const void*>(c_lo), bytes, &bar);
    }
    const SnakeCoef coef = make_coef<PRECISE>(alpha[c], alpha[c], beta[c], beta[c]);
    mbar_wait(&bar, 0);

    const int64_t i0 = t0 + (int64_t)threadIdx.x * NCL_PER_THREAD;
    const bool active = i0 < T;  // no early return: every thread reaches the staging barrier below

    // xr[j] = x[cl(i0 - 8 + j)]
    float xr[NCL_XR];
    if (active) {
        if (i0 - 8 >= lo && i0 + NCL_PER_THREAD + 8 <= hi) {  // interior thread: no clamping
            const int base = delta + static_cast<int>(i0 - 8 - lo);
            if (delta == 0) {
                NclIO<TIO>::load32(tile, base, xr);
            } else if (sizeof(TIO) == 4) {
                // fp32 row that is not 16-byte aligned (odd T): the tile was copied from the aligned address below the row
                // start, so this thread's window begins `delta` floats into a vector.  Nine aligned 128-bit reads and a
                // compile-time register selection per alignment class instead of 32 scalar reads.
                switch (delta) {
                    case 1: ncl_load32_shifted<1>(reinterpret_cast<const float*>(tile), base - 1, xr); break;
                    case 2: ncl_load32_shifted<2>(reinterpret_cast<const float*>(tile), base - 2, xr); break;
                    default: ncl_load32_shifted<3>(reinterpret_cast<const float*>(tile), base - 3, xr); break;
                }
            } else {  // bf16 row not 16-byte aligned: scalar shared-memory reads, still no index arithmetic per element
#pragma unroll
                for (int j = 0; j < NCL_XR; ++j) xr[j] = NclIO<TIO>::get(tile, base + j);
            }
        } else {
#pragma unroll
            for (int j = 0; j < NCL_XR; ++j) {
                int64_t idx = i0 - 8 + j;
                idx = idx < 0 ? 0 : (idx >= T ? T - 1 : idx);
                xr[j] = NclIO<TIO>::get(tile, delta + static_cast<int>(idx - lo));
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < NCL_XR; ++j) xr[j] = 0.f;
    }

    // interior threads (every activated pair m = i0-3 .. i0+18 lies inside the row) skip the replicate-padding logic of
    // the 2x signal: no 64-bit compares and selects per pair
    float out[NCL_PER_THREAD];
    if (i0 > 0 && i0 + NCL_PER_THREAD + 3 <= T)
        ncl_compute16<PRECISE, false>(xr, coef, i0, T, out);
    else
        ncl_compute16<PRECISE, true>(xr, coef, i0, T, out);
    TIO* yrow = y + (b * C + c) * T + i0;
    if (vec_store_ok) {
        if (active) {
            if (i0 + NCL_PER_THREAD <= T) {
                NclIO<TIO>::store16(yrow, out);
            } else {
#pragma unroll
                for (int q = 0; q < NCL_PER_THREAD; ++q)
                    if (i0 + q < T) NclIO<TIO>::store1(yrow + q, out[q]);
            }
        }
        return;
    }
    if (sizeof(TIO) == 4) {
        // fp32 rows that are not 16-byte aligned: every thread's 64-byte span still contains three aligned vectors
        if (active) {
            float* yr = reinterpret_cast<float*>(yrow);
            if (i0 + NCL_PER_THREAD <= T) {
                const int h = static_cast<int>((4 - ((reinterpret_cast<uintptr_t>(yr) >> 2) & 3)) & 3);
                switch (h) {
                    case 0: NclIO<float>::store16(yr, out); break;
                    case 1: ncl_store16_shifted<1>(yr, out); break;
                    case 2: ncl_store16_shifted<2>(yr, out); break;
                    default: ncl_store16_shifted<3>(yr, out); break;
                }
            } else {
#pragma unroll
                for (int q = 0; q < NCL_PER_THREAD; ++q)
                    if (i0 + q < T) yr[q] = out[q];
            }
        }
        return;
    }
    // bf16 rows that are not 16-byte aligned (odd T): per-thread stores would be 16 scattered scalar writes per lane, so the
    // tile is transposed through shared memory (one pad word per 32 keeps both phases bank-conflict free) and written
    // with fully coalesced scalar stores.
    __syncthreads();  // every thread has consumed its part of the input tile
    float* stage = reinterpret_cast<float*>(tile);
    if (active) {
#pragma unroll
        for (int q = 0; q < NCL_PER_THREAD; ++q) {
            const int w = threadIdx.x * NCL_PER_THREAD + q;
            stage[w + (w >> 5)] = out[q];
        }
    }
    __syncthreads();
    const int n_valid = static_cast<int>(min((int64_t)NCL_TILE, T - t0));
    TIO* ytile = y + (b * C + c) * T + t0;
    for (int w = threadIdx.x; w < n_valid; w += NCL_THREADS) NclIO<TIO>::store1(ytile + w, stage[w + (w >> 5)]);
}

// ------------------------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------------------------
template <typename TIn, typename TOut>
static int launch_nlc(const void* x, void* y, const float* alpha, const float* beta, int64_t B, int64_t T, int64_t C,
                      int precise, cudaStream_t st) {
    const int P = static_cast<int>(C / 2);
    int S = P;
    if (P > 256) {
        S = 1;
        for (int d = 256; d >= 1; --d)
            if (P % d == 0) { S = d; break; }
    }
    const int segs_per_block = S >= 256 ? 1 : 256 / S;
    // segment length: L + 6 must be a multiple of 7; prefer long segments (6 warm-up rows each) while keeping the
    // machine full
    static const int kL[] = {120, 64, 36, 22, 15, 8};
    int L = 8;
    for (int cand : kL) {
        if ((double)B * (double)ceil_div(T, cand) * (double)P >= 148.0 * 1536.0) { L = cand; break; }
    }
    const int64_t n_seg = ceil_div(T, L);
    dim3 grid(static_cast<unsigned>(ceil_div(n_seg, segs_per_block)), static_cast<unsigned>(P / S),
              static_cast<unsigned>(B));
    dim3 block(S * segs_per_block);
    // developer knob: unused dynamic shared memory per CTA caps the CTAs per SM (how many warps does the FMA pipe need?)
    static const int pad_kb = []() { const char* e = getenv("BVG_ACT_DIRECT_SMEM_KB"); return e ? atoi(e) : 0; }();
    const size_t pad = (size_t)pad_kb * 1024;
    if (pad > 0) {
        cudaFuncSetAttribute(act1d_nlc_kernel<TIn, TOut, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad);
        cudaFuncSetAttribute(act1d_nlc_kernel<TIn, TOut, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad);
    }
    if (precise)
        act1d_nlc_kernel<TIn, TOut, true><<<grid, block, pad, st>>>(static_cast<const TIn*>(x), static_cast<TOut*>(y),
                                                                     alpha, beta, T, (int)C, L, S, n_seg);
    else
        act1d_nlc_kernel<TIn, TOut, false><<<grid, block, pad, st>>>(static_cast<const TIn*>(x), static_cast<TOut*>(y),
                                                                      alpha, beta, T, (int)C, L, S, n_seg);
    BVG_LAUNCH_CHECK("act1d_nlc_kernel");
    return BVG_OK;
}

// Function attributes of one instantiation, set once per device (two driver calls per launch were a measurable part of
// the host time in front of every single-utterance forward).  Same L1/shared split as the conv kernel, so that CTAs of
// the two kernels can be resident on one SM at the same time.
template <typename TIn, typename TOut, int NH>
static int act_tma_attrs(bool precise) {
    static std::atomic<bool> done[2][64];
    int dev = 0;
    BVG_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && done[precise ? 1 : 0][dev].load(std::memory_order_acquire)) return BVG_OK;
    auto kern = precise ? act1d_nlc_tma_kernel<TIn, TOut, true, NH> : act1d_nlc_tma_kernel<TIn, TOut, false, NH>;
    BVG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
    BVG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    if (dev >= 0 && dev < 64) done[precise ? 1 : 0][dev].store(true, std::memory_order_release);
    return BVG_OK;
}

// TMA-pipelined launcher.  Returns BVG_OK after launching, or a positive value when the shape does not suit the TMA
// path (C not a multiple of 8, ...) and the caller should use the direct-load kernel.
template <typename TIn, typename TOut, int NH>
static int launch_nlc_tma(const void* x, const ActHeads<NH>& hd, int64_t B, int64_t T, int64_t C, int precise,
                          cudaStream_t st) {
    if (C % 8 != 0 || T >= (int64_t(1) << 31) - 4096 || reinterpret_cast<uintptr_t>(x) % 16 != 0) return 1;
    ActTmaParams p{};
    p.T = T;
    p.C = static_cast<int>(C);
    int CC = 0;
    for (int c = 64; c >= 8; c -= 8)
        if (C % c == 0) { CC = c; break; }
    p.CC = CC;
    p.P = CC / 2;
    p.nsegs = ACT_TMA_COMPUTE / p.P;
    p.n_cchunks = static_cast<int>(C / CC);
    const int sms = device_sm_count();
    // Segment length L (L + 6 a multiple of 7): minimise waves x (rows marched per unit incl. the 7 warm-up rows), i.e.
    // trade the per-segment warm-up (favours long segments) against the last-group padding and the tail wave of the
    // persistent grid (favour many short units).  Ties go to the longer segment.
    int L = 8;
    {
        const int64_t ctas = (int64_t)act_tma_min_ctas(NH) * sms;
        int64_t best = -1;
        for (int cand = 120; cand >= 8; cand -= 7) {
            const int64_t units_c = B * ceil_div(T, (int64_t)p.nsegs * cand) * p.n_cchunks;
            const int64_t waves = ceil_div(units_c, ctas);
            const int64_t cost = waves * (cand + 13);  // +6 rounding rows, +7 window rows
            if (best < 0 || cost < best) { best = cost; L = cand; }
        }
    }
    p.L = L;
    p.nbody = (L + 6) / 7;
    const int64_t groups = ceil_div(T, (int64_t)p.nsegs * L);
    const int64_t units = B * groups * p.n_cchunks;
    if (units >= (int64_t(1) << 31) || groups >= (int64_t(1) << 31)) return 1;
    p.groups_per_b = static_cast<uint32_t>(groups);
    p.n_units = static_cast<uint32_t>(units);
    p.row_bytes = static_cast<uint32_t>(CC * sizeof(TIn));
    static const int ring_kb = []() {
        const char* e = getenv("BVG_ACT_SMEM_KB");
        const int kb = e ? atoi(e) : 0;
        return (kb >= 16 && kb <= 100) ? kb : BVG_ACT_SMEM_KB_DEFAULT;
    }();
    // 7-row blocks per slab: with one block per slab the producer warp issues nsegs small TMA boxes (7 rows x <= 128 B)
    // per body and could not keep the ring full for the 21-segment C = 24 layout (ncu r01v: 13 % of the compute warps'
    // samples on the slab barrier); larger boxes cut the TMA issue and mbarrier traffic by rb at the same ring bytes.
    static const int rb_env = []() { const char* e = getenv("BVG_ACT_RB"); return e ? atoi(e) : 0; }();
    int stages = 0;
    p.rb = 1;
    for (int rb : {2, 4, 1}) {
        if (rb_env > 0 && rb != rb_env && rb != 1) continue;
        if (rb > p.nbody + 1) continue;
        const uint32_t seg_stride = static_cast<uint32_t>(round_up(7ll * rb * p.row_bytes, 128));
        const int st_c = static_cast<int>((ring_kb * 1024) / (p.nsegs * seg_stride));
        if (st_c >= 3 || rb == 1) {
            p.rb = rb;
            p.seg_stride = seg_stride;
            stages = st_c;
            break;
        }
    }
    p.stage_bytes = p.nsegs * p.seg_stride;
    p.tx_bytes = p.nsegs * 7u * p.rb * p.row_bytes;
    if (stages > ACT_TMA_MAX_STAGES) stages = ACT_TMA_MAX_STAGES;
    if (stages < 2) return 1;
    p.stages = stages;
    const uint32_t smem_bytes = stages * p.stage_bytes + 2 * ACT_TMA_MAX_STAGES * 8 + 128;

    EncodeTiledFn enc = get_encode_fn();
    BVG_REQUIRE(enc != nullptr, BVG_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    CUtensorMap map;
    const uint64_t es = sizeof(TIn);
    cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)T, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)C * es, (cuuint64_t)C * (cuuint64_t)T * es};
    cuuint32_t box[3] = {(cuuint32_t)CC, (cuuint32_t)(7 * p.rb), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&map, sizeof(TIn) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                     const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BVG_REQUIRE(r == CUDA_SUCCESS, BVG_ERR_CUDA, "cuTensorMapEncodeTiled(activation input) failed with CUresult %d", (int)r);

    auto kern = precise ? act1d_nlc_tma_kernel<TIn, TOut, true, NH> : act1d_nlc_tma_kernel<TIn, TOut, false, NH>;
    {
        const int rc_attr = act_tma_attrs<TIn, TOut, NH>(precise != 0);
        if (rc_attr != BVG_OK) return rc_attr;
    }
    static const int grid_mode = []() { const char* e = getenv("BVG_ACT_GRID"); return e ? atoi(e) : 0; }();
    const long long max_ctas = (long long)act_tma_min_ctas(NH) * sms;
    int grid = static_cast<int>(units < max_ctas ? units : max_ctas);
    if (grid_mode == 1) grid = static_cast<int>(units);                      // one unit per CTA: hardware load balancing
    if (grid_mode == 2) grid = static_cast<int>(units < sms ? units : sms);  // one CTA per SM
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(ACT_TMA_THREADS);
        cfg.dynamicSmemBytes = smem_bytes;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl_enabled() ? 1 : 0;
        BVG_CUDA(cudaLaunchKernelEx(&cfg, kern, map, hd, p));
    }
    BVG_LAUNCH_CHECK("act1d_nlc_tma_kernel");
    return BVG_OK;
}

// bvg_init(): attributes of every instantiation the forward plan launches
int activation1d_init() {
    for (int precise = 0; precise < 2; ++precise) {
        int rc = act_tma_attrs<float, __nv_bfloat16, 1>(precise);
        if (rc == BVG_OK) rc = act_tma_attrs<float, float, 1>(precise);
        if (rc == BVG_OK) rc = act_tma_attrs<float, Split6, 1>(precise);
        if (rc == BVG_OK) rc = act_tma_attrs<__nv_bfloat16, __nv_bfloat16, 1>(precise);
        if (rc == BVG_OK) rc = act_tma_attrs<__nv_bfloat16, float, 1>(precise);
        if (rc == BVG_OK) rc = act_tma_attrs<float, __nv_bfloat16, 3>(precise);
        if (rc == BVG_OK) rc = act_tma_attrs<float, float, 3>(precise);
        if (rc != BVG_OK) return rc;
    }
    return activation1d_umma_init();
}

static bool act_legacy_requested() {
    static const bool v = []() {
        const char* e = getenv("BVG_ACT_DIRECT");
        return e != nullptr && e[0] == '1';
    }();
    return v;
}

template <typename TIn, typename TOut>
static int launch_nlc_any(const void* x, void* y, const float* alpha, const float* beta, int64_t B, int64_t T, int64_t C,
                          int precise, cudaStream_t st) {
    if (!act_legacy_requested()) {
        ActHeads<1> hd;
        hd.y[0] = y;
        hd.alpha[0] = alpha;
        hd.beta[0] = beta;
        const int rc = launch_nlc_tma<TIn, TOut, 1>(x, hd, B, T, C, precise, st);
        if (rc <= 0) return rc;
    }
    return launch_nlc<TIn, TOut>(x, y, alpha, beta, B, T, C, precise, st);
}

int activation1d_nlc(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha, const float* beta,
                     int64_t B, int64_t T, int64_t C, int precise, cudaStream_t st) {
    BVG_REQUIRE(x && y && alpha && beta, BVG_ERR_INVALID, "activation1d_nlc: null pointer");
    BVG_REQUIRE(B >= 0 && T >= 0 && C > 0, BVG_ERR_INVALID, "activation1d_nlc: bad shape B=%lld T=%lld C=%lld",
                (long long)B, (long long)T, (long long)C);
    BVG_REQUIRE(C % 2 == 0, BVG_ERR_UNSUPPORTED, "activation1d_nlc: C=%lld must be even", (long long)C);
    BVG_REQUIRE(B <= 65535, BVG_ERR_UNSUPPORTED, "activation1d_nlc: B=%lld > 65535", (long long)B);
    if (B == 0 || T == 0) return BVG_OK;
    if (!precise && act_umma_requested(C)) {  // throughput mode: the tensor-core kernel (act1d_umma.cu)
        const int rc = activation1d_nlc_umma(x, x_dtype, y, y_dtype, alpha, beta, B, T, C, st);
        if (rc <= 0) return rc;
    }
    if (x_dtype == BVG_F32 && y_dtype == BVG_BF16)
        return launch_nlc_any<float, __nv_bfloat16>(x, y, alpha, beta, B, T, C, precise, st);
    if (x_dtype == BVG_F32 && y_dtype == BVG_F32) return launch_nlc_any<float, float>(x, y, alpha, beta, B, T, C, precise, st);
    if (x_dtype == BVG_F32 && y_dtype == BVG_BF16X3)  // fp32 parity mode on the tensor cores: split output [B, T, 6*C]
        return launch_nlc_any<float, Split6>(x, y, alpha, beta, B, T, C, precise, st);
    if (x_dtype == BVG_BF16 && y_dtype == BVG_BF16)
        return launch_nlc_any<__nv_bfloat16, __nv_bfloat16>(x, y, alpha, beta, B, T, C, precise, st);
    if (x_dtype == BVG_BF16 && y_dtype == BVG_F32)
        return launch_nlc_any<__nv_bfloat16, float>(x, y, alpha, beta, B, T, C, precise, st);
    return set_error(BVG_ERR_UNSUPPORTED, "activation1d_nlc: unsupported dtypes %d -> %d", x_dtype, y_dtype);
}

// Three Activation1d of the same tensor (the first activations of the three AMP blocks of a stage) in one launch; falls
// back to three launches for shapes the TMA kernel does not take.
template <typename TIn, typename TOut>
static int launch_nlc_heads3(const void* x, void* const* ys, const float* const* alphas, const float* const* betas, int64_t B,
                             int64_t T, int64_t C, int precise, cudaStream_t st) {
    if (!act_legacy_requested()) {
        ActHeads<3> hd;
        for (int h = 0; h < 3; ++h) {
            hd.y[h] = ys[h];
            hd.alpha[h] = alphas[h];
            hd.beta[h] = betas[h];
        }
        const int rc = launch_nlc_tma<TIn, TOut, 3>(x, hd, B, T, C, precise, st);
        if (rc <= 0) return rc;
    }
    for (int h = 0; h < 3; ++h) {
        const int rc = launch_nlc_any<TIn, TOut>(x, ys[h], alphas[h], betas[h], B, T, C, precise, st);
        if (rc != BVG_OK) return rc;
    }
    return BVG_OK;
}

int activation1d_nlc_multi(const void* x, int x_dtype, void* const* ys, int y_dtype, const float* const* alphas,
                           const float* const* betas, int n_heads, int64_t B, int64_t T, int64_t C, int precise,
                           cudaStream_t st) {
    BVG_REQUIRE(x && ys && alphas && betas && n_heads >= 1 && n_heads <= 8, BVG_ERR_INVALID,
                "activation1d_nlc_multi: null pointer or bad head count %d", n_heads);
    for (int h = 0; h < n_heads; ++h) {
        BVG_REQUIRE(ys[h] && alphas[h] && betas[h], BVG_ERR_INVALID, "activation1d_nlc_multi: null pointer (head %d)", h);
        BVG_REQUIRE(ys[h] != x, BVG_ERR_INVALID, "activation1d_nlc_multi: in-place is not supported");
    }
    BVG_REQUIRE(B >= 0 && T >= 0 && C > 0 && C % 2 == 0 && B <= 65535, BVG_ERR_INVALID,
                "activation1d_nlc_multi: bad shape B=%lld T=%lld C=%lld", (long long)B, (long long)T, (long long)C);
    if (B == 0 || T == 0) return BVG_OK;
    int h0 = 0;
    // fast-sin mode on shapes the tensor-core kernel takes: one launch per head (three of those beat the shared-up-sampler
    // SIMT kernel, and results stay bit-identical to single calls by construction)
    const bool per_head = !precise && act_umma_requested(C) && C % 8 == 0;
    for (; !per_head && h0 + 3 <= n_heads; h0 += 3) {
        int rc;
        if (x_dtype == BVG_F32 && y_dtype == BVG_BF16)
            rc = launch_nlc_heads3<float, __nv_bfloat16>(x, ys + h0, alphas + h0, betas + h0, B, T, C, precise, st);
        else if (x_dtype == BVG_F32 && y_dtype == BVG_F32)
            rc = launch_nlc_heads3<float, float>(x, ys + h0, alphas + h0, betas + h0, B, T, C, precise, st);
        else
            break;  // bf16-input combinations: no fused instantiation, one launch per head below
        if (rc != BVG_OK) return rc;
    }
    for (int h = h0; h < n_heads; ++h) {
        const int rc = activation1d_nlc(x, x_dtype, ys[h], y_dtype, alphas[h], betas[h], B, T, C, precise, st);
        if (rc != BVG_OK) return rc;
    }
    return BVG_OK;
}

template <typename TIO>
static int launch_ncl(const void* x, void* y, const float* alpha, const float* beta, int64_t B, int64_t C, int64_t T,
                      int64_t sb, int64_t sc, int precise, cudaStream_t st) {
    const int vec_ok = ((T * (int64_t)sizeof(TIO)) % 16 == 0) && (reinterpret_cast<uintptr_t>(y) % 16 == 0);
    dim3 grid(static_cast<unsigned>(ceil_div(T, NCL_TILE)), static_cast<unsigned>(C), static_cast<unsigned>(B));
    if (precise)
        act1d_ncl_kernel<TIO, true><<<grid, NCL_THREADS, 0, st>>>(static_cast<const TIO*>(x), static_cast<TIO*>(y),
                                                                   alpha, beta, C, T, sb, sc, vec_ok);
    else
        act1d_ncl_kernel<TIO, false><<<grid, NCL_THREADS, 0, st>>>(static_cast<const TIO*>(x), static_cast<TIO*>(y),
                                                                    alpha, beta, C, T, sb, sc, vec_ok);
    BVG_LAUNCH_CHECK("act1d_ncl_kernel");
    return BVG_OK;
}

}  // namespace bvg

using namespace bvg;

extern "C" int bvg_activation1d_fwd(const void* x, void* y, const float* alpha_log, const float* beta_log, int64_t B,
                                    int64_t C, int64_t T, int64_t x_stride_b, int64_t x_stride_c, int dtype,
                                    int precise, void* stream) {
    BVG_REQUIRE(x && y && alpha_log && beta_log, BVG_ERR_INVALID, "bvg_activation1d_fwd: null pointer");
    BVG_REQUIRE(B >= 0 && C >= 0 && T >= 0, BVG_ERR_INVALID, "bvg_activation1d_fwd: negative shape");
    BVG_REQUIRE(C <= 65535 && B <= 65535, BVG_ERR_UNSUPPORTED, "bvg_activation1d_fwd: B, C must be <= 65535");
    BVG_REQUIRE(dtype == BVG_F32 || dtype == BVG_BF16, BVG_ERR_UNSUPPORTED,
                "bvg_activation1d_fwd: dtype %d not supported (fp32 and bf16 only)", dtype);
    const int64_t es = dtype == BVG_F32 ? 4 : 2;
    BVG_REQUIRE(reinterpret_cast<uintptr_t>(x) % es == 0 && reinterpret_cast<uintptr_t>(y) % es == 0, BVG_ERR_INVALID,
                "bvg_activation1d_fwd: pointers not element aligned");
    BVG_REQUIRE(x_stride_c >= T && x_stride_b >= 0, BVG_ERR_INVALID, "bvg_activation1d_fwd: bad strides");
    if (B == 0 || C == 0 || T == 0) return BVG_OK;  // like the reference (.cu:193-196): empty input is a no-op
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == BVG_F32)
        return launch_ncl<float>(x, y, alpha_log, beta_log, B, C, T, x_stride_b, x_stride_c, precise, st);
    return launch_ncl<__nv_bfloat16>(x, y, alpha_log, beta_log, B, C, T, x_stride_b, x_stride_c, precise, st);
}

extern "C" int bvg_activation1d_nlc_fwd(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha_log,
                                        const float* beta_log, int64_t B, int64_t T, int64_t C, int precise,
                                        void* stream) {
    return activation1d_nlc(x, x_dtype, y, y_dtype, alpha_log, beta_log, B, T, C, precise,
                            static_cast<cudaStream_t>(stream));
}

extern "C" int bvg_activation1d_nlc_multi_fwd(const void* x, int x_dtype, void* const* ys, int y_dtype,
                                              const float* const* alphas_log, const float* const* betas_log, int n_heads,
                                              int64_t B, int64_t T, int64_t C, int precise, void* stream) {
    return activation1d_nlc_multi(x, x_dtype, ys, y_dtype, alphas_log, betas_log, n_heads, B, T, C, precise,
                                  static_cast<cudaStream_t>(stream));
}
