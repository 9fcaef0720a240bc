// Activation1d with both FIR filters on the tensor cores (tcgen05, A operand in tensor memory).
//
// Replaces the reference's fused extension kernel (alias_free_activation/cuda/anti_alias_activation_cuda.cu:43-179) and
// its torch path (alias_free_activation/torch/act.py:25-30) in the internal channels-last layout, fast-sin mode.  Same
// closed forms as act1d.cu, cl() = clamp to [0, T-1]:
//     u[2m]   = sum_j E_even[j] x[cl(m+j)]   j = -3..2        u[2m+1] = sum_j E_odd[j] x[cl(m+j)]   j = -2..3
//     v[n]    = u[n] + 1/(exp(beta)+1e-9) * sin^2(exp(alpha) * u[n])
//     y[i]    = sum_{k<12} f[k] v[clamp(2i+k-5, 0, 2T-1)]
//
// Why: the SIMT kernel (act1d.cu) is bound by the FP32 pipe -- 24 of its 29 packed operations per element pair are the two
// FIRs -- while the tensor pipe idles (ncu r01: FMA pipe 67 %, tensor pipe 0 %, 0.36 of the HBM roofline).  Here a FIR is
// a GEMM with a constant banded Toeplitz matrix:
//     U[signal, u-time] = X[signal, x-time] * Bup[u-time, x-time]^T        (M = 128 signals, N = 64, K = 48)
//     Y[signal, y-time] = V[signal, v-time] * Bdn[y-time, v-time]^T        (M = 128 signals, N = 32, K = 80)
// M is the SIGNAL axis: a TMEM lane is one independent 1-D signal (a channel of one time block), so a thread that reads its
// lane with tcgen05.ld (32x32b) owns consecutive time samples of ONE channel and keeps that channel's snake parameters
// in registers.  The signal operand lives in tensor memory (tcgen05.mma with A in TMEM: probed in
// tools/probes/probe_umma_ts.cu -- lane = row, one 32-bit column = two consecutive fp16 K elements, any 4-column
// aligned start), written with tcgen05.st by the same threads that produced it: x after conversion to fp16, v straight
// from the snake arithmetic.  The 2x signal therefore never touches shared or global memory, and the sliding FIR
// window is just a moving column offset into a ring of TMEM columns.  The Toeplitz matrices (8 KB each, fp16, K-major,
// 128-byte swizzle) are built once per CTA from the six filter taps.
//
// fp16, not bf16: 11 significant bits keep the rounding of x, v and the taps at 2^-12 (-72 dB), far below the bf16
// rounding of the kernel's own output (2^-9); products are exact and accumulate in fp32.  |x| must stay below ~3e4
// (fp16 range) -- true for audio-scale activations; the fp32 parity mode keeps the SIMT kernel.
//
// Work decomposition: a CTA processes units of 128 signals = (128 / cw time blocks) x (cw channels), cw = the largest
// power of two <= 128 dividing C, so every C % 8 == 0 fills all 128 lanes.  A lane marches along its time block in
// chunks of 32 output rows; per chunk c:  stage-in X(c) -> UP(c) -> snake(c) -> V(c) -> DOWN(c-1) -> out(c-1).
// Warp 0 issues the MMAs; two warpgroups (warps 1-4 and 5-8, each covering the 128 TMEM lanes) alternate chunks, so the
// snake of chunk c (MUFU-bound: one cosine per 2x sample) overlaps the MMAs and the global loads / stores of its
// neighbours.  All hand-offs are mbarriers (full/empty per ring slot); MMA completion arrives through tcgen05.commit.
#include "bvg_common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

#include <atomic>
#include <cstdlib>
#include <cuda_fp16.h>

namespace bvg {

constexpr int AU_THREADS = 288;        // warp 0: MMA issuer / TMEM owner; warps 1-4: group 0; warps 5-8: group 1
constexpr int AU_CHUNK = 32;           // output rows per chunk (64 samples of the 2x signal)
// tensor-memory columns
constexpr uint32_t AU_DU = 0;          // 2 accumulators x 64 columns: U of a chunk (fp32)
constexpr uint32_t AU_DY = 128;        // 2 accumulators x 32 columns: Y of a chunk (fp32)
constexpr uint32_t AU_X = 192;         // ring of 4 chunks x 16 columns (32 fp16 x-rows each) + 8 mirror columns
constexpr uint32_t AU_V = 272;         // ring of 4 chunks x 32 columns (64 fp16 v-rows each) + 8 mirror columns
constexpr uint32_t AU_TMEM_COLS = 512;
constexpr int AU_SMEM_BYTES = 1024 + 8192 + 8192 + 512;

struct ActUmmaParams {
    const void* x;
    void* y;
    const float* alpha;
    const float* beta;
    int64_t T;
    int C;
    int cw_shift;            // log2(channels per unit)
    int nc;                  // chunks (of 32 output rows) per time block
    uint32_t n_cchunks;      // C / cw
    uint32_t groups_per_b;   // time-block groups per batch item
    uint32_t n_units;
};

__device__ __forceinline__ void tmem_st_x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem descriptor]^T, fp16 operands, fp32 accumulation, one issuing thread.
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// kind::f16 instruction descriptor: D fp32, A fp16, B fp16, both K-major
__host__ __device__ constexpr uint32_t umma_idesc_f16_f32(uint32_t M, uint32_t N) {
    return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// 12-tap filter (bvg_common.cuh): f[k] = F[min(k, 11 - k)]
__device__ __forceinline__ float au_tap(int k) {
    const int i = k < 6 ? k : 11 - k;
    return i == 0 ? BVG_F0 : i == 1 ? BVG_F1 : i == 2 ? BVG_F2 : i == 3 ? BVG_F3 : i == 4 ? BVG_F4 : BVG_F5;
}
// Up-sampler Toeplitz entry.  Column n of a U chunk is 2x-sample 2*y0 + 64*c - 8 + n, i.e. m = y0 + 32*c - 4 + (n >> 1) with
// phase n & 1; K index k of the window is x row y0 + 32*c - 8 + k, so x[m + d] sits at k = (n >> 1) + 4 + d.
//   even: u = E0 x[m-3] + E2 x[m-2] + E4 x[m-1] + E5 x[m] + E3 x[m+1] + E1 x[m+2]
//   odd : u = E1 x[m-2] + E3 x[m-1] + E5 x[m] + E4 x[m+1] + E2 x[m+2] + E0 x[m+3]          (E = 2 f)
__device__ __forceinline__ float au_up_entry(int n, int k) {
    const int d = k - (n >> 1) - 4;
    int tap;
    if ((n & 1) == 0) {
        if (d < -3 || d > 2) return 0.f;
        tap = d == -3 ? 0 : d == -2 ? 2 : d == -1 ? 4 : d == 0 ? 5 : d == 1 ? 3 : 1;
    } else {
        if (d < -2 || d > 3) return 0.f;
        tap = d == -2 ? 1 : d == -1 ? 3 : d == 0 ? 5 : d == 1 ? 4 : d == 2 ? 2 : 0;
    }
    return 2.0f * au_tap(tap);
}
// Low-pass Toeplitz entry.  Output i of a Y chunk is row y0 + 32*c + i = sum_t f[t] v[2 (y0 + 32 c + i) + t - 5]; K index k
// of the window is 2x-sample 2*y0 + 64*c - 8 + k, so tap t sits at k = 2 i + t + 3.
__device__ __forceinline__ float au_dn_entry(int i, int k) {
    const int t = k - 2 * i - 3;
    return (t < 0 || t > 11) ? 0.f : au_tap(t);
}
// byte offset of element (row, k) in a K-major SWIZZLE_128B tile (rows of 64 fp16 = 128 B; 16-byte chunk index XOR row % 8)
__device__ __forceinline__ uint32_t au_sw128(int row, int k) {
    return static_cast<uint32_t>(row * 128 + ((((k >> 3) ^ row) & 7) << 4) + (k & 7) * 2);
}

template <typename T>
struct AuIO;
template <>
struct AuIO<float> {
    static __device__ __forceinline__ float load(const float* p) { return __ldg(p); }
    static __device__ __forceinline__ void store(float* p, float v) { *p = v; }
};
template <>
struct AuIO<__nv_bfloat16> {
    static __device__ __forceinline__ float load(const __nv_bfloat16* p) {
        return __uint_as_float(static_cast<uint32_t>(__ldg(reinterpret_cast<const unsigned short*>(p))) << 16);
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

__device__ __forceinline__ uint32_t au_pack_half2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(AU_THREADS, 1)
act1d_umma_kernel(const __grid_constant__ ActUmmaParams p) {
    extern __shared__ uint8_t au_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(au_smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sBup = smem;          // [64 rows = u-time][64 K] fp16: K = 48 used
    uint8_t* sBdn = smem + 8192;   // 2 K-blocks x [32 rows = y-time][64 K] fp16: K = 80 used
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 16384);
    uint64_t* x_full = bars;          // [4] count 4 (the producing group's warps)
    uint64_t* x_empty = bars + 4;     // [4] count 1 (tcgen05.commit)
    uint64_t* v_full = bars + 8;      // [4] count 4
    uint64_t* v_empty = bars + 12;    // [4] count 1
    uint64_t* du_full = bars + 16;    // [2] count 1
    uint64_t* du_empty = bars + 18;   // [2] count 4
    uint64_t* dy_full = bars + 20;    // [2] count 1
    uint64_t* dy_empty = bars + 22;   // [2] count 4
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 24);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    griddep_launch_dependents();  // PDL: the prologue below touches constants only
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) {
            mbar_init(&x_full[i], 4); mbar_init(&x_empty[i], 1);
            mbar_init(&v_full[i], 4); mbar_init(&v_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&du_full[i], 1); mbar_init(&du_empty[i], 4);
            mbar_init(&dy_full[i], 1); mbar_init(&dy_empty[i], 4);
        }
        fence_barrier_init();
    }
    // Toeplitz operands
    for (int e = threadIdx.x; e < 64 * 64; e += AU_THREADS) {
        const int n = e >> 6, k = e & 63;
        *reinterpret_cast<__half*>(sBup + au_sw128(n, k)) = __float2half_rn(k < 48 ? au_up_entry(n, k) : 0.f);
    }
    for (int e = threadIdx.x; e < 32 * 128; e += AU_THREADS) {
        const int i = e >> 7, k = e & 127;
        *reinterpret_cast<__half*>(sBdn + (k >> 6) * 4096 + au_sw128(i, k & 63)) = __float2half_rn(k < 80 ? au_dn_entry(i, k) : 0.f);
    }
    fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
    if (warp == 0) tmem_alloc(tmem_ptr, AU_TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;
    griddep_wait();  // x is produced by the previous kernel of the stream (and y may still be read by it)

    const int nc = p.nc;
    uint32_t G0 = 0;  // running chunk number at the start of the unit: identical in every role, drives all ring slots

    if (warp == 0) {
        // ------------------------------ MMA issuer ------------------------------
        const bool leader = elect_one();
        const uint32_t idesc_up = umma_idesc_f16_f32(128, 64), idesc_dn = umma_idesc_f16_f32(128, 32);
        const uint64_t bup = umma_desc_sw128_kmajor(smem_u32(sBup)), bdn = umma_desc_sw128_kmajor(smem_u32(sBdn));
        uint32_t dyc[2] = {0u, 0u};  // uses of each Y accumulator so far
        for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
            for (int j = 0; j <= nc + 1; ++j) {
                if (j <= nc) {  // UP(j): U chunk j from X chunk j and the first half of X chunk j + 1
                    const uint32_t G = G0 + j, slot = G & 3u, q = G & 1u;
                    mbar_wait_warp(&x_full[slot], (G >> 2) & 1u);
                    if (j < nc) mbar_wait_warp(&x_full[(G + 1) & 3u], ((G + 1) >> 2) & 1u);
                    mbar_wait_warp(&du_empty[q], ((G >> 1) & 1u) ^ 1u);
                    tc_fence_after_sync();
                    if (leader) {
                        const int ksteps = j < nc ? 3 : 2;  // the last chunk of a block needs rows of its own X chunk only
                        const uint32_t d = tmem_base + AU_DU + 64u * q, a = tmem_base + AU_X + 16u * slot;
                        for (int kk = 0; kk < ksteps; ++kk) umma_f16_ts(d, a + 8u * kk, bup + 2u * kk, idesc_up, kk > 0);
                        umma_commit(&du_full[q]);
                        umma_commit(&x_empty[slot]);
                    }
                    __syncwarp();
                }
                const int s = j - 2;
                if (s >= 0 && s < nc) {  // DOWN(s): Y chunk s from V chunk s and the first 16 samples of V chunk s + 1
                    const uint32_t G = G0 + s, slot = G & 3u, q = G & 1u;
                    mbar_wait_warp(&v_full[slot], (G >> 2) & 1u);
                    mbar_wait_warp(&v_full[(G + 1) & 3u], ((G + 1) >> 2) & 1u);
                    mbar_wait_warp(&dy_empty[q], (dyc[q] & 1u) ^ 1u);
                    tc_fence_after_sync();
                    if (leader) {
                        const uint32_t d = tmem_base + AU_DY + 32u * q, a = tmem_base + AU_V + 32u * slot;
                        for (int kk = 0; kk < 5; ++kk)
                            umma_f16_ts(d, a + 8u * kk, kk < 4 ? bdn + 2u * kk : bdn + (4096u >> 4), idesc_dn, kk > 0);
                        umma_commit(&dy_full[q]);
                        umma_commit(&v_empty[slot]);
                        if (s == nc - 1) umma_commit(&v_empty[(G + 1) & 3u]);  // V chunk nc has no DOWN of its own
                    }
                    __syncwarp();
                    ++dyc[q];
                }
            }
            G0 += static_cast<uint32_t>(nc) + 1u;
        }
    } else {
        // ------------------------------ compute groups ------------------------------
        const uint32_t wg = static_cast<uint32_t>(warp - 1) >> 2;       // 0 / 1: owns the chunks with G % 2 == wg
        const uint32_t quarter = static_cast<uint32_t>(warp) & 3u;      // TMEM lane quarter this warp may access
        const uint32_t sig = quarter * 32u + static_cast<uint32_t>(lane);  // signal (TMEM lane) of this thread
        const uint32_t tl = tmem_base + ((quarter * 32u) << 16);
        const int cw_mask = (1 << p.cw_shift) - 1;
        const int64_t T = p.T, C = p.C;
        const int64_t Lq = static_cast<int64_t>(nc) * AU_CHUNK;
        const int64_t ntb = 128 >> p.cw_shift;
        uint32_t dyc = 0;     // uses of this group's Y accumulator
        float vl = 0.f;       // last valid activated sample (replicate padding of the 2x signal on the right)

        for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
            const uint32_t cc = unit % p.n_cchunks;
            const uint32_t rest = unit / p.n_cchunks;
            const uint32_t b = rest / p.groups_per_b;
            const uint32_t g = rest - b * p.groups_per_b;
            const int ch = static_cast<int>(cc << p.cw_shift) + (static_cast<int>(sig) & cw_mask);
            const int64_t y0 = (static_cast<int64_t>(g) * ntb + (sig >> p.cw_shift)) * Lq;  // first output row of this lane
            const float a2 = 2.0f * expf(__ldg(p.alpha + ch));
            const float cb = 0.5f / (expf(__ldg(p.beta + ch)) + 1e-9f);
            const TIn* xb = static_cast<const TIn*>(p.x) + static_cast<int64_t>(b) * T * C + ch;
            TOut* yb = static_cast<TOut*>(p.y) + static_cast<int64_t>(b) * T * C + ch;
            const float2 a2v = make_float2(a2, a2), ncb = make_float2(-cb, -cb);

            float xr[32];  // prefetched x rows of the next chunk this group stages (as fp32)
            auto load_chunk = [&](int i) {
                const int64_t t0 = y0 - 8 + static_cast<int64_t>(i) * AU_CHUNK;
                const bool interior = (t0 >= 0) && (t0 + 31 < T);
                if (__all_sync(0xffffffffu, interior)) {
                    const TIn* pr = xb + t0 * C;
#pragma unroll
                    for (int r = 0; r < 32; ++r) xr[r] = AuIO<TIn>::load(pr + r * C);
                } else {  // replicate padding of x at the sequence ends (and finite data for lanes past the end)
#pragma unroll
                    for (int r = 0; r < 32; ++r) {
                        int64_t t = t0 + r;
                        t = t < 0 ? 0 : (t >= T ? T - 1 : t);
                        xr[r] = AuIO<TIn>::load(xb + t * C);
                    }
                }
            };

            const int i0 = ((G0 & 1u) == wg) ? 0 : 1;
            if (i0 <= nc) load_chunk(i0);
            for (int i = i0; i <= nc + 4; i += 2) {
                const uint32_t G = G0 + static_cast<uint32_t>(i);
                if (i <= nc) {
                    // ---- stage-in: X chunk i (32 rows) -> fp16 pairs -> 16 TMEM columns of ring slot G % 4
                    const uint32_t slot = G & 3u;
                    uint32_t pk[16];
#pragma unroll
                    for (int r = 0; r < 16; ++r) pk[r] = au_pack_half2(xr[2 * r], xr[2 * r + 1]);
                    mbar_wait_warp(&x_empty[slot], ((G >> 2) & 1u) ^ 1u);
                    tc_fence_after_sync();
                    tmem_st_x16(tl + AU_X + 16u * slot, pk);
                    if (slot == 0u) tmem_st_x8(tl + AU_X + 64u, pk);  // mirror: the window of a slot-3 chunk runs on into it
                    tmem_st_wait();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&x_full[slot]);
                }
                if (i + 2 <= nc) load_chunk(i + 2);
                const int j = i - 2;
                if (j >= 0 && j <= nc) {
                    // ---- snake: U chunk j (fp32, 64 columns) -> v = u - c cos(2 a u) -> fp16 -> V ring slot
                    const uint32_t Gj = G - 2u, slot = Gj & 3u;
                    mbar_wait_warp(&du_full[wg], (Gj >> 1) & 1u);
                    mbar_wait_warp(&v_empty[slot], ((Gj >> 2) & 1u) ^ 1u);
                    tc_fence_after_sync();
                    // column q of the chunk is 2x-sample 2*y0 + 64*j - 8 + q of this lane's sequence
                    const int64_t n0 = 2 * y0 + 64 * static_cast<int64_t>(j) - 8;
                    const bool left = (y0 == 0) && (j == 0);
                    const int64_t ql64 = 2 * T - 1 - n0;
                    const int qlast = ql64 > 64 ? 64 : (ql64 < -1 ? -1 : static_cast<int>(ql64));  // column of sample 2T-1
                    const bool edge = __any_sync(0xffffffffu, left || qlast < 63);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t uv[32];
                        tmem_ld_32x32b_x32(tl + AU_DU + 64u * wg + 32u * h, uv);
                        tmem_ld_wait();
                        if (h == 1) {  // both halves are in registers: hand the accumulator back to the MMA warp
                            tc_fence_before_sync();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&du_empty[wg]);
                        }
                        float v[32];
#pragma unroll
                        for (int e = 0; e < 32; e += 2) {
                            const float2 u = make_float2(__uint_as_float(uv[e]), __uint_as_float(uv[e + 1]));
                            const float2 t = __fmul2_rn(u, a2v);
                            const float2 cs = make_float2(__cosf(t.x), __cosf(t.y));
                            const float2 r = __ffma2_rn(ncb, cs, u);
                            v[e] = r.x;
                            v[e + 1] = r.y;
                        }
                        if (edge) {
                            if (left && h == 0) {  // samples before 0 repeat v[0] (column 8 of the first chunk)
#pragma unroll
                                for (int e = 0; e < 8; ++e) v[e] = v[8];
                            }
#pragma unroll
                            for (int e = 0; e < 32; ++e) {  // samples after 2T-1 repeat v[2T-1]
                                if (32 * h + e > qlast) v[e] = vl; else vl = v[e];
                            }
                        }
                        uint32_t pk[16];
#pragma unroll
                        for (int e = 0; e < 16; ++e) pk[e] = au_pack_half2(v[2 * e], v[2 * e + 1]);
                        tmem_st_x16(tl + AU_V + 32u * slot + 16u * h, pk);
                        if (h == 0 && slot == 0u) tmem_st_x8(tl + AU_V + 128u, pk);  // mirror (see X ring)
                    }
                    tmem_st_wait();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&v_full[slot]);
                }
                const int s = i - 4;
                if (s >= 0 && s < nc) {
                    // ---- out: Y chunk s (fp32, 32 columns) + c -> rows y0 + 32 s .. of this lane's channel
                    mbar_wait_warp(&dy_full[wg], dyc & 1u);
                    tc_fence_after_sync();
                    uint32_t dv[32];
                    tmem_ld_32x32b_x32(tl + AU_DY + 32u * wg, dv);
                    tmem_ld_wait();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&dy_empty[wg]);
                    ++dyc;
                    const int64_t t0 = y0 + static_cast<int64_t>(s) * AU_CHUNK;
                    const int64_t left_rows = T - t0;
                    const int n_ok = left_rows >= 32 ? 32 : (left_rows < 0 ? 0 : static_cast<int>(left_rows));
                    TOut* pr = yb + t0 * C;
                    if (__all_sync(0xffffffffu, n_ok == 32)) {
#pragma unroll
                        for (int r = 0; r < 32; ++r) AuIO<TOut>::store(pr + r * C, __uint_as_float(dv[r]) + cb);
                    } else {
#pragma unroll
                        for (int r = 0; r < 32; ++r)
                            if (r < n_ok) AuIO<TOut>::store(pr + r * C, __uint_as_float(dv[r]) + cb);
                    }
                }
            }
            G0 += static_cast<uint32_t>(nc) + 1u;
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, AU_TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
template <typename TIn, typename TOut>
static int act_umma_attrs() {
    static std::atomic<bool> done[64];
    int dev = 0;
    BVG_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && done[dev].load(std::memory_order_acquire)) return BVG_OK;
    BVG_CUDA(cudaFuncSetAttribute(act1d_umma_kernel<TIn, TOut>, cudaFuncAttributeMaxDynamicSharedMemorySize, AU_SMEM_BYTES));
    if (dev >= 0 && dev < 64) done[dev].store(true, std::memory_order_release);
    return BVG_OK;
}

// Returns BVG_OK after launching, a positive value when the shape does not suit this kernel (the caller then uses the
// SIMT kernel), a negative status on error.
template <typename TIn, typename TOut>
static int launch_act_umma(const void* x, void* y, const float* alpha, const float* beta, int64_t B, int64_t T, int64_t C,
                           cudaStream_t st) {
    if (C % 8 != 0 || C > (1 << 20) || T < 1 || T >= (int64_t(1) << 40)) return 1;
    ActUmmaParams p{};
    p.x = x; p.y = y; p.alpha = alpha; p.beta = beta; p.T = T; p.C = static_cast<int>(C);
    int sh = 3;
    while (sh < 7 && C % (int64_t(2) << sh) == 0) ++sh;  // cw = largest power of two <= 128 dividing C
    p.cw_shift = sh;
    const int64_t cw = int64_t(1) << sh, ntb = 128 / cw;
    p.n_cchunks = static_cast<uint32_t>(C / cw);
    const int sms = device_sm_count();
    // Chunks per time block: long blocks amortise the pipeline fill / drain (~4 chunk times per unit), short ones balance
    // the persistent grid.  cost = waves x (nc + 4).
    int nc = 1;
    {
        double best = -1.0;
        const int64_t nc_max = ceil_div(T, AU_CHUNK);
        for (int64_t cand = 1; cand <= nc_max && cand <= 4096; ++cand) {
            const int64_t groups = ceil_div(T, ntb * AU_CHUNK * cand);
            const int64_t units = B * p.n_cchunks * groups;
            const double cost = static_cast<double>(ceil_div(units, sms)) * static_cast<double>(cand + 4);
            if (best < 0 || cost <= best) { best = cost; nc = static_cast<int>(cand); }
        }
    }
    p.nc = nc;
    const int64_t groups = ceil_div(T, ntb * AU_CHUNK * nc);
    const int64_t units = B * p.n_cchunks * groups;
    if (units >= (int64_t(1) << 31) || groups >= (int64_t(1) << 31)) return 1;
    p.groups_per_b = static_cast<uint32_t>(groups);
    p.n_units = static_cast<uint32_t>(units);
    {
        const int rc = act_umma_attrs<TIn, TOut>();
        if (rc != BVG_OK) return rc;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(units < sms ? units : sms));
    cfg.blockDim = dim3(AU_THREADS);
    cfg.dynamicSmemBytes = AU_SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    BVG_CUDA(cudaLaunchKernelEx(&cfg, act1d_umma_kernel<TIn, TOut>, p));
    BVG_LAUNCH_CHECK("act1d_umma_kernel");
    return BVG_OK;
}

// BVG_ACT_UMMA=0 keeps the SIMT kernel (A/B runs)
bool act_umma_requested() {
    static const bool v = []() {
        const char* e = getenv("BVG_ACT_UMMA");
        return !(e != nullptr && e[0] == '0');
    }();
    return v;
}

int activation1d_nlc_umma(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha, const float* beta, int64_t B,
                          int64_t T, int64_t C, cudaStream_t st) {
    if (x_dtype == BVG_F32 && y_dtype == BVG_BF16) return launch_act_umma<float, __nv_bfloat16>(x, y, alpha, beta, B, T, C, st);
    if (x_dtype == BVG_BF16 && y_dtype == BVG_BF16)
        return launch_act_umma<__nv_bfloat16, __nv_bfloat16>(x, y, alpha, beta, B, T, C, st);
    if (x_dtype == BVG_F32 && y_dtype == BVG_F32) return launch_act_umma<float, float>(x, y, alpha, beta, B, T, C, st);
    return 1;
}

int activation1d_umma_init() {
    int rc = act_umma_attrs<float, __nv_bfloat16>();
    if (rc == BVG_OK) rc = act_umma_attrs<__nv_bfloat16, __nv_bfloat16>();
    if (rc == BVG_OK) rc = act_umma_attrs<float, float>();
    return rc;
}

}  // namespace bvg
