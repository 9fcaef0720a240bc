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
// Warp 0 issues the up-sampler MMAs and warp 17 the low-pass MMAs (an issuer spends more time on its barrier waits than on
// its 3 / 5 instructions, so the two chains run in parallel); two groups of eight compute warps alternate chunks, and
// inside a group two warps per TMEM lane quarter split a chunk's rows / columns in halves, so the snake of chunk c
// (MUFU-bound: one cosine per 2x sample) overlaps the MMAs and the global loads / stores of its neighbours.  All hand-offs are mbarriers (full/empty per ring slot); MMA completion arrives
// through tcgen05.commit.
#include "bvg_common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

#include <atomic>
#include <cstdlib>
#include <cuda_fp16.h>

namespace bvg {

constexpr int AU_THREADS = 576;        // warp 0: UP issuer / TMEM owner; warps 1-16: compute; warp 17: DOWN issuer
constexpr int AU_CHUNK = 32;           // output rows per chunk (64 samples of the 2x signal)
// tensor-memory columns
constexpr uint32_t AU_DU = 0;          // 2 accumulators x 64 columns: U of a chunk (fp32)
constexpr uint32_t AU_DY = 128;        // 2 accumulators x 32 columns: Y of a chunk (fp32)
constexpr uint32_t AU_X = 192;         // ring of 4 chunks x 16 columns (32 fp16 x-rows each) + 8 mirror columns
constexpr uint32_t AU_V = 272;         // ring of 4 chunks x 32 columns (64 fp16 v-rows each) + 8 mirror columns
constexpr uint32_t AU_TMEM_COLS = 512;
constexpr int AU_SMEM_BYTES = 1024 + 8192 + 8192 + 256 + 4096 + 64;

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
    int dbg;                 // developer attribution toggles (BVG_ACT_DBG): 1 no cosine, 2 no global loads, 4 no stores,
                             // 8 one K-step per MMA chain (results are wrong with any of them set)
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

// Wait for up to three barriers at once (whole warp): the try_waits are issued back to back so their latencies overlap
// instead of adding up (an already-complete try_wait still costs ~90 cycles).
__device__ __forceinline__ void mbar_wait_warp3(uint64_t* b0, uint32_t p0, uint64_t* b1, uint32_t p1, uint64_t* b2, uint32_t p2) {
    uint32_t spins = 0;
    long long t0 = 0;
    bool d0 = false, d1 = (b1 == nullptr), d2 = false;
    for (;;) {
        if (!d0) d0 = mbar_try_wait(b0, p0);
        if (!d1) d1 = mbar_try_wait(b1, p1);
        if (!d2) d2 = mbar_try_wait(b2, p2);
        if (__all_sync(0xffffffffu, d0 && d1 && d2)) return;
        mbar_backoff(spins, t0);
    }
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

template <typename TOut>
struct AuOut;
template <>
struct AuOut<float> {  // two consecutive rows (stride elements apart) of one channel
    static __device__ __forceinline__ void store2(float* p, int stride, float2 v, bool ok0, bool ok1) {
        if (ok0) p[0] = v.x;
        if (ok1) p[stride] = v.y;
    }
};
template <>
struct AuOut<__nv_bfloat16> {
    static __device__ __forceinline__ void store2(__nv_bfloat16* p, int stride, float2 v, bool ok0, bool ok1) {
        const __nv_bfloat162 h = __float22bfloat162_rn(v);
        if (ok0) p[0] = __low2bfloat16(h);
        if (ok1) p[stride] = __high2bfloat16(h);
    }
};

// CS: channel count (row stride in elements) as a compile-time constant, 0 = runtime.  With a constant stride every row
// access of a lane is one instruction with an immediate offset; with a runtime stride each costs a 64-bit add on top
// (ncu r02f: 166 IMAD + 89 IADD3 of 1100 instructions per chunk were address arithmetic).
template <typename TIn, typename TOut, int CS>
__global__ void __launch_bounds__(AU_THREADS, 1)
act1d_umma_kernel(const __grid_constant__ ActUmmaParams p) {
    extern __shared__ uint8_t au_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(au_smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sBup = smem;          // [64 rows = u-time][64 K] fp16: K = 48 used
    uint8_t* sBdn = smem + 8192;   // 2 K-blocks x [32 rows = y-time][64 K] fp16: K = 80 used
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 16384);
    uint64_t* x_full = bars;          // [4] count 4 (the IO warps of the staging group)
    uint64_t* x_empty = bars + 4;     // [4] count 1 (tcgen05.commit)
    uint64_t* v_full = bars + 8;      // [4] count 4 (the snake warps)
    uint64_t* v_empty = bars + 12;    // [4] count 1 (tcgen05.commit)
    uint64_t* du_full = bars + 16;    // [2] count 1
    uint64_t* du_empty = bars + 18;   // [2] count 4
    uint64_t* dy_full = bars + 20;    // [2] count 1
    uint64_t* dy_empty = bars + 22;   // [2] count 4
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 24);
    // last valid activated sample per signal (replicate padding of the 2x signal on the right), handed from the snake warp
    // of one chunk to the snake warp of the next (the other group): [V ring slot][128 signals]
    float* vl_end = reinterpret_cast<float*>(smem + 16384 + 256);

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
    for (int e = threadIdx.x; e < 4 * 128; e += AU_THREADS) vl_end[e] = 0.f;
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
        // ------------------------------ UP issuer: U chunk j from X chunk j and the first half of X chunk j + 1 ----------
        const bool leader = elect_one();
        const uint32_t idesc_up = umma_idesc_f16_f32(128, 64);
        const uint64_t bup = umma_desc_sw128_kmajor(smem_u32(sBup));
        for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
            for (int j = 0; j <= nc; ++j) {
                const uint32_t G = G0 + j, slot = G & 3u, q = G & 1u;
                mbar_wait_warp3(&x_full[slot], (G >> 2) & 1u, j < nc ? &x_full[(G + 1) & 3u] : nullptr, ((G + 1) >> 2) & 1u,
                                &du_empty[q], ((G >> 1) & 1u) ^ 1u);
                tc_fence_after_sync();
                if (leader) {
                    const int ksteps = (p.dbg & 8) ? 1 : (j < nc ? 3 : 2);  // the last chunk needs rows of its own X chunk only
                    const uint32_t d = tmem_base + AU_DU + 64u * q, a = tmem_base + AU_X + 16u * slot;
                    for (int kk = 0; kk < ksteps; ++kk) umma_f16_ts(d, a + 8u * kk, bup + 2u * kk, idesc_up, kk > 0);
                    umma_commit(&du_full[q]);
                    umma_commit(&x_empty[slot]);
                }
                __syncwarp();
            }
            G0 += static_cast<uint32_t>(nc) + 1u;
        }
    } else if (warp == AU_THREADS / 32 - 1) {
        // ------------------------------ DOWN issuer: Y chunk s from V chunk s and the first 16 samples of V chunk s + 1 ---
        const bool leader = elect_one();
        const uint32_t idesc_dn = umma_idesc_f16_f32(128, 32);
        const uint64_t bdn = umma_desc_sw128_kmajor(smem_u32(sBdn));
        uint32_t dyc0 = 0u, dyc1 = 0u;  // uses of each Y accumulator so far
        for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
            for (int s = 0; s < nc; ++s) {
                const uint32_t G = G0 + s, slot = G & 3u, q = G & 1u;
                const uint32_t use = q ? dyc1 : dyc0;
                mbar_wait_warp3(&v_full[slot], (G >> 2) & 1u, &v_full[(G + 1) & 3u], ((G + 1) >> 2) & 1u, &dy_empty[q],
                                (use & 1u) ^ 1u);
                tc_fence_after_sync();
                if (leader) {
                    const uint32_t d = tmem_base + AU_DY + 32u * q, a = tmem_base + AU_V + 32u * slot;
                    for (int kk = 0; kk < ((p.dbg & 8) ? 1 : 5); ++kk)
                        umma_f16_ts(d, a + 8u * kk, kk < 4 ? bdn + 2u * kk : bdn + (4096u >> 4), idesc_dn, kk > 0);
                    umma_commit(&dy_full[q]);
                    umma_commit(&v_empty[slot]);
                    if (s == nc - 1) umma_commit(&v_empty[(G + 1) & 3u]);  // V chunk nc has no DOWN of its own
                }
                __syncwarp();
                if (q) ++dyc1; else ++dyc0;
            }
            G0 += static_cast<uint32_t>(nc) + 1u;
        }
    } else {
        // ------------------------------ compute warps ------------------------------
        // 16 warps = 2 groups (chunk parity) x 2 roles x 4 TMEM lane quarters.  Role 0 ("IO") loads a chunk's x rows, stages
        // them into the X ring and writes the output rows; role 1 ("snake") turns U into V.  Four warps per scheduler in
        // different phases (global loads, MUFU-bound snake, stores) keep the issue slots busy.
        const uint32_t widx = static_cast<uint32_t>(warp - 1);
        const uint32_t wg = widx >> 3;                                   // owns the chunks with G % 2 == wg
        const uint32_t role = (widx >> 2) & 1u;
        const uint32_t quarter = static_cast<uint32_t>(warp) & 3u;       // TMEM lane quarter this warp may access
        const uint32_t sig = quarter * 32u + static_cast<uint32_t>(lane);   // signal (TMEM lane) of this thread
        const uint32_t tl = tmem_base + ((quarter * 32u) << 16);
        const int cw_mask = (1 << p.cw_shift) - 1;
        const int T = static_cast<int>(p.T);          // the launcher guarantees T * C * 4 < 2^31: 32-bit row arithmetic
        const int Ci = CS > 0 ? CS : p.C;
        const int Lq = nc * AU_CHUNK;
        const int ntb = 128 >> p.cw_shift;
        uint32_t dyc = 0;     // uses of this group's Y accumulator

        for (uint32_t unit = blockIdx.x; unit < p.n_units; unit += gridDim.x) {
            const uint32_t cc = unit % p.n_cchunks;
            const uint32_t rest = unit / p.n_cchunks;
            const uint32_t b = rest / p.groups_per_b;
            const uint32_t g = rest - b * p.groups_per_b;
            const int ch = static_cast<int>(cc << p.cw_shift) + (static_cast<int>(sig) & cw_mask);
            // first output row of this lane; lanes whose block starts past the end are parked on the last block's rows
            // (their loads stay in bounds, their stores are masked)
            const int64_t y0l = (static_cast<int64_t>(g) * ntb + (sig >> p.cw_shift)) * Lq;
            const bool lane_live = y0l < T;
            const int y0 = lane_live ? static_cast<int>(y0l) : 0;
            const int i0 = ((G0 & 1u) == wg) ? 0 : 1;

            if (role == 0u) {
                // ================= IO warp: stage-in X chunk i, prefetch chunk i + 2, write Y chunk i - 4 =================
                const float cb = 0.5f / (expf(__ldg(p.beta + ch)) + 1e-9f);
                const float2 pcb = make_float2(cb, cb);
                const TIn* xb = static_cast<const TIn*>(p.x) + static_cast<int64_t>(b) * T * Ci + ch;
                TOut* yb = static_cast<TOut*>(p.y) + static_cast<int64_t>(b) * T * Ci + ch;
                float xr[32];  // prefetched x rows of the next chunk this warp stages (as fp32)
                auto load_chunk = [&](int i) {
                    const int t0 = y0 - 8 + i * AU_CHUNK;
                    const bool interior = (t0 >= 0) && (t0 + 31 < T);
                    if (p.dbg & 2) {
#pragma unroll
                        for (int r = 0; r < 32; ++r) xr[r] = static_cast<float>(r + lane) * 0.01f;
                    } else if (__all_sync(0xffffffffu, interior)) {
                        const TIn* pr = xb + static_cast<int64_t>(t0) * Ci;
#pragma unroll
                        for (int r = 0; r < 32; ++r) xr[r] = AuIO<TIn>::load(pr + r * Ci);
                    } else {  // replicate padding of x at the sequence ends (and in-bounds data for rows past the end)
#pragma unroll
                        for (int r = 0; r < 32; ++r) {
                            int t = t0 + r;
                            t = t < 0 ? 0 : (t >= T ? T - 1 : t);
                            xr[r] = AuIO<TIn>::load(xb + static_cast<int64_t>(t) * Ci);
                        }
                    }
                };
                // The register prefetch keeps one chunk (4 KB per warp, 32 KB per SM) in flight -- at ~1.2 us of loaded DRAM
                // latency that is 4 TB/s for the whole chip (measured: exactly that).  An L2 prefetch of the chunk after
                // next costs one instruction per row and no registers, and turns the later loads into L2 hits.
                auto prefetch_chunk = [&](int i) {
                    const int t0 = y0 - 8 + i * AU_CHUNK;
                    const bool interior = (t0 >= 0) && (t0 + 31 < T);
                    if (!(p.dbg & 2) && __all_sync(0xffffffffu, interior)) {
                        const TIn* pr = xb + static_cast<int64_t>(t0) * Ci;
#pragma unroll
                        for (int r = 0; r < 32; ++r) asm volatile("prefetch.global.L2 [%0];" ::"l"(pr + r * Ci));
                    }
                };
                if (i0 <= nc) load_chunk(i0);
                if (i0 + 2 <= nc) prefetch_chunk(i0 + 2);
                for (int i = i0; i <= nc + 4; i += 2) {
                    const uint32_t G = G0 + static_cast<uint32_t>(i);
                    if (i + 4 <= nc) prefetch_chunk(i + 4);
                    if (i <= nc) {
                        const uint32_t slot = G & 3u;
                        uint32_t pk[16];
#pragma unroll
                        for (int r = 0; r < 16; ++r) pk[r] = au_pack_half2(xr[2 * r], xr[2 * r + 1]);
                        mbar_wait_warp(&x_empty[slot], ((G >> 2) & 1u) ^ 1u);
                        tc_fence_after_sync();
                        tmem_st_x16(tl + AU_X + 16u * slot, pk);
                        if (slot == 0u) tmem_st_x8(tl + AU_X + 64u, pk);  // mirror: a slot-3 window runs on into it
                        tmem_st_wait();
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&x_full[slot]);
                    }
                    if (i + 2 <= nc) load_chunk(i + 2);
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
                        const int t0 = y0 + s * AU_CHUNK;
                        const int left_rows = lane_live ? T - t0 : 0;
                        const int n_ok = (p.dbg & 4) ? 0 : (left_rows >= 32 ? 32 : (left_rows < 0 ? 0 : left_rows));
                        TOut* pr = yb + static_cast<int64_t>(t0) * Ci;
                        if (__all_sync(0xffffffffu, n_ok == 32)) {
#pragma unroll
                            for (int r = 0; r < 32; r += 2) {
                                const float2 o = __fadd2_rn(make_float2(__uint_as_float(dv[r]), __uint_as_float(dv[r + 1])), pcb);
                                AuOut<TOut>::store2(pr + r * Ci, Ci, o, true, true);
                            }
                        } else {
#pragma unroll
                            for (int r = 0; r < 32; r += 2) {
                                const float2 o = __fadd2_rn(make_float2(__uint_as_float(dv[r]), __uint_as_float(dv[r + 1])), pcb);
                                AuOut<TOut>::store2(pr + r * Ci, Ci, o, r < n_ok, r + 1 < n_ok);
                            }
                        }
                    }
                }
            } else {
                // ================= snake warp: U chunk j (fp32, 64 columns) -> v = u - c cos(2 a u) -> fp16 -> V ring slot ===
                const float a2 = 2.0f * expf(__ldg(p.alpha + ch));
                const float cb = 0.5f / (expf(__ldg(p.beta + ch)) + 1e-9f);
                const float2 a2v = make_float2(a2, a2), ncb = make_float2(-cb, -cb);
                for (int j = i0; j <= nc; j += 2) {
                    const uint32_t Gj = G0 + static_cast<uint32_t>(j), slot = Gj & 3u;
                    mbar_wait_warp(&du_full[wg], (Gj >> 1) & 1u);
                    tc_fence_after_sync();
                    // column q of the chunk is 2x-sample 2*y0 + 64*j - 8 + q of this lane's sequence
                    const int n0 = 2 * y0 + 64 * j - 8;
                    const bool left = lane_live && (y0 == 0) && (j == 0);
                    const int qd = 2 * T - 1 - n0;
                    const int qlast = !lane_live ? 64 : (qd > 64 ? 64 : (qd < -1 ? -1 : qd));  // column of sample 2T-1
                    const bool edge = __any_sync(0xffffffffu, left || qlast <= 63);
                    float vl = 0.f;
                    if (edge && j > 0) {
                        // Replicate padding on the right: samples after 2T-1 repeat v[2T-1], which the previous chunk's
                        // snake warp (other group) may have produced: it travels through shared memory, ordered by that
                        // chunk's v_full barrier.
                        mbar_wait_warp(&v_full[(Gj - 1u) & 3u], ((Gj - 1u) >> 2) & 1u);
                        vl = vl_end[((Gj - 1u) & 3u) * 128u + sig];
                    }
                    mbar_wait_warp(&v_empty[slot], ((Gj >> 2) & 1u) ^ 1u);
                    tc_fence_after_sync();
                    // 32 columns at a time; each pair is activated and packed at once (registers: 96 per thread with 18
                    // warps -- five warps share one scheduler's register file)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t uv[32];
                        tmem_ld_32x32b_x32(tl + AU_DU + 64u * wg + 32u * h, uv);
                        tmem_ld_wait();
                        if (h == 1) {  // U is in registers: hand the accumulator back to the UP issuer
                            tc_fence_before_sync();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&du_empty[wg]);
                        }
                        uint32_t pk[16];
                        if (p.dbg & 1) {
#pragma unroll
                            for (int e = 0; e < 32; e += 2) pk[e >> 1] = au_pack_half2(__uint_as_float(uv[e]), __uint_as_float(uv[e + 1]));
                        } else if (!edge) {
#pragma unroll
                            for (int e = 0; e < 32; e += 2) {
                                const float2 u = make_float2(__uint_as_float(uv[e]), __uint_as_float(uv[e + 1]));
                                const float2 t = __fmul2_rn(u, a2v);
                                const float2 cs = make_float2(__cosf(t.x), __cosf(t.y));
                                const float2 r = __ffma2_rn(ncb, cs, u);
                                pk[e >> 1] = au_pack_half2(r.x, r.y);
                            }
                        } else {
                            float v0 = 0.f;  // v[0] of the sequence = column 8 of the first chunk (left replicate padding)
                            if (left && h == 0) {
                                const float u8 = __uint_as_float(uv[8]);
                                v0 = fmaf(-cb, __cosf(u8 * a2), u8);
                            }
#pragma unroll
                            for (int e = 0; e < 32; e += 2) {
                                const float2 u = make_float2(__uint_as_float(uv[e]), __uint_as_float(uv[e + 1]));
                                const float2 t = __fmul2_rn(u, a2v);
                                const float2 cs = make_float2(__cosf(t.x), __cosf(t.y));
                                float2 r = __ffma2_rn(ncb, cs, u);
                                if (left && h == 0 && e < 8) r = make_float2(v0, v0);
                                if (32 * h + e > qlast) r.x = vl; else vl = r.x;      // samples after 2T-1 repeat v[2T-1]
                                if (32 * h + e + 1 > qlast) r.y = vl; else vl = r.y;
                                pk[e >> 1] = au_pack_half2(r.x, r.y);
                            }
                        }
                        tmem_st_x16(tl + AU_V + 32u * slot + 16u * h, pk);
                        if (h == 0 && slot == 0u) tmem_st_x8(tl + AU_V + 128u, pk);  // mirror (see X ring)
                    }
                    if (edge) vl_end[slot * 128u + sig] = vl;
                    tmem_st_wait();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&v_full[slot]);
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
template <typename TIn, typename TOut, int CS>
static int act_umma_attrs() {
    static std::atomic<bool> done[64];
    int dev = 0;
    BVG_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && done[dev].load(std::memory_order_acquire)) return BVG_OK;
    BVG_CUDA(cudaFuncSetAttribute(act1d_umma_kernel<TIn, TOut, CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, AU_SMEM_BYTES));
    if (dev >= 0 && dev < 64) done[dev].store(true, std::memory_order_release);
    return BVG_OK;
}

template <typename TIn, typename TOut, int CS>
static int launch_act_umma_cs(const ActUmmaParams& p, int grid, cudaStream_t st) {
    const int rc = act_umma_attrs<TIn, TOut, CS>();
    if (rc != BVG_OK) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(grid));
    cfg.blockDim = dim3(AU_THREADS);
    cfg.dynamicSmemBytes = AU_SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    BVG_CUDA(cudaLaunchKernelEx(&cfg, act1d_umma_kernel<TIn, TOut, CS>, p));
    BVG_LAUNCH_CHECK("act1d_umma_kernel");
    return BVG_OK;
}

// Returns BVG_OK after launching, a positive value when the shape does not suit this kernel (the caller then uses the
// SIMT kernel), a negative status on error.  WIDTHS: the generator's stage widths get a compile-time row stride.
template <typename TIn, typename TOut, bool WIDTHS>
static int launch_act_umma(const void* x, void* y, const float* alpha, const float* beta, int64_t B, int64_t T, int64_t C,
                           cudaStream_t st) {
    // 32-bit row arithmetic inside the kernel
    if (C % 8 != 0 || C > (1 << 16) || T < 1 || T >= (int64_t(1) << 28) || T * C >= (int64_t(1) << 29)) return 1;
    ActUmmaParams p{};
    p.x = x; p.y = y; p.alpha = alpha; p.beta = beta; p.T = T; p.C = static_cast<int>(C);
    static const int dbg_env = []() { const char* e = getenv("BVG_ACT_DBG"); return e ? atoi(e) : 0; }();
    p.dbg = dbg_env;
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
    const int grid = static_cast<int>(units < sms ? units : sms);
    if (WIDTHS) {
        switch (C) {
            case 24: return launch_act_umma_cs<TIn, TOut, 24>(p, grid, st);
            case 48: return launch_act_umma_cs<TIn, TOut, 48>(p, grid, st);
            case 96: return launch_act_umma_cs<TIn, TOut, 96>(p, grid, st);
            case 192: return launch_act_umma_cs<TIn, TOut, 192>(p, grid, st);
            case 384: return launch_act_umma_cs<TIn, TOut, 384>(p, grid, st);
            case 768: return launch_act_umma_cs<TIn, TOut, 768>(p, grid, st);
            default: break;
        }
    }
    return launch_act_umma_cs<TIn, TOut, 0>(p, grid, st);
}

// Opt-in (BVG_ACT_UMMA=1).  Measured on B200 at the cfg2 stage shapes (profiles/r02_act1d_simt_vs_umma.jsonl): level with
// the SIMT kernel for fp32 input at C = 96..384 (126-127 vs 129-130 us), 4-6 % behind for bf16 input, and well behind at
// C = 24 (188 vs 131 us: 48-byte rows make this kernel's per-lane global accesses scalar and strided).  The empty
// pipeline (no cosine, no global traffic) already takes 79 us -- every U and Y element crosses the tensor-memory read
// port -- so the SIMT kernel stays the default; DESIGN.md 4 has the attribution.
bool act_umma_requested(int64_t C) {
    (void)C;
    static const bool v = []() {
        const char* e = getenv("BVG_ACT_UMMA");
        return e != nullptr && e[0] == '1';
    }();
    return v;
}

int activation1d_nlc_umma(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha, const float* beta, int64_t B,
                          int64_t T, int64_t C, cudaStream_t st) {
    if (x_dtype == BVG_F32 && y_dtype == BVG_BF16)
        return launch_act_umma<float, __nv_bfloat16, true>(x, y, alpha, beta, B, T, C, st);
    if (x_dtype == BVG_BF16 && y_dtype == BVG_BF16)
        return launch_act_umma<__nv_bfloat16, __nv_bfloat16, true>(x, y, alpha, beta, B, T, C, st);
    if (x_dtype == BVG_F32 && y_dtype == BVG_F32)  // activation_post only: one width, runtime stride
        return launch_act_umma<float, float, false>(x, y, alpha, beta, B, T, C, st);
    return 1;
}

template <typename TIn, typename TOut>
static int act_umma_attrs_widths() {
    int rc = act_umma_attrs<TIn, TOut, 0>();
    if (rc == BVG_OK) rc = act_umma_attrs<TIn, TOut, 24>();
    if (rc == BVG_OK) rc = act_umma_attrs<TIn, TOut, 48>();
    if (rc == BVG_OK) rc = act_umma_attrs<TIn, TOut, 96>();
    if (rc == BVG_OK) rc = act_umma_attrs<TIn, TOut, 192>();
    if (rc == BVG_OK) rc = act_umma_attrs<TIn, TOut, 384>();
    if (rc == BVG_OK) rc = act_umma_attrs<TIn, TOut, 768>();
    return rc;
}

int activation1d_umma_init() {
    int rc = act_umma_attrs_widths<float, __nv_bfloat16>();
    if (rc == BVG_OK) rc = act_umma_attrs_widths<__nv_bfloat16, __nv_bfloat16>();
    if (rc == BVG_OK) rc = act_umma_attrs<float, float, 0>();
    return rc;
}

}  // namespace bvg
