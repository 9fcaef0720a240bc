// Dilated Conv1d / polyphase ConvTranspose1d as an implicit GEMM on tcgen05 tensor cores (sm_100a).
//
// Replaces cuDNN/ATen behind nn.Conv1d (reference bigvgan.py:58-66,75-83,285-287) and nn.ConvTranspose1d
// (bigvgan.py:305-313).  Layout is channels-last ("nlc"): activations [B, T, Cin] bf16, packed weights
// [taps][Nout][Cin] bf16, so that
//   D[M = 128 time rows, N = output channels] += A[M, K = 64 input channels] * B[N, K]^T   per tap, per K-block,
// with both operands K-major in shared memory under the 128-byte swizzle.
//
// One halo'd TMA tile serves every tap: the A tile holds rows t0+base .. t0+base+(k-1)*d+127 of the input (TMA
// zero-fills rows outside [0,T) -> the conv's zero padding, per batch item because the tensor map is 3-D
// [Cin, T, B]); tap j's operand is the SAME shared memory addressed through a UMMA descriptor whose start address
// is advanced by j*d rows (j*d*128 bytes).  The swizzle XOR is a function of absolute smem address bits, so the
// row-shifted descriptor reads the row-shifted operand (verified on B200, profiles/r01_probe_umma_row_shift.log).
//
// A transposed convolution is the same kernel: phase r of the stride-u output is an ordinary 2-tap conv (K/u taps)
// whose output "channels" are n = r*Cout + co of a [B, Tin, u*Cout] tensor -- which is bit-for-bit the
// [B, u*Tin, Cout] result.  Each N tile lies inside one phase and uses that phase's tap base, so no zero-stuffed
// tensor and no wasted MACs.
//
// Warp-specialised persistent CTAs (1 per SM):  warp 0 = TMA producer, warp 1 = UMMA issuer (+ TMEM allocator),
// warps 2..5 = epilogue (tcgen05.ld -> +bias +residual, *scale, +accum -> swizzled smem -> TMA tensor stores),
// warp 6 = TMA loader of the fp32 residual / accum chunks the epilogue consumes.
// Pipelines: A ring (one halo tile per K-block), B ring (one weight tile per K-block x tap), two TMEM accumulators
// so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "bvg_common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

#include <atomic>
#include <cstdlib>

#ifndef BVG_CONV_SMEM_KB_DEFAULT
#define BVG_CONV_SMEM_KB_DEFAULT 227
#endif

namespace bvg {

constexpr int UMMA_BLOCK_M = 128;
constexpr int UMMA_BLOCK_K = 64;  // bf16 elements per 128-byte swizzled row
constexpr int UMMA_THREADS = 224;   // 7 warps: TMA producer, UMMA issuer, 4 epilogue, residual loader
constexpr int UMMA_THREADS2 = 384;  // dual pipeline: + a second UMMA issuer (warp 7) and 4 more epilogue warps (8..11)
constexpr int UMMA_MAX_STAGES = 8;
constexpr int EPI_CHUNK = 32;                  // accumulator columns per epilogue step (= one 128-byte fp32 row segment)
constexpr int EPI_LD_TILE_BYTES = 128 * 128;   // one residual / accum chunk: 128 rows x 32 fp32, 128B-swizzled
constexpr int EPI_ST32_TILE_BYTES = 32 * 128;  // per-warp fp32 output chunk: 32 rows x 32 fp32, 128B-swizzled
constexpr int EPI_ST16_TILE_BYTES = 32 * 64;   // per-warp bf16 output chunk: 32 rows x 32 bf16, 64B-swizzled

struct ConvUmmaParams {
    const float* bias;      // [cout_phase] or null
    const float* cond;      // per-utterance bias [B, cout_phase] or null (v1 generator conditioning)
    int cluster;            // 1: clusters of two CTAs share every streamed weight stage by TMA multicast (see launcher)
    const float* residual;  // [B,T,N] fp32 or null
    const float* accum;     // [B,T,N] fp32 or null
    float* y_f32;           // [B,T,N] or null
    __nv_bfloat16* y_bf16;  // [B,T,N] or null
    int64_t T;              // rows per batch item
    int B;
    int Cin;
    int N;           // total output channels (u*Cout for a transposed conv)
    int cout_phase;  // channels per phase (= N for a plain conv)
    int taps;        // taps per phase
    int dil;
    int tap_base[BVG_MAX_PHASES];  // input row offset of tap 0 for each phase
    int n_tile;                    // UMMA N (multiple of 16, <= 256)
    int tiles_per_phase;           // ceil(cout_phase / n_tile)
    int num_phases;
    int64_t m_tiles;  // ceil(T / 128)
    int64_t total_tiles;
    int a_stages, b_stages;
    uint32_t a_stage_bytes, b_stage_bytes;
    // pair mode (mp = 2, weight-streaming layers): a tile is 256 rows = two 128-row sub-tiles that share every weight
    // tile of the B ring (two accumulators per TMEM buffer), which halves the L2 -> shared-memory weight traffic; an A
    // stage then holds two overlapping halo boxes 16 KB apart (a_box_bytes each)
    int mp;
    uint32_t a_box_bytes;
    uint32_t tmem_cols;
    // pipeline groups (1 or 2): group g owns the CTA's tiles i with i % ng == g, its own UMMA issuer warp, 4 epilogue
    // warps, 2 TMEM accumulators and output staging tiles; the TMA producer / loader warps feed both in tile order.
    // For small-channel layers a tile is ~1 us of single-warp instruction issue in each role, so two groups double it.
    int ng;
    // weight-stationary mode: every (n tile, K block, tap) weight tile is loaded into shared memory ONCE per CTA and
    // only the activation halo tiles stream (small-channel layers, where re-streaming weights per 128-row tile made
    // the kernel TMA-latency bound)
    int stationary;
    uint32_t w_bytes;  // bytes of the stationary weight region
    // epilogue I/O is all TMA: residual / accum chunks are prefetched by a loader warp, outputs leave through
    // per-warp swizzled smem tiles and bulk tensor stores
    int has_res, has_acc, has_y32, has_y16;
    int ld_stages;     // residual/accum chunk ring depth (<= UMMA_MAX_STAGES)
    unsigned long long* trace;  // developer timeline buffer (bvg_debug_conv_trace) or null
    int trace_n;
    uint32_t bias_smem;  // bytes of the zero-padded bias copy in shared memory (0: read the bias from global memory)
    float out_scale;
    // dynamic tile scheduling: a zero-initialised counter in global memory (one per launch, owned by the caller's plan).  A CTA
    // starts on tile blockIdx.x and takes tile gridDim.x + atomicAdd(counter, 1) next, so the kernel ends on the mean CTA
    // instead of the slowest static share.  nullptr: static striding (tile = blockIdx.x + i * gridDim.x).
    unsigned int* tile_ctr;
};

// Developer timeline: event e of pipeline role `slot` at the role's i-th tile (CTA 0 only).  Compiled in only with
// -DBVG_CONV_TRACE (tools/build_variant.sh): the shipped library carries no trace pointer and no global state for it.
__device__ __forceinline__ void trace_ev(const ConvUmmaParams& p, int slot, uint32_t i) {
#ifdef BVG_CONV_TRACE
    if (p.trace != nullptr && blockIdx.x == 0 && i < static_cast<uint32_t>(p.trace_n))
        p.trace[static_cast<size_t>(slot) * p.trace_n + i] = clock64();
#else
    (void)p; (void)slot; (void)i;
#endif
}

// Tile index -> (n tile, m tile, batch item), 32-bit arithmetic (total_tiles < 2^31 is checked by the launcher).
struct TileCoord {
    int nt, mt, b;
};
__device__ __forceinline__ TileCoord decode_tile(uint32_t tile, uint32_t n_tiles, uint32_t m_tiles) {
    TileCoord c;
    uint32_t rest = tile;
    c.nt = 0;
    if (n_tiles > 1) {
        rest = tile / n_tiles;
        c.nt = static_cast<int>(tile - rest * n_tiles);
    }
    const uint32_t b = rest / m_tiles;
    c.mt = static_cast<int>(rest - b * m_tiles);
    c.b = static_cast<int>(b);
    return c;
}
// A role's walk over its tiles (first, first + step, ...): the two divisions are paid once, every further tile is a
// mixed-radix add with carries (a tile is only ~1000 cycles for the small-channel layers, two 32-bit divisions per
// tile and role were a measurable part of it).
struct TileWalker {
    TileCoord c, d;
    int n_tiles, m_tiles;
    __device__ __forceinline__ TileWalker(uint32_t first, uint32_t step, uint32_t n_tiles_, uint32_t m_tiles_)
        : c(decode_tile(first, n_tiles_, m_tiles_)), d(decode_tile(step, n_tiles_, m_tiles_)),
          n_tiles(static_cast<int>(n_tiles_)), m_tiles(static_cast<int>(m_tiles_)) {}
    __device__ __forceinline__ void next() {
        c.nt += d.nt;
        c.mt += d.mt;
        c.b += d.b;
        if (c.nt >= n_tiles) { c.nt -= n_tiles; ++c.mt; }
        if (c.mt >= m_tiles) { c.mt -= m_tiles; ++c.b; }
    }
};

// The tile sequence of a CTA as its consumer roles (UMMA issuer, epilogue warps, residual loader) see it.
// Static: tile = first + i * step, decoded incrementally.  Dynamic (ConvUmmaParams::tile_ctr): the producer warp fetches
// tile indices from the global counter, decodes them once and publishes {tile, n tile, m tile, batch} in a small
// shared-memory queue; entry i belongs to pipeline group i % ng, the queue ends with one TQ_END entry per group.
constexpr int TQ_SLOTS = 8;  // multiple of the number of pipeline groups
constexpr uint32_t TQ_END = 0xffffffffu;
struct TileFeed {
    const uint4* q;
    uint64_t* full;
    uint64_t* empty;
    uint32_t qi, qstride;  // dynamic: queue cursor of this role
    uint32_t tile, step, total;
    bool dyn, started;
    TileWalker tw;
    __device__ __forceinline__ TileFeed(bool dyn_, const uint4* q_, uint64_t* full_, uint64_t* empty_, uint32_t first_entry,
                                        uint32_t entry_stride, uint32_t first, uint32_t step_, uint32_t total_,
                                        uint32_t n_tiles, uint32_t m_tiles)
        : q(q_), full(full_), empty(empty_), qi(first_entry), qstride(entry_stride), tile(first), step(step_), total(total_),
          dyn(dyn_), started(false), tw(first, step_, n_tiles, m_tiles) {}
    // whole-warp call; false when the sequence is exhausted
    __device__ __forceinline__ bool next(TileCoord& c, int lane) {
        if (!dyn) {
            if (started) { tile += step; tw.next(); }
            started = true;
            if (tile >= total) return false;
            c = tw.c;
            return true;
        }
        const uint32_t slot = qi % TQ_SLOTS;
        mbar_wait_warp(&full[slot], (qi / TQ_SLOTS) & 1u);
        const uint4 e = q[slot];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        qi += qstride;
        if (e.x == TQ_END) return false;
        tile = e.x;
        c.nt = static_cast<int>(e.y);
        c.mt = static_cast<int>(e.z);
        c.b = static_cast<int>(e.w);
        return true;
    }
};

// tcgen05.mma with the 64-bit shared-memory descriptors assembled from a 32-bit cursor (start-address field, i.e.
// byte address >> 4) and the constant upper half: advancing an operand is one 32-bit uniform add.
__device__ __forceinline__ void umma_f16_lo(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi,
                                            uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %5, 0;\n"
        "mov.b64 da, {%1, %3};\n"
        "mov.b64 db, {%2, %3};\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
        "}\n"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

// The UMMA issue loop of warp 1.  For the small-channel layers one tap is only 16-48 tensor-pipe cycles per UMMA, so
// this single-warp instruction stream IS the critical path (ncu r01c: 200 cycles per tap before this rewrite): it is
// specialised on the weight mode, keeps every cursor in 32-bit (uniform) registers and touches no kernel parameter
// inside the tap loop.
template <bool STAT>
__device__ __forceinline__ void umma_issue_loop(const ConvUmmaParams& p, uint32_t tmem_base, uint32_t sA_d, uint32_t sB_d,
                                                uint64_t* a_full, uint64_t* a_empty, uint64_t* b_full, uint64_t* b_empty,
                                                uint64_t* acc_full, uint64_t* acc_empty, int k_blocks, int n_tiles,
                                                int group, uint32_t cta_first, uint32_t cta_step, const uint4* tq,
                                                uint64_t* tq_full, uint64_t* tq_empty, uint32_t m_tiles) {
    const uint32_t idesc = umma_idesc_bf16_f32(UMMA_BLOCK_M, p.n_tile);
    const uint32_t desc_hi = static_cast<uint32_t>(umma_desc_sw128_kmajor(0) >> 32);
    const uint32_t tap_step = static_cast<uint32_t>(p.dil) * 8u;  // dil rows * 128 B >> 4
    const uint32_t b_step = p.b_stage_bytes >> 4;
    const uint32_t a_stage_d = p.a_stage_bytes >> 4;
    const int taps = p.taps;
    const int a_stages = p.a_stages / p.ng, b_stages = p.b_stages;  // A stages per pipeline group
    const int last_steps = ((p.Cin - (k_blocks - 1) * UMMA_BLOCK_K) + 15) / 16;
    const uint32_t n_tile = p.n_tile;
    const uint32_t w_tile_d = static_cast<uint32_t>(k_blocks * taps) * b_step;  // stationary weights of one n tile
    const uint32_t total = static_cast<uint32_t>(p.total_tiles);
    const bool leader = elect_one();
    // each group consumes its own sub-ring of A stages (slots [group*a_stages, (group+1)*a_stages)), so a slot is
    // always refilled for the same consumer and the parity protocol stays unambiguous
    const int ng = p.ng;
    const uint32_t mp = static_cast<uint32_t>(p.mp);
    acc_full += 2 * group;
    acc_empty += 2 * group;
    a_full += group * a_stages;
    a_empty += group * a_stages;
    sA_d += static_cast<uint32_t>(group * a_stages) * a_stage_d;
    int sa = 0, sb = 0;
    uint32_t acc = 0, pa = 0, pb = 0, pacc = 0;
    if (STAT) mbar_wait_warp(&b_full[0], 0);
    uint32_t it = 0;
    const bool cluster = p.cluster != 0;
    const int lane = static_cast<int>(threadIdx.x & 31u);
    TileFeed tf(p.tile_ctr != nullptr, tq, tq_full, tq_empty, group, ng, cta_first + group * cta_step, ng * cta_step, total,
                n_tiles, m_tiles);
    TileCoord tcd;
    for (; tf.next(tcd, lane); ++it) {
        mbar_wait_warp(&acc_empty[acc], pacc ^ 1);
        tc_fence_after_sync();
        if (leader) trace_ev(p, 1 + 6 * group, it);
        const uint32_t d_tmem = tmem_base + ((2 * group + acc) * mp) * n_tile;
        uint32_t accumulate = 0;
        uint32_t w_d = sB_d;  // stationary weight cursor
        if (STAT && n_tiles > 1) w_d += static_cast<uint32_t>(tcd.nt) * w_tile_d;
        for (int kb = 0; kb < k_blocks; ++kb) {
            const int k_steps = (kb + 1 == k_blocks) ? last_steps : 4;
            const bool k2 = k_steps > 1, k3 = k_steps > 2, k4 = k_steps > 3;
            mbar_wait_warp(&a_full[sa], pa);
            tc_fence_after_sync();
            if (leader && kb == 0) trace_ev(p, 2 + 6 * group, it);
            uint32_t a_d = sA_d + sa * a_stage_d;
#pragma unroll 1
            for (int j = 0; j < taps; ++j) {
                uint32_t b_d;
                if (STAT) {
                    b_d = w_d;
                    w_d += b_step;
                } else {
                    mbar_wait_warp(&b_full[sb], pb);
                    tc_fence_after_sync();
                    b_d = sB_d + sb * b_step;
                }
                if (leader) {
                    umma_f16_lo(d_tmem, a_d, b_d, desc_hi, idesc, accumulate);
                    if (k2) umma_f16_lo(d_tmem, a_d + 2, b_d + 2, desc_hi, idesc, 1);
                    if (k3) umma_f16_lo(d_tmem, a_d + 4, b_d + 4, desc_hi, idesc, 1);
                    if (k4) umma_f16_lo(d_tmem, a_d + 6, b_d + 6, desc_hi, idesc, 1);
                    if (mp == 2) {  // second 128-row sub-tile: rows +128 (16 KB further), next accumulator, same weights
                        const uint32_t a2 = a_d + 1024u, d2 = d_tmem + n_tile;
                        umma_f16_lo(d2, a2, b_d, desc_hi, idesc, accumulate);
                        if (k2) umma_f16_lo(d2, a2 + 2, b_d + 2, desc_hi, idesc, 1);
                        if (k3) umma_f16_lo(d2, a2 + 4, b_d + 4, desc_hi, idesc, 1);
                        if (k4) umma_f16_lo(d2, a2 + 6, b_d + 6, desc_hi, idesc, 1);
                    }
                    if (!STAT) {
                        if (cluster) umma_commit_mc(&b_empty[sb], 3);  // both CTAs refill this slot together
                        else umma_commit(&b_empty[sb]);
                    }
                }
                accumulate = 1;
                a_d += tap_step;
                if (!STAT) {
                    if (++sb == b_stages) { sb = 0; pb ^= 1; }
                }
            }
            if (leader) umma_commit(&a_empty[sa]);
            if (++sa == a_stages) { sa = 0; pa ^= 1; }
        }
        if (leader) umma_commit(&acc_full[acc]);
        if (leader) trace_ev(p, 3 + 6 * group, it);
        acc ^= 1;
        if (acc == 0) pacc ^= 1;
    }
}

// -DBVG_CONV_MAXNREG=96 (developer A/B, tools/build_variant.sh): cap the registers so that a 384-thread conv CTA and a
// 288-thread Activation1d CTA (96 registers) fit one SM's register file together.
#ifdef BVG_CONV_MAXNREG
__global__ void __maxnreg__(BVG_CONV_MAXNREG)
#else
__global__ void __launch_bounds__(UMMA_THREADS2, 1)
#endif
conv_umma_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_res, const __grid_constant__ CUtensorMap map_acc,
                 const __grid_constant__ CUtensorMap map_y32, const __grid_constant__ CUtensorMap map_y16,
                 const __grid_constant__ ConvUmmaParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    // carve-up: [A ring][B ring | stationary weights][residual/accum chunk ring][fp32 out tiles][bf16 out tiles]
    //           [barriers][tmem ptr]
    uint8_t* sA = smem;
    uint8_t* sB = sA + p.a_stages * p.a_stage_bytes;
    uint8_t* sLd = sB + (p.stationary ? p.w_bytes : p.b_stages * p.b_stage_bytes);
    const int n_ld = p.has_res + p.has_acc;
    uint8_t* sSt32 = sLd + p.ld_stages * n_ld * EPI_LD_TILE_BYTES;
    uint8_t* sSt16 = sSt32 + (p.has_y32 ? p.ng * 4 * 2 * EPI_ST32_TILE_BYTES : 0);
    float* sBias = reinterpret_cast<float*>(sSt16 + (p.has_y16 ? p.ng * 4 * 2 * EPI_ST16_TILE_BYTES : 0));
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sBias) + p.bias_smem);
    uint64_t* a_full = bars;
    uint64_t* a_empty = bars + UMMA_MAX_STAGES;
    uint64_t* b_full = bars + 2 * UMMA_MAX_STAGES;   // [0] doubles as "weights resident" in stationary mode
    uint64_t* b_empty = bars + 3 * UMMA_MAX_STAGES;
    uint64_t* acc_full = bars + 4 * UMMA_MAX_STAGES;  // [2 per group]
    uint64_t* acc_empty = acc_full + 4;               // [2 per group]
    uint64_t* ld_full = acc_empty + 4;                // [UMMA_MAX_STAGES]
    uint64_t* ld_empty = ld_full + UMMA_MAX_STAGES;   // [UMMA_MAX_STAGES]
    uint64_t* tq_full = ld_empty + UMMA_MAX_STAGES;   // [TQ_SLOTS] tile queue (dynamic scheduling)
    uint64_t* tq_empty = tq_full + TQ_SLOTS;          // [TQ_SLOTS]
    uint4* tq = reinterpret_cast<uint4*>(tq_empty + TQ_SLOTS);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tq + TQ_SLOTS);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

#ifdef BVG_CONV_TRACE
    if (p.trace != nullptr && threadIdx.x == 0 && gridDim.x <= static_cast<unsigned>(p.trace_n)) {
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        p.trace[13 * static_cast<size_t>(p.trace_n) + blockIdx.x] = ns;  // CTA start (ns)
    }
#endif
    // PDL: the next kernel of the stream may become resident as soon as this CTA's resources free up; everything this
    // kernel does before griddep_wait() below (barrier init, TMEM allocation, bias copy, stationary weight loads) only
    // touches constants, so it overlaps the tail of the previous kernel.
    griddep_launch_dependents();
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_w);
        for (int i = 0; i < UMMA_MAX_STAGES; ++i) {
            mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1);
            mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], p.cluster ? 2 : 1);
            mbar_init(&ld_full[i], 1); mbar_init(&ld_empty[i], 4);
        }
        for (int i = 0; i < 4; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
        // a queue entry is read by its group's UMMA issuer, its four epilogue warps and (if it runs) the chunk loader
        for (int i = 0; i < TQ_SLOTS; ++i) {
            mbar_init(&tq_full[i], 1);
            mbar_init(&tq_empty[i], 5 + ((p.has_res + p.has_acc) > 0 ? 1 : 0));
        }
        fence_barrier_init();
    }
    if (p.bias_smem) {
        const int nb = static_cast<int>(p.bias_smem / 4);
        for (int i = threadIdx.x; i < nb; i += blockDim.x) sBias[i] = i < p.cout_phase ? __ldg(p.bias + i) : 0.0f;
    }
    if (warp == 1) tmem_alloc(tmem_ptr, p.tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    if (p.cluster) cluster_barrier_sync();  // the peer's barriers exist before any multicast / remote arrive targets them
    const uint32_t tmem_base = *tmem_ptr;

    const int k_blocks = (p.Cin + UMMA_BLOCK_K - 1) / UMMA_BLOCK_K;
    const int n_tiles = p.num_phases * p.tiles_per_phase;
    const uint32_t total_tiles = static_cast<uint32_t>(p.total_tiles);
    const uint32_t m_tiles = static_cast<uint32_t>(p.m_tiles);
    // Cluster mode: the tile index space is one of row-tile PAIRS (p.m_tiles = pairs, set by the launcher); the two CTAs
    // of a cluster walk the same pairs and take row tile 2 * pair + rank, so they need the same weight stages in the
    // same order.  An odd row-tile count leaves a ghost tile past T: its loads are zero-filled, its stores clipped.
    const uint32_t crank = p.cluster ? (blockIdx.x & 1u) : 0u;
    const uint32_t cta_first = p.cluster ? (blockIdx.x >> 1) : blockIdx.x;
    const uint32_t cta_step = p.cluster ? (gridDim.x >> 1) : gridDim.x;
    const int mt_mul = p.cluster ? 2 : 1;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        // The whole warp runs the control flow (so every value stays warp-uniform and lives in uniform registers);
        // one elected lane issues the TMA operations.
        if (p.stationary) {
            if (elect_one()) mbar_arrive_expect_tx(&b_full[0], p.w_bytes);
            __syncwarp();
            for (int nt = 0; nt < n_tiles; ++nt) {
                const int n0 = (nt / p.tiles_per_phase) * p.cout_phase + (nt % p.tiles_per_phase) * p.n_tile;
                for (int kb = 0; kb < k_blocks; ++kb)
                    for (int j = 0; j < p.taps; ++j)
                        if (elect_one())
                            tma_load_3d_hint(sB + ((nt * k_blocks + kb) * p.taps + j) * p.b_stage_bytes, &map_w, &b_full[0],
                                             kb * UMMA_BLOCK_K, n0, j, kPolicyEvictLast);
            }
        }
        griddep_wait();  // activations are produced by the previous kernels of the stream
        // per-group cursors into the groups' A sub-rings (tiles alternate between the groups)
        const int a_per = p.a_stages / p.ng;
        int sa = 0, sa_other = 0, sb = 0, g = 0;
        uint32_t pa = 0, pa_other = 0, pb = 0, it = 0;
        TileWalker tw(cta_first, cta_step, n_tiles, m_tiles);
        const uint32_t b_half = p.b_stage_bytes >> 1;
        const bool dyn = p.tile_ctr != nullptr;
        uint32_t tile = cta_first, nxt = 0, qi = 0;
        int ends = 0;  // dynamic: end markers published (one per pipeline group)
        for (;; ++it) {
            TileCoord tc;
            if (dyn) {
                const bool valid = tile < total_tiles;
                // the next tile index is fetched now: the atomic's round trip overlaps this tile's loads
                if (valid && lane == 0) nxt = atomicAdd(p.tile_ctr, 1u) + gridDim.x;
                tc = decode_tile(valid ? tile : 0u, n_tiles, m_tiles);
                const uint32_t slot = qi % TQ_SLOTS;
                mbar_wait_warp(&tq_empty[slot], ((qi / TQ_SLOTS) & 1u) ^ 1u);
                if (lane == 0) {
                    tq[slot] = make_uint4(valid ? tile : TQ_END, static_cast<uint32_t>(tc.nt), static_cast<uint32_t>(tc.mt),
                                          static_cast<uint32_t>(tc.b));
                    mbar_arrive(&tq_full[slot]);
                }
                __syncwarp();
                ++qi;
                if (!valid) {
                    if (++ends == p.ng) break;
                    continue;
                }
            } else {
                if (tile >= total_tiles) break;
                tc = tw.c;
            }
            const int nt = tc.nt, mt = tc.mt * mt_mul + static_cast<int>(crank), b = tc.b;
            const int slot0 = g * a_per;
            const int phase = nt / p.tiles_per_phase;
            const int n0 = phase * p.cout_phase + (nt % p.tiles_per_phase) * p.n_tile;
            const int row0 = mt * UMMA_BLOCK_M * p.mp + p.tap_base[phase];
            for (int kb = 0; kb < k_blocks; ++kb) {
                mbar_wait_warp(&a_empty[slot0 + sa], pa ^ 1);
                if (kb == 0 && lane == 0) trace_ev(p, 0, it);
                if (elect_one()) {
                    mbar_arrive_expect_tx(&a_full[slot0 + sa], p.mp * p.a_box_bytes);
                    uint8_t* dst = sA + (slot0 + sa) * p.a_stage_bytes;
                    tma_load_3d(dst, &map_x, &a_full[slot0 + sa], kb * UMMA_BLOCK_K, row0, b);
                    // pair mode: the second halo box starts 128 rows (16 KB) further; the overlapping rows are
                    // simply written twice with the same data
                    if (p.mp == 2)
                        tma_load_3d(dst + 16384, &map_x, &a_full[slot0 + sa], kb * UMMA_BLOCK_K, row0 + UMMA_BLOCK_M, b);
                }
                if (++sa == a_per) { sa = 0; pa ^= 1; }
                if (!p.stationary) {
                    for (int j = 0; j < p.taps; ++j) {
                        mbar_wait_warp(&b_empty[sb], pb ^ 1);
                        if (elect_one()) {
                            mbar_arrive_expect_tx(&b_full[sb], p.b_stage_bytes);
                            if (p.cluster)  // this CTA fetches its half of the stage and multicasts it to both
                                tma_load_3d_mc_hint(sB + sb * p.b_stage_bytes + crank * b_half, &map_w, &b_full[sb],
                                                    kb * UMMA_BLOCK_K, n0 + static_cast<int>(crank) * (p.n_tile >> 1), j, 3,
                                                    kPolicyEvictLast);
                            else
                                tma_load_3d_hint(sB + sb * p.b_stage_bytes, &map_w, &b_full[sb], kb * UMMA_BLOCK_K, n0, j,
                                                 kPolicyEvictLast);
                        }
                        if (++sb == p.b_stages) { sb = 0; pb ^= 1; }
                    }
                }
            }
            if (p.ng == 2) {  // next tile belongs to the other group
                int ts = sa; sa = sa_other; sa_other = ts;
                uint32_t tp = pa; pa = pa_other; pa_other = tp;
                g ^= 1;
            }
            if (dyn) {
                tile = __shfl_sync(0xffffffffu, nxt, 0);
            } else {
                tile += cta_step;
                tw.next();
            }
        }
    } else if (warp == 1 || warp == 7) {
        // ------------------------------ UMMA issuer (one per pipeline group) ------------------------------
        const int group = warp == 7 ? 1 : 0;
        const uint32_t sA_d = smem_u32(sA) >> 4, sB_d = smem_u32(sB) >> 4;
        if (p.stationary)
            umma_issue_loop<true>(p, tmem_base, sA_d, sB_d, a_full, a_empty, b_full, b_empty, acc_full, acc_empty, k_blocks,
                                  n_tiles, group, cta_first, cta_step, tq, tq_full, tq_empty, m_tiles);
        else
            umma_issue_loop<false>(p, tmem_base, sA_d, sB_d, a_full, a_empty, b_full, b_empty, acc_full, acc_empty,
                                   k_blocks, n_tiles, group, cta_first, cta_step, tq, tq_full, tq_empty, m_tiles);
    } else if (warp == 6) {
        // ------------------------------ residual / accum chunk loader ------------------------------
        if (n_ld > 0) {
            griddep_wait();  // residual / running-sum tensors are produced by previous kernels
            // like the A ring, the chunk ring is split into one sub-ring per pipeline group: a slot is then always
            // refilled for the same consumer warps, which keeps the parity protocol unambiguous
            const int ld_per = p.ld_stages / p.ng;
            int s = 0, s_other = 0, g = 0;
            uint32_t ps = 0, ps_other = 0;
            TileFeed tf(p.tile_ctr != nullptr, tq, tq_full, tq_empty, 0u, 1u, cta_first, cta_step, total_tiles, n_tiles, m_tiles);
            TileCoord tc;
            while (tf.next(tc, lane)) {
                const int nt = tc.nt, t00 = (tc.mt * mt_mul + static_cast<int>(crank)) * UMMA_BLOCK_M * p.mp, b = tc.b;
                const int phase = nt / p.tiles_per_phase;
                const int c0 = (nt % p.tiles_per_phase) * p.n_tile;
                const int slot0 = g * ld_per;
                for (int sc = 0; sc < p.mp * p.n_tile; sc += EPI_CHUNK) {  // (sub-tile, chunk) in the epilogue's order
                    const int sub = sc >= p.n_tile ? 1 : 0;
                    const int c = sc - sub * p.n_tile;
                    const int t0 = t00 + sub * UMMA_BLOCK_M;
                    const int slot = slot0 + s;
                    mbar_wait_warp(&ld_empty[slot], ps ^ 1);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&ld_full[slot], n_ld * EPI_LD_TILE_BYTES);
                        uint8_t* dst = sLd + slot * n_ld * EPI_LD_TILE_BYTES;
                        if (p.has_res) tma_load_4d(dst, &map_res, &ld_full[slot], c0 + c, phase, t0, b);
                        if (p.has_acc)
                            tma_load_4d(dst + p.has_res * EPI_LD_TILE_BYTES, &map_acc, &ld_full[slot], c0 + c, phase, t0, b);
                    }
                    if (++s == ld_per) { s = 0; ps ^= 1; }
                }
                if (p.ng == 2) {  // next tile belongs to the other group
                    int ts = s; s = s_other; s_other = ts;
                    uint32_t tp = ps; ps = ps_other; ps_other = tp;
                    g ^= 1;
                }
            }
        }
    } else {
        // ------------------------------ epilogue (warps 2..5; group 1: warps 8..11) ------------------------------
        // thread = accumulator row.  tcgen05.ld -> +bias (+residual chunk from smem) * scale (+accum chunk) -> swizzled
        // per-warp smem tiles -> TMA bulk tensor stores.  No direct global loads/stores except the bias broadcast.
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;          // row inside the 128-row tile
        const uint32_t sw128 = static_cast<uint32_t>(row & 7);
        const uint32_t sw64 = static_cast<uint32_t>((lane >> 1) & 3);
        const int group = warp >= 8 ? 1 : 0;
        uint8_t* st32 = sSt32 + (group * 4 + q) * 2 * EPI_ST32_TILE_BYTES;
        uint8_t* st16 = sSt16 + (group * 4 + q) * 2 * EPI_ST16_TILE_BYTES;
        acc_full += 2 * group;
        acc_empty += 2 * group;
        const int ld_per = p.ld_stages / p.ng;  // this group's sub-ring of residual / accum chunk slots
        const uint8_t* sLdG = sLd + group * ld_per * n_ld * EPI_LD_TILE_BYTES;
        ld_full += group * ld_per;
        ld_empty += group * ld_per;
        int acc = 0, ls = 0, ob = 0;
        uint32_t pacc = 0, pls = 0;
        const bool tracer = (q == 0 && lane == 0);
        uint32_t it = 0;
        TileFeed tf(p.tile_ctr != nullptr, tq, tq_full, tq_empty, static_cast<uint32_t>(group), static_cast<uint32_t>(p.ng),
                    cta_first + group * cta_step, p.ng * cta_step, total_tiles, n_tiles, m_tiles);
        TileCoord tc;
        for (; tf.next(tc, lane); ++it) {
            const int nt = tc.nt, t00 = (tc.mt * mt_mul + static_cast<int>(crank)) * UMMA_BLOCK_M * p.mp, b = tc.b;
            const int phase = nt / p.tiles_per_phase;
            const int c0 = (nt % p.tiles_per_phase) * p.n_tile;  // channel offset inside the phase

            mbar_wait_warp(&acc_full[acc], pacc);
            tc_fence_after_sync();
            if (tracer) trace_ev(p, 4 + 6 * group, it);
            for (int sub = 0; sub < p.mp; ++sub) {
            const int t0 = t00 + sub * UMMA_BLOCK_M;
            const uint32_t t_addr = tmem_base + ((static_cast<uint32_t>(q) * 32u) << 16) +
                                    ((2 * group + acc) * p.mp + sub) * p.n_tile;
            for (int c = 0; c < p.n_tile; c += EPI_CHUNK) {
                const int width = min(EPI_CHUNK, p.n_tile - c);  // 32 or 16
                float o[32];
                {
                    uint32_t v[32];
                    if (width == EPI_CHUNK) {
                        tmem_ld_32x32b_x32(t_addr + c, v);
                    } else {
                        uint32_t h[16];
                        tmem_ld_32x32b_x16(t_addr + c, h);
#pragma unroll
                        for (int e = 0; e < 16; ++e) { v[e] = h[e]; v[16 + e] = 0u; }
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 32; ++e) o[e] = __uint_as_float(v[e]);
                }
                if (c + EPI_CHUNK >= p.n_tile && sub + 1 == p.mp) {  // last TMEM read of this tile: hand the accumulator back early
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[acc]);
                    if (tracer) trace_ev(p, 5 + 6 * group, it);
                }
                if (p.bias) {
                    const int chb = c0 + c;
                    if (p.bias_smem) {  // zero-padded copy in shared memory: broadcast reads, no bounds checks
                        const float4* bs = reinterpret_cast<const float4*>(sBias + chb);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 bv = bs[j];
                            o[4 * j] += bv.x; o[4 * j + 1] += bv.y; o[4 * j + 2] += bv.z; o[4 * j + 3] += bv.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (chb + 4 * j < p.cout_phase) {
                                const float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + chb + 4 * j));
                                o[4 * j] += bv.x; o[4 * j + 1] += bv.y; o[4 * j + 2] += bv.z; o[4 * j + 3] += bv.w;
                            }
                        }
                    }
                }
#ifndef BVG_NO_COND  // developer A/B only (tools/build_variant.sh): the epilogue without the conditioning branch
                if (p.cond) {  // warp-uniform addresses: one broadcast transaction per float4
                    const int chb = c0 + c;
                    const float* cv = p.cond + static_cast<size_t>(b) * p.cout_phase + chb;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (chb + 4 * j < p.cout_phase) {
                            const float4 bv = __ldg(reinterpret_cast<const float4*>(cv + 4 * j));
                            o[4 * j] += bv.x; o[4 * j + 1] += bv.y; o[4 * j + 2] += bv.z; o[4 * j + 3] += bv.w;
                        }
                    }
                }
#endif
                if (n_ld > 0) {
                    mbar_wait_warp(&ld_full[ls], pls);
                    const uint8_t* ld = sLdG + ls * n_ld * EPI_LD_TILE_BYTES + row * 128;
                    if (p.has_res) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 rv = *reinterpret_cast<const float4*>(ld + ((j ^ sw128) << 4));
                            o[4 * j] += rv.x; o[4 * j + 1] += rv.y; o[4 * j + 2] += rv.z; o[4 * j + 3] += rv.w;
                        }
                    }
                    if (p.out_scale != 1.0f) {
#pragma unroll
                        for (int e = 0; e < 32; ++e) o[e] *= p.out_scale;
                    }
                    if (p.has_acc) {
                        const uint8_t* la = ld + p.has_res * EPI_LD_TILE_BYTES;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 rv = *reinterpret_cast<const float4*>(la + ((j ^ sw128) << 4));
                            o[4 * j] += rv.x; o[4 * j + 1] += rv.y; o[4 * j + 2] += rv.z; o[4 * j + 3] += rv.w;
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ld_empty[ls]);
                    if (++ls == ld_per) { ls = 0; pls ^= 1; }
                } else if (p.out_scale != 1.0f) {
#pragma unroll
                    for (int e = 0; e < 32; ++e) o[e] *= p.out_scale;
                }
                // the smem tile written two chunks ago must have been read by its TMA store
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
                if (p.has_y32) {
                    uint8_t* dst = st32 + ob * EPI_ST32_TILE_BYTES + lane * 128;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<float4*>(dst + ((j ^ (lane & 7)) << 4)) =
                            make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                }
                if (p.has_y16) {
                    uint8_t* dst = st16 + ob * EPI_ST16_TILE_BYTES + lane * 64;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        __nv_bfloat162 h0 = __floats2bfloat162_rn(o[8 * j], o[8 * j + 1]);
                        __nv_bfloat162 h1 = __floats2bfloat162_rn(o[8 * j + 2], o[8 * j + 3]);
                        __nv_bfloat162 h2 = __floats2bfloat162_rn(o[8 * j + 4], o[8 * j + 5]);
                        __nv_bfloat162 h3 = __floats2bfloat162_rn(o[8 * j + 6], o[8 * j + 7]);
                        *reinterpret_cast<uint4*>(dst + ((j ^ sw64) << 4)) =
                            make_uint4(*reinterpret_cast<unsigned int*>(&h0), *reinterpret_cast<unsigned int*>(&h1),
                                       *reinterpret_cast<unsigned int*>(&h2), *reinterpret_cast<unsigned int*>(&h3));
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (p.has_y32) tma_store_4d(&map_y32, st32 + ob * EPI_ST32_TILE_BYTES, c0 + c, phase, t0 + q * 32, b);
                    if (p.has_y16) tma_store_4d(&map_y16, st16 + ob * EPI_ST16_TILE_BYTES, c0 + c, phase, t0 + q * 32, b);
                    tma_store_commit();
                }
                ob ^= 1;
            }
            }  // sub
            if (++acc == 2) { acc = 0; pacc ^= 1; }
            if (tracer) trace_ev(p, 6 + 6 * group, it);
        }
        if (lane == 0) tma_store_wait<0>();
    }

    tc_fence_before_sync();
    __syncthreads();
#ifdef BVG_CONV_TRACE
    if (p.trace != nullptr && threadIdx.x == 0 && gridDim.x <= static_cast<unsigned>(p.trace_n)) {
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        p.trace[14 * static_cast<size_t>(p.trace_n) + blockIdx.x] = ns;  // CTA end (ns)
    }
#endif
    if (p.cluster) cluster_barrier_sync();  // the peer's last multicast commits land in this CTA's barriers before it exits
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, p.tmem_cols);
    }
}


// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
static int encode_3d_bf16(CUtensorMap* map, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t b0,
                          uint32_t b1, uint32_t b2, const char* what) {
    EncodeTiledFn enc = get_encode_fn();
    BVG_REQUIRE(enc != nullptr, BVG_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t dims[3] = {d0, d1, d2};
    cuuint64_t strides[2] = {d0 * 2, d0 * d1 * 2};
    cuuint32_t box[3] = {b0, b1, b2};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BVG_REQUIRE(r == CUDA_SUCCESS, BVG_ERR_CUDA,
                "cuTensorMapEncodeTiled(%s) failed with CUresult %d (dims %llu x %llu x %llu, box %u x %u x %u)", what,
                (int)r, (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2, b0, b1, b2);
    return BVG_OK;
}

// [B,T,N] tensor viewed as [cout_phase, phases, T, B] so that column clipping is per phase.
static int encode_io_4d(CUtensorMap* map, const void* ptr, bool bf16, uint64_t cp, uint64_t phases, uint64_t T, uint64_t B,
                        uint32_t box_rows, const char* what) {
    EncodeTiledFn enc = get_encode_fn();
    BVG_REQUIRE(enc != nullptr, BVG_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    const uint64_t es = bf16 ? 2 : 4;
    const uint64_t N = cp * phases;
    cuuint64_t dims[4] = {cp, phases, T, B};
    cuuint64_t strides[3] = {cp * es, N * es, T * N * es};
    cuuint32_t box[4] = {32, 1, box_rows, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                     const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     bf16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BVG_REQUIRE(r == CUDA_SUCCESS, BVG_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, (int)r);
    return BVG_OK;
}

#ifdef BVG_CONV_TRACE
static unsigned long long* g_conv_trace = nullptr;
static int g_conv_trace_n = 0;
#endif

// Shared-memory budget of one CTA.  Below the 227 KB maximum an Activation1d CTA of another AMP block can be resident
// on the same SM (forward.cu runs the blocks on separate streams), trading pipeline depth for FP32/tensor overlap.
static int64_t conv_smem_budget() {
    static const int64_t v = []() -> int64_t {
        const char* e = getenv("BVG_CONV_SMEM_KB");
        int kb = e ? atoi(e) : 0;
        if (kb < 96 || kb > 227) kb = BVG_CONV_SMEM_KB_DEFAULT;
        return (int64_t)kb * 1024;
    }();
    return v;
}

// Opt in to the large dynamic shared memory carve-out once per device (idempotent calls; the flag is atomic so that
// concurrent host threads race benignly: at worst both set the same attributes).  Also reachable through bvg_init().
int conv_umma_init() {
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    BVG_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        BVG_CUDA(cudaFuncSetAttribute(conv_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        BVG_CUDA(cudaFuncSetAttribute(conv_umma_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                      cudaSharedmemCarveoutMaxShared));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    return BVG_OK;
}

// Generic launcher: plain conv (num_phases = 1) or polyphase transposed conv.
int conv_umma_launch(const void* x, const void* w_packed, const float* bias, const float* residual, const float* accum,
                     float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t N, int64_t cout_phase,
                     int taps, int dil, const int* tap_base, int num_phases, float out_scale, cudaStream_t st,
                     const float* cond, unsigned int* tile_ctr) {
    BVG_REQUIRE(x && w_packed && (y_f32 || y_bf16), BVG_ERR_INVALID, "conv_umma: null pointer");
    BVG_REQUIRE(Cin % 8 == 0 && cout_phase % 8 == 0 && N == cout_phase * num_phases, BVG_ERR_UNSUPPORTED,
                "conv_umma: Cin=%lld and Cout=%lld must be multiples of 8", (long long)Cin, (long long)cout_phase);
    BVG_REQUIRE(taps >= 1 && taps <= 11 && dil >= 1 && (taps - 1) * dil <= 120, BVG_ERR_UNSUPPORTED,
                "conv_umma: taps=%d dilation=%d outside the supported range", taps, dil);
    BVG_REQUIRE(num_phases >= 1 && num_phases <= BVG_MAX_PHASES, BVG_ERR_UNSUPPORTED, "conv_umma: %d phases", num_phases);
    BVG_REQUIRE(T < (int64_t(1) << 31) - 256 && B < 65536, BVG_ERR_UNSUPPORTED, "conv_umma: T or B too large");
    BVG_REQUIRE((reinterpret_cast<uintptr_t>(x) % 16 == 0) && (reinterpret_cast<uintptr_t>(w_packed) % 16 == 0),
                BVG_ERR_INVALID, "conv_umma: x / w must be 16-byte aligned");
    if (B == 0 || T == 0) return BVG_OK;

    ConvUmmaParams p{};
    p.bias = bias; p.cond = cond; p.residual = residual; p.accum = accum; p.y_f32 = y_f32;
    p.y_bf16 = static_cast<__nv_bfloat16*>(y_bf16);
    p.T = T; p.B = (int)B; p.Cin = (int)Cin; p.N = (int)N; p.cout_phase = (int)cout_phase;
    p.taps = taps; p.dil = dil; p.num_phases = num_phases; p.out_scale = out_scale;
    for (int r = 0; r < num_phases; ++r) p.tap_base[r] = tap_base[r];
    // N tile: cover a phase with as few equal tiles as possible, <= 256 wide, multiple of 16
    {
        const int cp = (int)cout_phase;
        int nt = (cp + 255) / 256;
        // several tiles per phase: 32-column granularity so that epilogue chunks never straddle tiles
        int tile = (int)round_up(ceil_div(cp, nt), nt > 1 ? 32 : 16);
        // Small problems (a single utterance, a time shard): with the widest tile the grid would leave most SMs idle,
        // so narrow the N tile until the tiles fill the machine.  Cost model: waves x max(tile, 96) -- below ~96
        // columns a tile stops getting cheaper (the A operand reads and the UMMA issue rate dominate).
        const int64_t mt_all = ceil_div(T, UMMA_BLOCK_M) * B * num_phases;
        const int sms_l = device_sm_count();
        if (mt_all * ceil_div(cp, tile) < sms_l && cp >= 64) {
            int64_t best_cost = ceil_div(mt_all * ceil_div(cp, tile), sms_l) * (tile > 96 ? tile : 96);
            for (int nt_c = nt + 1; nt_c <= cp / 32; ++nt_c) {
                const int tile_c = (int)round_up(ceil_div(cp, nt_c), 32);
                const int64_t tiles_c = mt_all * ceil_div(cp, tile_c);
                const int64_t cost = ceil_div(tiles_c, sms_l) * (tile_c > 96 ? tile_c : 96);
                if (cost < best_cost) { best_cost = cost; tile = tile_c; }
            }
        }
        // Pair mode for the weight-streaming mid-channel layers (C = 96, 192 with k >= 7): their B ring is bound by
        // L2 -> shared-memory weight traffic (every 128-row tile re-reads all k * Cin * Cout weights), so two 128-row
        // sub-tiles share each weight tile.  Needs four accumulators in TMEM (n_tile <= 128) and whole 32-column chunks.
        p.mp = 1;
        static const int pair_env = []() { const char* e = getenv("BVG_CONV_PAIR"); return e ? atoi(e) : 1; }();
        const int kb_l = (int)ceil_div(Cin, UMMA_BLOCK_K);
        if (pair_env && num_phases == 1 && Cin <= 192 && cp <= 128 &&
            (int64_t)ceil_div(cp, tile) * kb_l * taps * round_up((int64_t)tile * 128, 1024) > 112 * 1024) {
            // only where the N tile stays the whole layer: halving a 192-wide tile to fit four accumulators re-reads the
            // activations and was measured slower (stage 2: 192 -> 210 us for k = 7), gpurun r01x
            const int t2 = (cp <= 128 && cp % 32 == 0) ? cp : 0;
            if (t2 > 0 && ceil_div(T, 2 * UMMA_BLOCK_M) * B * (cp / t2) >= 2 * (int64_t)sms_l) {
                tile = t2;
                p.mp = 2;
            }
        }
        p.n_tile = tile;
        p.tiles_per_phase = (int)ceil_div(cp, tile);
    }
    p.m_tiles = ceil_div(T, UMMA_BLOCK_M * p.mp);
    p.total_tiles = p.m_tiles * B * p.num_phases * p.tiles_per_phase;
    BVG_REQUIRE(p.total_tiles < (int64_t(1) << 31) - 1024, BVG_ERR_UNSUPPORTED, "conv_umma: %lld tiles exceed the 32-bit tile index",
                (long long)p.total_tiles);
    const int a_rows = (int)round_up(UMMA_BLOCK_M + (taps - 1) * dil, 8);
    p.a_box_bytes = (uint32_t)round_up((int64_t)a_rows * 128, 1024);
    p.a_stage_bytes = p.mp == 2 ? 16384u + p.a_box_bytes : p.a_box_bytes;
    p.b_stage_bytes = (uint32_t)round_up((int64_t)p.n_tile * 128, 1024);
    p.ng = 1;
#ifdef BVG_CONV_TRACE
    p.trace = g_conv_trace;
    p.trace_n = g_conv_trace_n;
#endif
    p.has_res = residual != nullptr; p.has_acc = accum != nullptr;
    p.has_y32 = y_f32 != nullptr; p.has_y16 = y_bf16 != nullptr;
    const int n_ld = p.has_res + p.has_acc;
    BVG_REQUIRE(reinterpret_cast<uintptr_t>(residual) % 16 == 0 && reinterpret_cast<uintptr_t>(accum) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(y_f32) % 16 == 0 && reinterpret_cast<uintptr_t>(y_bf16) % 16 == 0,
                BVG_ERR_INVALID, "conv_umma: residual / accum / outputs must be 16-byte aligned");
    // shared memory plan: 227 KB - alignment slack - barriers - epilogue tiles is what the operand pipelines get
    const int k_blocks = (int)ceil_div(Cin, UMMA_BLOCK_K);
    const int64_t w_all = (int64_t)p.num_phases * p.tiles_per_phase * k_blocks * taps * p.b_stage_bytes;
    // dual pipeline when the weights are small enough to be stationary and four accumulators fit TMEM
    static const int ng_env = []() { const char* e = getenv("BVG_CONV_GROUPS"); return e ? atoi(e) : 0; }();
    if (ng_env != 1 && 4 * p.n_tile <= 512 && w_all <= 96 * 1024) p.ng = 2;
    // bias copy in shared memory (zero-padded by one epilogue chunk so the chunk reads need no bounds checks)
    p.bias_smem = (bias != nullptr && cout_phase <= 768) ? (uint32_t)round_up((cout_phase + EPI_CHUNK) * 4, 128) : 0u;
    bool planned = false;
    uint32_t st_bytes = 0;
    for (int ng_try = p.ng; ng_try >= 1 && !planned; --ng_try) {
        p.ng = ng_try;
        st_bytes = p.ng * ((p.has_y32 ? 8 * EPI_ST32_TILE_BYTES : 0) + (p.has_y16 ? 8 * EPI_ST16_TILE_BYTES : 0));
        const uint32_t fixed = 1024 + 1024 + st_bytes + p.bias_smem;
        // best effort: plan inside the (possibly reduced) budget first, fall back to the full 227 KB when a layer needs it
        const int64_t budgets[2] = {conv_smem_budget(), 227 * 1024};
        for (int attempt = 0; attempt < 2 && !planned; ++attempt) {
            const int64_t avail = budgets[attempt] - (int64_t)fixed;
            const int64_t ld_stage_bytes = (int64_t)n_ld * EPI_LD_TILE_BYTES;
            if (w_all <= 112 * 1024 && avail - w_all >= 2 * (int64_t)p.a_stage_bytes + ld_stage_bytes) {
                // weight-stationary: all weight tiles resident; the rest is split between a deep activation ring and a deep
                // residual/accum ring (both hide DRAM latency, a tile is only ~1 us of work here)
                p.stationary = 1;
                p.w_bytes = (uint32_t)w_all;
                p.b_stages = 0;
                const int64_t rem = avail - w_all;
                // both rings are split evenly between the pipeline groups, so their depths are multiples of ng
                const int ld_min = n_ld ? p.ng : 0;
                for (int ld = n_ld ? UMMA_MAX_STAGES : 0;; ld -= p.ng) {
                    int64_t as = (rem - ld * ld_stage_bytes) / p.a_stage_bytes;
                    if (as > UMMA_MAX_STAGES) as = UMMA_MAX_STAGES;
                    as -= as % p.ng;
                    if (as >= 4 || (ld <= ld_min && as >= 2)) { p.a_stages = (int)as; p.ld_stages = ld; planned = true; break; }
                    if (ld <= ld_min) break;
                }
            }
            if (!planned) {
                p.stationary = 0;
                p.w_bytes = 0;
                for (int ld = n_ld ? 4 : 0;; --ld) {
                    const int64_t rem = avail - ld * ld_stage_bytes;
                    int as = 3;
                    int64_t bs = (rem - 3 * (int64_t)p.a_stage_bytes) / p.b_stage_bytes;
                    if (bs < 4) { as = 2; bs = (rem - 2 * (int64_t)p.a_stage_bytes) / p.b_stage_bytes; }
                    if (bs > UMMA_MAX_STAGES) bs = UMMA_MAX_STAGES;
                    if (bs >= 4 || (ld <= 1 && bs >= 2)) { p.a_stages = as; p.b_stages = (int)bs; p.ld_stages = ld; planned = true; break; }
                    if (ld <= 1) break;
                }
            }
        }
        // the dual pipeline needs resident weights and at least two A stages per group (measured: with one stage per group
        // it is slower than the single pipeline with its deeper rings, gpurun_out/r01o)
        if (planned && p.ng == 2 && (!p.stationary || p.a_stages < 4)) planned = false;
    }
    BVG_REQUIRE(planned, BVG_ERR_UNSUPPORTED, "conv_umma: tile does not fit shared memory");
    BVG_REQUIRE(p.ng == 1 || p.stationary, BVG_ERR_UNSUPPORTED, "conv_umma: dual pipeline planned without stationary weights");
    BVG_REQUIRE(p.mp == 1 || (!p.stationary && p.ng == 1 && 4 * p.n_tile <= 512 && p.n_tile % 32 == 0), BVG_ERR_UNSUPPORTED,
                "conv_umma: pair mode planned for an unsuitable layer");
    {
        uint32_t tm = 32;
        while (tm < (uint32_t)(2 * p.ng * p.mp * p.n_tile)) tm <<= 1;
        p.tmem_cols = tm;
    }
    const uint32_t smem_bytes = p.a_stages * p.a_stage_bytes + (p.stationary ? p.w_bytes : p.b_stages * p.b_stage_bytes) +
                                p.ld_stages * n_ld * EPI_LD_TILE_BYTES + st_bytes + p.bias_smem + 1024 + 1024;
    BVG_REQUIRE(smem_bytes <= 227 * 1024, BVG_ERR_UNSUPPORTED, "conv_umma: %u bytes of shared memory needed", smem_bytes);

    // Cluster mode for SMALL-M weight-streaming layers (one utterance at stages 0-1: 14-54 row tiles of C = 768 / 384).
    // Every row tile streams the whole weight slice of its N tile from L2; with ~112 CTAs doing that at once the layer is
    // bound by aggregate L2 bandwidth (99 cycles per MMA against the 56-66-cycle issue floor, tools/conv_first_tile.py).
    // Two CTAs that work on neighbouring row tiles of the same N tile form a cluster: each fetches half of every weight
    // stage and multicasts it into both shared memories, halving the L2 reads at an unchanged CTA count.
    // BVG_CONV_CLUSTER=1 switches it on.
    static const int cluster_env = []() { const char* e = getenv("BVG_CONV_CLUSTER"); return e ? atoi(e) : 0; }();
    p.cluster = 0;
    // BVG_CONV_CLUSTER=2: every weight-streaming single-pipeline layer, whatever its size (experiment)
    // BVG_CONV_CLUSTER=3 (experiment): also the pair-mode layers (C = 96 class with k >= 7), whose 264 KB of weights per
    // 256-row tile make them L2-bound (gpurun r02t: 96 cycles per MMA against a 56-66-cycle issue floor)
    if (cluster_env && !p.stationary && p.ng == 1 && (p.mp == 1 || cluster_env >= 3) && num_phases == 1 &&
        p.n_tile % 16 == 0 && p.m_tiles >= 2 &&
        (cluster_env >= 2 || (Cin >= 256 && p.total_tiles <= 2 * (int64_t)device_sm_count()))) {
        p.cluster = 1;
        p.m_tiles = ceil_div(p.m_tiles, 2);  // the kernel walks row-tile PAIRS
        p.total_tiles = p.m_tiles * B * p.num_phases * p.tiles_per_phase;
    }

    // Dynamic tile scheduling when the caller's plan provides a zeroed counter (forward.cu) and BVG_CONV_DYNAMIC=1 asks for it.
    // Off by default: measured neutral to slightly slower on every cfg2 layer (profiles/r02_conv_dynamic_ab.log: conv sum
    // 23.2 vs 22.9 ms) -- the 5-9 % spread of CTA durations is smaller than the cost of the queue hand-off and the atomics.
    // Not in cluster mode (the two CTAs of a cluster must walk the same tiles) and pointless when every CTA has one tile.
    static const int dyn_env = []() { const char* e = getenv("BVG_CONV_DYNAMIC"); return e ? atoi(e) : 0; }();
    p.tile_ctr = (dyn_env && !p.cluster && p.total_tiles > device_sm_count()) ? tile_ctr : nullptr;

    CUtensorMap map_x, map_w, map_res, map_acc, map_y32, map_y16;
    // the A box is a_stage_bytes/128 rows so that expect_tx matches the bytes TMA writes
    const uint32_t a_box_rows = p.a_box_bytes / 128;
    BVG_REQUIRE(a_box_rows <= 256, BVG_ERR_UNSUPPORTED, "conv_umma: halo tile of %u rows exceeds the TMA box limit", a_box_rows);
    int rc = encode_3d_bf16(&map_x, x, (uint64_t)Cin, (uint64_t)T, (uint64_t)B, UMMA_BLOCK_K, a_box_rows, 1, "x");
    if (rc != BVG_OK) return rc;
    const uint32_t b_box_rows = p.b_stage_bytes / 128 / (p.cluster ? 2 : 1);  // cluster mode: each CTA fetches half a stage
    rc = encode_3d_bf16(&map_w, w_packed, (uint64_t)Cin, (uint64_t)N, (uint64_t)taps, UMMA_BLOCK_K, b_box_rows, 1, "w");
    if (rc != BVG_OK) return rc;

    {
        const int rc_init = conv_umma_init();
        if (rc_init != BVG_OK) return rc_init;
    }
    const int sms = device_sm_count();
    const int grid = p.cluster ? 2 * (int)(p.total_tiles < sms / 2 ? p.total_tiles : sms / 2)
                               : (int)(p.total_tiles < sms ? p.total_tiles : sms);
    map_res = map_acc = map_y32 = map_y16 = map_x;  // placeholders for unused maps
    if (p.has_res && (rc = encode_io_4d(&map_res, residual, false, cout_phase, num_phases, T, B, 128, "residual")) != BVG_OK) return rc;
    if (p.has_acc && (rc = encode_io_4d(&map_acc, accum, false, cout_phase, num_phases, T, B, 128, "accum")) != BVG_OK) return rc;
    if (p.has_y32 && (rc = encode_io_4d(&map_y32, y_f32, false, cout_phase, num_phases, T, B, 32, "y_f32")) != BVG_OK) return rc;
    if (p.has_y16 && (rc = encode_io_4d(&map_y16, y_bf16, true, cout_phase, num_phases, T, B, 32, "y_bf16")) != BVG_OK) return rc;
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(p.ng == 2 ? UMMA_THREADS2 : UMMA_THREADS);
        cfg.dynamicSmemBytes = smem_bytes;
        cfg.stream = st;
        cudaLaunchAttribute attr[2];
        int na = 0;
        if (pdl_enabled()) {
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        if (p.cluster) {
            attr[na].id = cudaLaunchAttributeClusterDimension;
            attr[na].val.clusterDim.x = 2;
            attr[na].val.clusterDim.y = 1;
            attr[na].val.clusterDim.z = 1;
            ++na;
        }
        cfg.attrs = attr;
        cfg.numAttrs = na;
        BVG_CUDA(cudaLaunchKernelEx(&cfg, conv_umma_kernel, map_x, map_w, map_res, map_acc, map_y32, map_y16, p));
    }
    BVG_LAUNCH_CHECK("conv_umma_kernel");
    return BVG_OK;
}

}  // namespace bvg

// Developer timeline switch.  Only a library built with -DBVG_CONV_TRACE keeps the pointer (process-global, developer
// builds only); the shipped build exports the symbol as a no-op that reports BVG_ERR_UNSUPPORTED.
extern "C" int bvg_debug_conv_trace(void* buf, int n_tiles) {
#ifdef BVG_CONV_TRACE
    bvg::g_conv_trace = static_cast<unsigned long long*>(buf);
    bvg::g_conv_trace_n = buf ? n_tiles : 0;
    return BVG_OK;
#else
    (void)n_tiles;
    if (buf == nullptr) return BVG_OK;
    return bvg::set_error(BVG_ERR_UNSUPPORTED, "bvg_debug_conv_trace: this libbvg.so was built without -DBVG_CONV_TRACE");
#endif
}
