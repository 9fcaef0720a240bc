// Hardware probe (developer tool, not part of libbvg.so): how many cycles does one tcgen05.mma (M = 128, K = 16, bf16,
// both operands in shared memory, SWIZZLE_128B K-major) cost for small N, issued back to back by one thread?  The
// small-channel convs (C <= 96) are bound by this number times taps * ceil(C / 16) per 128-row tile.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I. -o probe_umma_rate probe_umma_rate.cu && ./probe_umma_rate
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include "sm100_ptx.cuh"

using namespace bvg;

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);     \
            exit(2);                                                                            \
        }                                                                                       \
    } while (0)

__device__ __forceinline__ void mma_lo(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                       uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %5, 0;\n"
        "mov.b64 da, {%1, %3};\n"
        "mov.b64 db, {%2, %3};\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
        "}\n" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

// mode 0: every MMA reads the same A / B tiles (pure issue + pipe rate)
// mode 1: conv-like walk: 4 K steps (+32 B) per tap, A advanced by `dil` rows per tap, B by one weight tile per tap
// mode 2: like 1, two accumulators alternating per MMA (two independent dependency chains in TMEM)
// mode 3: like 0 with TWO issuing threads (n_mma / 2 each, own accumulator): cycles per MMA of the pair
__global__ void __launch_bounds__(128, 1)
rate_kernel(long long* out, int N, int n_mma, int mode, int dil) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;               // 256 rows x 128 B
    uint8_t* sB = smem + 256 * 128;   // 11 taps x 96 rows x 128 B at most = 132 KB
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 11 * 96 * 128);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar + 1);
    for (int i = threadIdx.x; i < (256 * 128 + 11 * 96 * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    if (threadIdx.x / 32 == 0) tmem_alloc(tmem_ptr, 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;
    if (mode == 3) {
        // two issuing threads (warps 0 and 1) on one SM, separate accumulators and mbarriers: is the ~56-cycle cost per
        // instruction a property of the issuing thread or of the SM's tensor pipe?
        __shared__ uint64_t bar2[2];
        __shared__ long long tt[2][2];
        if (threadIdx.x == 0) {
            mbar_init(&bar2[0], 1);
            mbar_init(&bar2[1], 1);
            fence_barrier_init();
        }
        __syncthreads();
        const int w = threadIdx.x / 32;
        if (w < 2 && threadIdx.x % 32 == 0) {
            const uint32_t idesc = umma_idesc_bf16_f32(128, N);
            const uint32_t desc_hi = static_cast<uint32_t>(umma_desc_sw128_kmajor(0) >> 32);
            const uint32_t a0 = ((smem_u32(sA) & 0x3FFFF) >> 4) + w * 1024u, b0 = (smem_u32(sB) & 0x3FFFF) >> 4;
            const long long t0 = clock64();
#pragma unroll 8
            for (int i = 0; i < n_mma / 2; ++i) mma_lo(tmem_base + w * 256u, a0, b0, desc_hi, idesc, i > 0);
            umma_commit(&bar2[w]);
            mbar_wait(&bar2[w], 0);
            tt[w][0] = t0;
            tt[w][1] = clock64();
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            const long long lo = tt[0][0] < tt[1][0] ? tt[0][0] : tt[1][0], hi = tt[0][1] > tt[1][1] ? tt[0][1] : tt[1][1];
            out[0] = hi - lo;
            out[1] = hi - lo;
        }
    } else if (threadIdx.x == 0) {
        const uint32_t idesc = umma_idesc_bf16_f32(128, N);
        const uint32_t desc_hi = static_cast<uint32_t>(umma_desc_sw128_kmajor(0) >> 32);
        const uint32_t a0 = (smem_u32(sA) & 0x3FFFF) >> 4, b0 = (smem_u32(sB) & 0x3FFFF) >> 4;
        const uint32_t b_step = (96 * 128) >> 4, tap_step = dil * 8u;
        const long long t0 = clock64();
        if (mode == 0) {
#pragma unroll 8
            for (int i = 0; i < n_mma; ++i) mma_lo(tmem_base, a0, b0, desc_hi, idesc, i > 0);
        } else {
            int i = 0;
            while (i < n_mma) {
                uint32_t a_d = a0, b_d = b0;
                for (int j = 0; j < 11 && i < n_mma; ++j) {
                    const uint32_t d = tmem_base + ((mode == 2 && (j & 1)) ? 256u : 0u);
                    mma_lo(d, a_d, b_d, desc_hi, idesc, i > 1);
                    mma_lo(d, a_d + 2, b_d + 2, desc_hi, idesc, 1);
                    mma_lo(d, a_d + 4, b_d + 4, desc_hi, idesc, 1);
                    mma_lo(d, a_d + 6, b_d + 6, desc_hi, idesc, 1);
                    i += 4;
                    a_d += tap_step;
                    b_d += b_step;
                }
            }
        }
        const long long t1 = clock64();
        umma_commit(bar);
        mbar_wait(bar, 0);
        const long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x / 32 == 0) tmem_dealloc(tmem_base, 512);
}

int main() {
    long long* dout;
    CK(cudaMalloc(&dout, 16));
    const int smem_bytes = 256 * 128 + 11 * 96 * 128 + 64 + 1024;
    CK(cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    const int n_mma = 1760;
    for (int mode = 0; mode < 4; ++mode)
        for (int N : {8, 24, 32, 48, 64, 96, 128, 256}) {
            if (mode > 0 && N > 96) continue;
            long long best[2] = {1LL << 60, 1LL << 60}, h[2];
            for (int rep = 0; rep < 5; ++rep) {
                rate_kernel<<<1, 128, smem_bytes>>>(dout, N, n_mma, mode, 3);
                CK(cudaGetLastError());
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(h, dout, 16, cudaMemcpyDeviceToHost));
                if (h[1] < best[1]) { best[0] = h[0]; best[1] = h[1]; }
            }
            printf("mode %d  N=%3d : issue %.1f cyc/MMA, completion %.1f cyc/MMA  (floor 128*N/256 = %d)\n", mode, N,
                   (double)best[0] / n_mma, (double)best[1] / n_mma, N / 2);
        }
    return 0;
}
