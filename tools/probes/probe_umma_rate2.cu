// Hardware probe (developer tool, not part of libbvg.so), companion of probe_umma_rate.cu: the same back-to-back issue
// test for a CTA PAIR -- tcgen05.mma.cta_group::2, M = 256 (128 rows in each CTA), K = 16, bf16, operands in shared
// memory.  Question for the round-2 plan: does one cta_group::2 instruction (two SMs, 256 rows) cost the same ~56
// dispatch cycles as one cta_group::1 instruction (one SM, 128 rows) at small N?
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I. -o probe_umma_rate2 probe_umma_rate2.cu && ./probe_umma_rate2
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

__device__ __forceinline__ void mma2_lo(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %5, 0;\n"
        "mov.b64 da, {%1, %3};\n"
        "mov.b64 db, {%2, %3};\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n"
        "}\n" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
rate2_kernel(long long* out, int N, int n_mma, int mode, int dil) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;               // 256 rows x 128 B (this CTA's 128 rows of A + tap shifts)
    uint8_t* sB = smem + 256 * 128;   // 11 taps x 64 rows x 128 B (this CTA's half of B)
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB + 11 * 64 * 128);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar + 1);
    uint32_t rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    for (int i = threadIdx.x; i < (256 * 128 + 11 * 64 * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    if (threadIdx.x / 32 == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before_sync();
    cluster_sync_all();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;
    if (threadIdx.x == 0 && rank == 0) {
        const uint32_t idesc = umma_idesc_bf16_f32(256, N);
        const uint32_t desc_hi = static_cast<uint32_t>(umma_desc_sw128_kmajor(0) >> 32);
        const uint32_t a0 = (smem_u32(sA) & 0x3FFFF) >> 4, b0 = (smem_u32(sB) & 0x3FFFF) >> 4;
        const uint32_t b_step = (64 * 128) >> 4, tap_step = dil * 8u;
        const long long t0 = clock64();
        if (mode == 0) {
#pragma unroll 8
            for (int i = 0; i < n_mma; ++i) mma2_lo(tmem_base, a0, b0, desc_hi, idesc, i > 0);
        } else {
            int i = 0;
            while (i < n_mma) {
                uint32_t a_d = a0, b_d = b0;
                for (int j = 0; j < 11 && i < n_mma; ++j) {
                    mma2_lo(tmem_base, a_d, b_d, desc_hi, idesc, i > 1);
                    mma2_lo(tmem_base, a_d + 2, b_d + 2, desc_hi, idesc, 1);
                    mma2_lo(tmem_base, a_d + 4, b_d + 4, desc_hi, idesc, 1);
                    mma2_lo(tmem_base, a_d + 6, b_d + 6, desc_hi, idesc, 1);
                    i += 4;
                    a_d += tap_step;
                    b_d += b_step;
                }
            }
        }
        const long long t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                         smem_u32(bar)),
                     "h"((uint16_t)1)
                     : "memory");
        mbar_wait(bar, 0);
        const long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
    }
    tc_fence_before_sync();
    cluster_sync_all();
    if (threadIdx.x / 32 == 0)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

int main() {
    long long* dout;
    CK(cudaMalloc(&dout, 16));
    const int smem_bytes = 256 * 128 + 11 * 64 * 128 + 64 + 1024;
    CK(cudaFuncSetAttribute(rate2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    const int n_mma = 1760;
    for (int mode = 0; mode < 2; ++mode)
        for (int N : {32, 48, 64, 96, 128, 256}) {
            if (mode > 0 && N > 128) continue;
            long long best[2] = {1LL << 60, 1LL << 60}, h[2];
            for (int rep = 0; rep < 5; ++rep) {
                rate2_kernel<<<2, 128, smem_bytes>>>(dout, N, n_mma, mode, 3);
                CK(cudaGetLastError());
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(h, dout, 16, cudaMemcpyDeviceToHost));
                if (h[1] < best[1]) { best[0] = h[0]; best[1] = h[1]; }
            }
            printf("cta_group::2 M=256 mode %d  N=%3d : issue %.1f cyc/MMA, completion %.1f cyc/MMA  (pipe floor 128*N/512 = %d)\n",
                   mode, N, (double)best[0] / n_mma, (double)best[1] / n_mma, N / 4);
        }
    return 0;
}
