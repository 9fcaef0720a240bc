// Hardware probe (developer tool): throughput of the legacy warp-level mma.sync.m16n8k16 (fp16 in, fp32 accumulate) on
// sm_100a -- the register-level tensor-core path considered for Activation1d's FIR filters (no TMEM round trip).
// Prints cycles per MMA per scheduler for 1..8 warps per scheduler, with 1, 2 and 4 independent accumulator chains.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/probe_mma_sync_rate tools/probes/probe_mma_sync_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma16816(float (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int CHAINS>
__global__ void rate_kernel(long long* cyc, float* sink, int reps) {
    unsigned a[4] = {0x3c003c00u + threadIdx.x, 0x3c003c00u, 0x38003800u, 0x34003400u};
    unsigned b[2] = {0x3c003c00u, 0x38003800u + threadIdx.x};
    float d[CHAINS][4];
    for (int c = 0; c < CHAINS; ++c) for (int i = 0; i < 4; ++i) d[c][i] = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) mma16816(d[c], a, b);
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int c = 0; c < CHAINS; ++c) for (int i = 0; i < 4; ++i) s += d[c][i];
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    long long* dc; float* ds;
    cudaMalloc(&dc, 8 * 1024); cudaMalloc(&ds, 4 * 1024 * 1024);
    const int reps = 2048;
    for (int warps : {4, 8, 16, 32}) {
        for (int chains : {1, 2, 4}) {
            long long h = 0;
            if (chains == 1) rate_kernel<1><<<148, warps * 32>>>(dc, ds, reps);
            if (chains == 2) rate_kernel<2><<<148, warps * 32>>>(dc, ds, reps);
            if (chains == 4) rate_kernel<4><<<148, warps * 32>>>(dc, ds, reps);
            if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 2; }
            cudaMemcpy(&h, dc, 8, cudaMemcpyDeviceToHost);
            const double per_sched = double(h) / (double(reps) * chains * (warps / 4.0));
            printf("mma.sync m16n8k16 f16->f32: %2d warps/SM (%d per scheduler), %d chains: %.2f cycles per MMA per scheduler "
                   "(%.0f dense FMA/clk/SM)\n", warps, warps / 4, chains, per_sched, 4.0 * 2048.0 / per_sched);
        }
    }
    return 0;
}
