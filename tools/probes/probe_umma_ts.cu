// Hardware probe (developer tool, not part of libbvg.so) for the tensor-core Activation1d kernel (csrc/act1d_umma.cu):
// tcgen05.mma with the A operand in TENSOR MEMORY ("TS" form) and fp16 inputs.
//
//   1. Layout of a 16-bit A operand in TMEM: lane = M row, one 32-bit column = two consecutive K elements (low half =
//      even k), written by tcgen05.st.32x32b from packed half2 registers; a K = 16 step consumes 8 columns.
//   2. Which column offsets the A operand may start at -- the activation kernel slides its FIR window over a ring of
//      columns.  Measured (profiles/r02_probe_umma_ts.log): any multiple of 4 columns (16 bytes per lane) works, an
//      offset of 2 raises "misaligned address".
//   3. N = 32 / 64 with M = 128, fp16 x fp16 -> fp32 (idesc formats 0 / 0 / 1), B K-major SW128 written by threads.
//   4. Issue cost: cycles per tcgen05.mma for back-to-back TS MMAs (N = 32, 64, 128) from one thread.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I index-tts_b200/csrc -o build/probe_umma_ts tools/probes/probe_umma_ts.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
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

constexpr int KDIM = 64;
constexpr int MAXN = 128;
constexpr int A_BASE_COL = 256;  // A region starts here (D occupies columns [0, 128))

__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__host__ __device__ constexpr uint32_t idesc_f16_f32(uint32_t M, uint32_t N) {
    return (1u << 4) | (0u << 7) | (0u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);  // D f32, A f16, B f16, K-major both
}

// a: [128][KDIM] fp16 row-major; b: [MAXN][KDIM] fp16 row-major (row n = output column n).
// mode 0: correctness (out[128][MAXN] = A * B^T, A read from TMEM columns A_BASE_COL + acol ...)
// mode 1: timing of `reps` back-to-back TS MMAs, cycles written to cyc[0]
// mode 2: fp16 ACCUMULATOR (idesc c_format = 0): raw TMEM columns [0, 64) of every lane are written to out (as bit patterns)
__global__ void __launch_bounds__(128, 1)
probe_kernel(const __half* a, const __half* b, float* out, long long* cyc, int N, int acol, int mode, int reps) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sB = smem;  // MAXN rows x 128 B, K-major SW128
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + MAXN * 128);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2);
    const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;

    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_ptr, 512);
    // B tile: row n at n*128, 16-byte chunk (k/8) XOR (n & 7)
    for (int i = threadIdx.x; i < MAXN * KDIM / 8; i += 128) {
        const int n = i / (KDIM / 8), ch = i % (KDIM / 8);
        const uint4 v = *reinterpret_cast<const uint4*>(b + n * KDIM + ch * 8);
        *reinterpret_cast<uint4*>(sB + n * 128 + ((ch ^ (n & 7)) << 4)) = v;
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    // A: thread (warp, lane) owns TMEM lane 32*warp + lane = row m; 32 columns = 64 fp16
    {
        const int m = threadIdx.x;
        const uint32_t* arow = reinterpret_cast<const uint32_t*>(a + m * KDIM);
        uint32_t v[16];
        for (int h = 0; h < 2; ++h) {
            for (int j = 0; j < 16; ++j) v[j] = arow[16 * h + j];
            tmem_st_32x32b_x16(tmem_base + ((warp * 32u) << 16) + A_BASE_COL + acol + 16 * h, v);
        }
        tmem_st_wait();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();

    if (threadIdx.x == 0) {
        const uint32_t idesc = mode == 2 ? (idesc_f16_f32(128, N) & ~(3u << 4)) : idesc_f16_f32(128, N);
        if (mode == 0 || mode == 2) {
            for (int ks = 0; ks < KDIM / 16; ++ks) {
                const uint64_t db = umma_desc_sw128_kmajor(smem_u32(sB) + ks * 32, 0);
                umma_f16_ts(tmem_base, tmem_base + A_BASE_COL + acol + 8 * ks, db, idesc, ks > 0);
            }
            umma_commit(&bars[0]);
        } else {
            const long long t0 = clock64();
            for (int r = 0; r < reps; ++r) {
                const int ks = r & 3;
                const uint64_t db = umma_desc_sw128_kmajor(smem_u32(sB) + ks * 32, 0);
                umma_f16_ts(tmem_base, tmem_base + A_BASE_COL + 8 * ks, db, idesc, r > 0);
            }
            umma_commit(&bars[0]);
            mbar_wait(&bars[0], 0);
            cyc[0] = clock64() - t0;
        }
    }
    __syncwarp();
    mbar_wait(&bars[0], 0);
    tc_fence_after_sync();
    if (mode == 2) {
        uint32_t v[32];
        for (int c = 0; c < 64; c += 32) {
            tmem_ld_32x32b_x32(tmem_base + ((warp * 32u) << 16) + c, v);
            tmem_ld_wait();
            for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * MAXN + c + j] = __uint_as_float(v[j]);
        }
    }
    if (mode == 0) {
        uint32_t v[32];
        for (int c = 0; c < N; c += 32) {
            tmem_ld_32x32b_x32(tmem_base + ((warp * 32u) << 16) + c, v);
            tmem_ld_wait();
            for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * MAXN + c + j] = __uint_as_float(v[j]);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 512);
}

int main() {
    std::vector<__half> ha(128 * KDIM), hb(MAXN * KDIM);
    std::vector<float> fa(128 * KDIM), fb(MAXN * KDIM);
    srand(5);
    for (int i = 0; i < 128 * KDIM; ++i) { fa[i] = float(rand() % 9 - 4) * 0.5f; ha[i] = __float2half(fa[i]); }
    for (int i = 0; i < MAXN * KDIM; ++i) { fb[i] = float(rand() % 7 - 3) * 0.25f; hb[i] = __float2half(fb[i]); }
    __half *da, *db;
    float* dout;
    long long* dcyc;
    CK(cudaMalloc(&da, ha.size() * 2));
    CK(cudaMalloc(&db, hb.size() * 2));
    CK(cudaMalloc(&dout, 128 * MAXN * 4));
    CK(cudaMalloc(&dcyc, 8));
    CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
    const int smem_bytes = MAXN * 128 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    std::vector<float> hout(128 * MAXN);

    int ok = 0, total = 0;
    for (int N : {32, 64, 128})
        for (int acol : {0, 8, 16, 40, 4, 12, 20}) {  // offsets that are not multiples of 4 columns fault (misaligned address)
            CK(cudaMemset(dout, 0xff, hout.size() * 4));
            probe_kernel<<<1, 128, smem_bytes>>>(da, db, dout, dcyc, N, acol, 0, 0);
            CK(cudaGetLastError());
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("TS A-in-TMEM: N=%3d acol=%2d : LAUNCH ERROR %s\n", N, acol, cudaGetErrorString(e)); return 3; }
            CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
            int bad = 0;
            for (int m = 0; m < 128; ++m)
                for (int n = 0; n < N; ++n) {
                    float ex = 0.f;
                    for (int k = 0; k < KDIM; ++k) ex += fa[m * KDIM + k] * fb[n * KDIM + k];
                    if (hout[m * MAXN + n] != ex) ++bad;
                }
            ++total;
            if (!bad) ++ok;
            printf("TS A-in-TMEM fp16: N=%3d acol=%2d : %s (%d mismatches)  d[0][0..3] = %g %g %g %g\n", N, acol,
                   bad ? "MISMATCH" : "ok", bad, hout[0], hout[1], hout[2], hout[3]);
        }
    printf("SUMMARY TS-mode A operand: %d/%d ok\n", ok, total);

    // ---- fp16 accumulator: where do the N results of a row live?
    for (int N : {32, 64}) {
        CK(cudaMemset(dout, 0, hout.size() * 4));
        probe_kernel<<<1, 128, smem_bytes>>>(da, db, dout, dcyc, N, 0, 2, 0);
        CK(cudaGetLastError());
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("fp16 accumulator: N=%d LAUNCH ERROR %s\n", N, cudaGetErrorString(e)); return 3; }
        CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
        int bad_packed = 0, bad_low = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < N; ++n) {
                float ex = 0.f;
                for (int k = 0; k < KDIM; ++k) ex += fa[m * KDIM + k] * fb[n * KDIM + k];
                const uint32_t* raw = reinterpret_cast<const uint32_t*>(&hout[m * MAXN]);
                const uint32_t wp = raw[n >> 1];
                const __half hp = __ushort_as_half((unsigned short)((n & 1) ? (wp >> 16) : (wp & 0xffff)));
                const __half hl = __ushort_as_half((unsigned short)(raw[n] & 0xffff));
                if (__half2float(hp) != ex) ++bad_packed;
                if (__half2float(hl) != ex) ++bad_low;
            }
        const uint32_t* r0 = reinterpret_cast<const uint32_t*>(&hout[0]);
        printf("fp16 accumulator N=%d: packed-pairs hypothesis %d mismatches, low-half hypothesis %d mismatches; lane 0 raw: %08x %08x %08x %08x\n",
               N, bad_packed, bad_low, r0[0], r0[1], r0[2], r0[3]);
    }

    for (int N : {16, 32, 64, 128}) {
        const int reps = 512;
        probe_kernel<<<1, 128, smem_bytes>>>(da, db, dout, dcyc, N, 0, 1, reps);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        long long c = 0;
        CK(cudaMemcpy(&c, dcyc, 8, cudaMemcpyDeviceToHost));
        printf("TS issue rate: N=%3d : %.1f cycles per tcgen05.mma (%d back to back, one issuer)\n", N, double(c) / reps, reps);
    }
    return 0;
}
