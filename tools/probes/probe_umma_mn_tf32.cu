// TF32 variant of probe_umma_mn.cu (kind::tf32, fp32 operands in shared memory, K = 8 per MMA, 32-element atoms).
// RESULT on B200 (profiles/r01_probe_umma_mn_tf32.log): kind::tf32 works with both operands K-major, but with the B
// operand flagged Major-MN the MMA returns all zeros -- like wgmma, tcgen05 transposes 16-bit operands only.  A TF32
// FIR-as-GEMM kernel therefore needs its activations time-contiguous ([B, C, T]), not channels-last.
// Hardware probe (developer tool, not part of libbvg.so) for the round-2 plan of running Activation1d's FIR filters on
// the tensor cores as banded Toeplitz GEMMs:
//
//   1. MN-major B operand.  A channels-last activation tile [time rows][64 channels] loaded by TMA with SWIZZLE_128B is
//      bit for bit the canonical "Major-MN, SW128" UMMA layout with MN = channels and K = time (atom: 64 MN elements
//      = 128 B contiguous, 8 K rows 128 B apart; LBO = distance between 64-channel atoms, SBO = distance between 8-row
//      K groups).  D[m, n] = sum_k F[m, k] * X[row0 + shift + k, n] is computed with F K-major and X MN-major, the K
//      step advancing the X start address by 16 rows, and `shift` an arbitrary (not 8-aligned) row offset.
//   2. M = 64 MMAs (to skip the empty half of a band block): which TMEM lanes receive the 64 result rows.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I. -o probe_umma_mn probe_umma_mn_tf32.cu && ./probe_umma_mn_tf32
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include <cmath>
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

constexpr int X_ROWS = 256;   // time rows of the X tile in smem (per 64-channel atom)
constexpr int KDIM = 32;      // K = time taps covered by one F tile (32 fp32 = one 128-byte row)
constexpr int MAXN = 128;

// Major-MN SWIZZLE_128B descriptor: start address, LBO (between 64-element MN atoms), SBO (between 8-row K groups).
__device__ __forceinline__ uint64_t umma_desc_sw128_mnmajor(uint32_t byte_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((byte_addr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= static_cast<uint64_t>(1) << 46;  // descriptor version (sm_100)
    d |= static_cast<uint64_t>(2) << 61;  // SWIZZLE_128B
    return d;
}

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap map_f, const __grid_constant__ CUtensorMap map_x, float* out, int M, int N,
             int row0, int shift) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sF = smem;                          // 128 x 128 B   (F: [M rows][K = 64], K-major)
    uint8_t* sX = smem + 128 * 128;              // 2 atoms x 256 rows x 128 B (X: [time][64 ch] per atom)
    uint64_t* bars = reinterpret_cast<uint64_t*>(sX + 2 * X_ROWS * 128);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2);
    const int warp = threadIdx.x / 32;

    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_ptr, 128);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bars[0], 128 * 128 + 2 * X_ROWS * 128);
        tma_load_2d(sF, &map_f, &bars[0], 0, 0);
        tma_load_2d(sX, &map_x, &bars[0], 0, row0);                      // channels 0..63
        tma_load_2d(sX + X_ROWS * 128, &map_x, &bars[0], 32, row0);      // channels 32..63
        mbar_wait(&bars[0], 0);
        tc_fence_after_sync();
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 16) | ((uint32_t)(N >> 3) << 17) |
                               ((uint32_t)(M >> 4) << 24);  // D f32, A/B tf32, b_major = MN
        const bool kmajor_sanity = (shift == -1000);
        const uint32_t idesc_k = idesc & ~(1u << 16);
        for (int ks = 0; ks < KDIM / 8; ++ks) {
            const uint64_t da = umma_desc_sw128_kmajor(smem_u32(sF) + ks * 32, 0);
            const uint64_t db = kmajor_sanity ? umma_desc_sw128_kmajor(smem_u32(sF) + ks * 32, 0)
                                              : umma_desc_sw128_mnmajor(smem_u32(sX) + (shift + 8 * ks) * 128, X_ROWS * 128, 1024);
            {
                const uint32_t accumulate = ks > 0;
                asm volatile(
                    "{\n"
                    ".reg .pred p;\n"
                    "setp.ne.b32 p, %4, 0;\n"
                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
                    "}\n" ::"r"(tmem_base), "l"(da), "l"(db), "r"(kmajor_sanity ? idesc_k : idesc), "r"(accumulate)
                    : "memory");
            }
        }
        umma_commit(&bars[1]);
    }
    __syncwarp();
    mbar_wait(&bars[1], 0);
    tc_fence_after_sync();
    uint32_t v[32];
    for (int c = 0; c < N; c += 32) {
        tmem_ld_32x32b_x32(tmem_base + ((warp * 32u) << 16) + c, v);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[(warp * 32 + threadIdx.x % 32) * MAXN + c + j] = __uint_as_float(v[j]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 128);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int R = 1024, C = 64;
    std::vector<float> hx(R * C), hf(128 * KDIM);
    std::vector<float> fx(R * C), ff(128 * KDIM);
    srand(3);
    for (int i = 0; i < R * C; ++i) { fx[i] = float(rand() % 9 - 4); hx[i] = fx[i]; }
    for (int i = 0; i < 128 * KDIM; ++i) { ff[i] = float(rand() % 7 - 3); hf[i] = ff[i]; }
    float *dx, *df;
    float* dout;
    CK(cudaMalloc(&dx, hx.size() * 4));
    CK(cudaMalloc(&df, hf.size() * 4));
    CK(cudaMalloc(&dout, 128 * MAXN * 4));
    CK(cudaMemcpy(dx, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(df, hf.data(), hf.size() * 4, cudaMemcpyHostToDevice));

    EncodeTiledFn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres));
    if (!encode) { printf("no cuTensorMapEncodeTiled\n"); return 2; }
    CUtensorMap map_f, map_x;
    {
        cuuint32_t es[2] = {1, 1};
        cuuint64_t dimsf[2] = {KDIM, 128};
        cuuint64_t stridesf[1] = {KDIM * 4};
        cuuint32_t boxf[2] = {KDIM, 128};
        CUresult r = encode(&map_f, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, df, dimsf, stridesf, boxf, es,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode F failed %d\n", (int)r); return 2; }
        cuuint64_t dimsx[2] = {(cuuint64_t)C, (cuuint64_t)R};
        cuuint64_t stridesx[1] = {(cuuint64_t)C * 4};
        cuuint32_t boxx[2] = {32, X_ROWS};
        r = encode(&map_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dx, dimsx, stridesx, boxx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode X failed %d\n", (int)r); return 2; }
    }
    const int smem_bytes = 128 * 128 + 2 * X_ROWS * 128 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    std::vector<float> hout(128 * MAXN);

    {   // ---- 0. sanity: kind::tf32 with both operands K-major: D = F F^T (N = 64 rows of F)
        CK(cudaMemset(dout, 0xff, hout.size() * 4));
        probe_kernel<<<1, 128, smem_bytes>>>(map_f, map_x, dout, 128, 64, 0, -1000);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 64; ++n) {
                float e = 0.f;
                for (int k = 0; k < KDIM; ++k) e += ff[m * KDIM + k] * ff[n * KDIM + k];
                if (hout[m * MAXN + n] != e) ++bad;
            }
        printf("TF32 K-major sanity (D = F F^T): %s (%d mismatches)\n", bad ? "MISMATCH" : "ok", bad);
    }
    // ---- 1. MN-major B, N = 64 (one atom) and N = 128 (two atoms, LBO), row shifts
    int ok = 0, total = 0;
    const int shifts[] = {0, 1, 3, 8, 13, 40, 101};
    const int row0s[] = {0, 300, -5};
    for (int N : {32, 64})
        for (int row0 : row0s)
            for (int s : shifts) {
                CK(cudaMemset(dout, 0xff, hout.size() * 4));
                probe_kernel<<<1, 128, smem_bytes>>>(map_f, map_x, dout, 128, N, row0, s);
                CK(cudaGetLastError());
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
                int bad = 0;
                for (int m = 0; m < 128; ++m)
                    for (int n = 0; n < N; ++n) {
                        float e = 0.f;
                        for (int k = 0; k < KDIM; ++k) {
                            const int r = row0 + s + k;
                            if (r >= 0 && r < R) e += ff[m * KDIM + k] * fx[r * C + n];
                        }
                        if (hout[m * MAXN + n] != e) ++bad;
                    }
                ++total;
                if (!bad) ++ok;
                printf("TF32 MN-major B: N=%3d row0=%4d shift=%3d : %s (%d mismatches)\n", N, row0, s, bad ? "MISMATCH" : "ok", bad);
            }
    printf("SUMMARY TF32 MN-major B operand: %d/%d ok\n", ok, total);

    {   // ---- 1b. decode: F = one-hot (F[m][k] = (m == k)), X[r][c] = r * 64 + c  ->  D[m][n] shows which X element the
        //           hardware reads as B[k = m][n]
        std::vector<float> f1(128 * KDIM, 0.f), x1(R * C);
        for (int m = 0; m < KDIM; ++m) f1[m * KDIM + m] = 1.f;
        for (int r = 0; r < R; ++r) for (int c = 0; c < C; ++c) x1[r * C + c] = float(r * 64 + c);
        CK(cudaMemcpy(df, f1.data(), f1.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dx, x1.data(), x1.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemset(dout, 0xff, hout.size() * 4));
        probe_kernel<<<1, 128, smem_bytes>>>(map_f, map_x, dout, 128, 64, 0, 0);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
        for (int m = 0; m < 4; ++m) {
            printf("decode k=%2d:", m);
            for (int n = 0; n < 64; n += 1) { float v = hout[m * MAXN + n]; printf(" %d.%d", (int)v / 64, (int)v % 64); if (n == 11) n = 27; if (n == 35) break; }
            printf("\n");
        }
        CK(cudaMemcpy(df, hf.data(), hf.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dx, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice));
    }
    // ---- 2. M = 64: where do the 64 result rows land in TMEM?
    {
        const int N = 32, row0 = 0, s = 0;
        CK(cudaMemset(dout, 0xff, hout.size() * 4));
        probe_kernel<<<1, 128, smem_bytes>>>(map_f, map_x, dout, 64, N, row0, s);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(hout.data(), dout, hout.size() * 4, cudaMemcpyDeviceToHost));
        std::vector<int> lane_of(64, -1);
        int found = 0;
        for (int m = 0; m < 64; ++m) {
            for (int l = 0; l < 128 && lane_of[m] < 0; ++l) {
                bool match = true;
                for (int n = 0; n < N && match; ++n) {
                    float e = 0.f;
                    for (int k = 0; k < KDIM; ++k) e += ff[m * KDIM + k] * fx[(row0 + s + k) * C + n];
                    if (hout[l * MAXN + n] != e) match = false;
                }
                if (match) lane_of[m] = l;
            }
            if (lane_of[m] >= 0) ++found;
        }
        printf("M=64: %d/64 result rows found; row -> TMEM lane:", found);
        for (int m = 0; m < 64; ++m) printf(" %d", lane_of[m]);
        printf("\n");
    }
    return 0;
}
