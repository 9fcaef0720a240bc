// Hardware probe (developer tool, not part of libbvg.so): does a UMMA shared-memory descriptor whose start address
// is advanced by a whole number of 128-byte rows inside a SWIZZLE_128B K-major tile read the row-shifted operand?
// This is the property the implicit-GEMM conv kernel relies on to serve all k dilated taps from ONE halo'd TMA tile.
// Also checks TMA out-of-bounds zero fill with negative coordinates.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_umma_shift probe_umma_shift.cu && ./probe_umma_shift
#include <cstdio>
#include <cstdlib>
#include <vector>
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

constexpr int A_ROWS = 256;  // rows of the A tile in smem
constexpr int N = 64;
constexpr int M = 128;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, float* out,
             int row0, int shift, int base_off_mode) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                       // 256 x 128 B
    uint8_t* sB = smem + A_ROWS * 128;        // 64 x 128 B
    uint64_t* bars = reinterpret_cast<uint64_t*>(sB + N * 128);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2);
    const int warp = threadIdx.x / 32;

    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_ptr, 64);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bars[0], A_ROWS * 128 + N * 128);
        tma_load_2d(sA, &map_a, &bars[0], 0, row0);
        tma_load_2d(sB, &map_b, &bars[0], 0, 0);
        mbar_wait(&bars[0], 0);
        tc_fence_after_sync();
        const uint32_t idesc = umma_idesc_bf16_f32(M, N);
        const uint32_t a_addr = smem_u32(sA) + shift * 128;
        const uint32_t bo = base_off_mode ? (shift & 7) : 0;
        for (int ks = 0; ks < 4; ++ks) {
            uint64_t da = umma_desc_sw128_kmajor(a_addr + ks * 32, bo);
            uint64_t db = umma_desc_sw128_kmajor(smem_u32(sB) + ks * 32, 0);
            umma_f16(tmem_base, da, db, idesc, ks > 0);
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
        for (int j = 0; j < 32; ++j) out[(warp * 32 + threadIdx.x % 32) * N + c + j] = __uint_as_float(v[j]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 64);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int R = 512;
    std::vector<__nv_bfloat16> hx(R * 64), hw(N * 64);
    std::vector<float> fx(R * 64), fw(N * 64);
    srand(1);
    for (int i = 0; i < R * 64; ++i) { fx[i] = float(rand() % 9 - 4); hx[i] = __float2bfloat16(fx[i]); }
    for (int i = 0; i < N * 64; ++i) { fw[i] = float(rand() % 7 - 3); hw[i] = __float2bfloat16(fw[i]); }
    __nv_bfloat16 *dx, *dw;
    float* dout;
    CK(cudaMalloc(&dx, hx.size() * 2));
    CK(cudaMalloc(&dw, hw.size() * 2));
    CK(cudaMalloc(&dout, M * N * 4));
    CK(cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice));

    EncodeTiledFn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres));
    if (!encode) { printf("no cuTensorMapEncodeTiled\n"); return 2; }

    CUtensorMap map_a, map_b;
    {
        cuuint64_t dims[2] = {64, (cuuint64_t)R};
        cuuint64_t strides[1] = {128};
        cuuint32_t box[2] = {64, A_ROWS};
        cuuint32_t es[2] = {1, 1};
        CUresult r = encode(&map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dx, dims, strides, box, es,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode A failed %d\n", (int)r); return 2; }
        cuuint64_t dimsb[2] = {64, (cuuint64_t)N};
        cuuint32_t boxb[2] = {64, N};
        r = encode(&map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dw, dimsb, strides, boxb, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode B failed %d\n", (int)r); return 2; }
    }
    const int smem_bytes = A_ROWS * 128 + N * 128 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));

    const int shifts[] = {0, 1, 2, 3, 5, 7, 8, 9, 10, 15, 25, 50, 100};
    const int row0s[] = {0, 64, -3, 300};
    std::vector<float> hout(M * N);
    int ok_mode[2] = {0, 0}, total = 0;
    for (int mode = 0; mode < 2; ++mode) {
        for (int row0 : row0s) {
            for (int s : shifts) {
                CK(cudaMemset(dout, 0xff, M * N * 4));
                probe_kernel<<<1, 128, smem_bytes>>>(map_a, map_b, dout, row0, s, mode);
                CK(cudaGetLastError());
                CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(hout.data(), dout, M * N * 4, cudaMemcpyDeviceToHost));
                int bad = 0;
                for (int m = 0; m < M; ++m)
                    for (int n = 0; n < N; ++n) {
                        int r = row0 + m + s;
                        float e = 0.f;
                        if (r >= 0 && r < R)
                            for (int k = 0; k < 64; ++k) e += fx[r * 64 + k] * fw[n * 64 + k];
                        if (hout[m * N + n] != e) ++bad;
                    }
                if (mode == 0) ++total;
                if (!bad) ++ok_mode[mode];
                printf("base_offset_mode=%d row0=%4d shift=%3d : %s (%d mismatches)\n", mode, row0, s,
                       bad ? "MISMATCH" : "ok", bad);
            }
        }
    }
    printf("SUMMARY base_offset=0: %d/%d ok ; base_offset=(shift&7): %d/%d ok\n", ok_mode[0], total, ok_mode[1], total);
    return 0;
}
