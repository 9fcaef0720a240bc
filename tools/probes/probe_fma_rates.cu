// Hardware probe (developer tool, not part of libbvg.so): issue rates of the FP32 / FP16 FMA forms the SIMT Activation1d
// kernel (csrc/act1d.cu) can be built from, per SM sub-partition, and whether they overlap with each other and with MUFU.
//
// ncu of act1d_nlc_tma_kernel (profiles/r02_p_act_c384.ncu.txt): fma-pipe INSTRUCTIONS 36 % of peak but fma-pipe CYCLES
// 67 % (fmaheavy 62 %): a packed fma.rn.f32x2 holds the pipe for two cycles.  Question: is there a form (scalar FFMA with
// an immediate tap, HFMA2, a mix of packed and scalar) that retires more FIR taps per clock?
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o build/probe_fma_rates tools/probes/probe_fma_rates.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_fp16.h>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);     \
            exit(2);                                                                            \
        }                                                                                       \
    } while (0)

constexpr int CH = 8;       // independent dependency chains per thread
constexpr int ITERS = 2048;

#define FFMA_RR(a, b, c) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a) : "f"(b), "f"(c))
#define FFMA_IMM(a, c) asm volatile("fma.rn.f32 %0, %0, 0f3F7FF000, %1;" : "+f"(a) : "f"(c))
#define FFMA2_RR(a, b, c) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a) : "l"(b), "l"(c))
#define HFMA2_RR(a, b, c) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define HFMA2_IMM(a, c) asm volatile("{.reg .b32 t; mov.b32 t, 0x3BFF3BFF; fma.rn.f16x2 %0, %0, t, %1;}" : "+r"(a) : "r"(c))
#define BFMA2_RR(a, b, c) asm volatile("fma.rn.bf16x2 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c))
#define FFMA2_D3(d, a, b) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b))
#define FFMA2_DIMM(d, a) asm volatile("{.reg .b64 t; mov.b64 t, 0x3F7FF0003F7FF000; fma.rn.f32x2 %0, %1, t, %0;}" : "+l"(d) : "l"(a))
#define FFMA2_D4(d, a, b, c) asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c))
#define MUFU_COS(a) asm volatile("cos.approx.ftz.f32 %0, %0;" : "+f"(a))
#define FADD2_RR(a, b) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(b))
#define FMUL2_RR(a, b) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(b))

// mode: which instruction mix one loop iteration issues, CH times each
template <int mode>
__global__ void __launch_bounds__(1024, 1) rate_kernel(float seed, long long* cyc, float* sink) {
    float f[CH], g[CH];
    unsigned long long p[CH];
    unsigned h[CH];
    unsigned long long q[CH], r[CH], w[CH];
    const float b = seed * 0.999f, c = seed * 1e-3f;
    unsigned long long pb, pc;
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(pb) : "f"(b));
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(pc) : "f"(c));
    const unsigned hb = 0x3BFF3BFFu ^ (seed == 2.f ? 1u : 0u), hc = 0x14001400u ^ (seed == 3.f ? 1u : 0u);  // runtime values
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        f[i] = seed + i + threadIdx.x;
        g[i] = seed * 0.5f + i;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(f[i]), "f"(g[i]));
        h[i] = 0x3C003C00u + i;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(q[i]) : "f"(f[i] * 1e-3f), "f"(g[i] * 1e-3f));
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(r[i]) : "f"(g[i] * 1e-3f), "f"(f[i] * 1e-3f));
        w[i] = 0;
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            switch (mode) {
                case 0: FFMA_RR(f[i], b, c); break;
                case 1: FFMA_IMM(f[i], c); break;
                case 2: FFMA2_RR(p[i], pb, pc); break;
                case 3: HFMA2_RR(h[i], hb, hc); break;
                case 4: HFMA2_IMM(h[i], hc); break;
                case 5: FFMA2_RR(p[i], pb, pc); FFMA_RR(f[i], b, c); break;
                case 6: FFMA2_RR(p[i], pb, pc); HFMA2_RR(h[i], hb, hc); break;
                case 7: MUFU_COS(f[i]); break;
                case 8: FFMA2_RR(p[i], pb, pc); FFMA2_RR(p[i], pc, pb); FFMA2_RR(p[i], pb, pc); MUFU_COS(f[i]); break;
                case 9: FFMA_RR(f[i], b, c); HFMA2_RR(h[i], hb, hc); break;
                case 10: BFMA2_RR(h[i], hb, hc); break;
                case 11: FFMA2_RR(p[i], pb, pc); FFMA_IMM(f[i], c); break;
                case 12: FADD2_RR(p[i], pc); break;
                case 13: FMUL2_RR(p[i], pb); break;
                case 14: FFMA_RR(f[i], b, c); FFMA_RR(g[i], c, b); break;
                case 15: FFMA2_D3(p[i], q[i], r[i]); break;                    // three distinct register pairs
                case 16: FFMA2_DIMM(p[i], q[i]); break;                        // two register pairs + constant
                case 17: FFMA2_D3(p[i], q[i], pb); break;                      // multiplier shared by all chains
                case 18: FFMA2_D4(w[i], q[i], r[i], p[i]); break;              // four distinct pairs, no dependency
                default: break;
            }
        }
    }
    const long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        float lo, hi;
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));
        acc += f[i] + g[i] + lo + hi + __uint_as_float(h[i]);
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(w[i]));
        acc += lo + hi;
    }
    if (acc == 123.456f) sink[0] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void launch(int sms, int threads, long long* dcyc, float* dsink) {
    rate_kernel<MODE><<<sms, threads>>>(1.0f, dcyc, dsink);
}
static void launch_mode(int mode, int sms, int threads, long long* dcyc, float* dsink) {
    switch (mode) {
        case 0: launch<0>(sms, threads, dcyc, dsink); break;
        case 1: launch<1>(sms, threads, dcyc, dsink); break;
        case 2: launch<2>(sms, threads, dcyc, dsink); break;
        case 3: launch<3>(sms, threads, dcyc, dsink); break;
        case 4: launch<4>(sms, threads, dcyc, dsink); break;
        case 5: launch<5>(sms, threads, dcyc, dsink); break;
        case 6: launch<6>(sms, threads, dcyc, dsink); break;
        case 7: launch<7>(sms, threads, dcyc, dsink); break;
        case 8: launch<8>(sms, threads, dcyc, dsink); break;
        case 9: launch<9>(sms, threads, dcyc, dsink); break;
        case 10: launch<10>(sms, threads, dcyc, dsink); break;
        case 11: launch<11>(sms, threads, dcyc, dsink); break;
        case 12: launch<12>(sms, threads, dcyc, dsink); break;
        case 13: launch<13>(sms, threads, dcyc, dsink); break;
        case 14: launch<14>(sms, threads, dcyc, dsink); break;
        case 15: launch<15>(sms, threads, dcyc, dsink); break;
        case 16: launch<16>(sms, threads, dcyc, dsink); break;
        case 17: launch<17>(sms, threads, dcyc, dsink); break;
        case 18: launch<18>(sms, threads, dcyc, dsink); break;
        default: break;
    }
}

int main() {
    long long* dcyc;
    float* dsink;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaMalloc(&dcyc, sizeof(long long) * sms));
    CK(cudaMalloc(&dsink, 4));
    struct M { int mode; const char* name; double fma_per_group; int instr_per_group; };
    const M modes[] = {
        {0, "FFMA reg,reg (scalar fp32)", 1, 1},
        {1, "FFMA with an immediate multiplier", 1, 1},
        {2, "FFMA2 (fma.rn.f32x2)", 2, 1},
        {15, "FFMA2 d = a*b + d, three distinct pairs", 2, 1},
        {16, "FFMA2 d = a*const + d", 2, 1},
        {17, "FFMA2 d = a*B + d, B shared", 2, 1},
        {3, "HFMA2 reg,reg (fma.rn.f16x2)", 2, 1},
        {4, "HFMA2 with an immediate multiplier", 2, 1},
        {10, "HFMA2.BF16 (fma.rn.bf16x2)", 2, 1},
        {12, "FADD2 (add.rn.f32x2)", 2, 1},
        {13, "FMUL2 (mul.rn.f32x2)", 2, 1},
        {14, "FFMA + FFMA (two scalar chains)", 2, 2},
        {5, "FFMA2 + FFMA interleaved", 3, 2},
        {11, "FFMA2 + FFMA-imm interleaved", 3, 2},
        {6, "FFMA2 + HFMA2 interleaved", 4, 2},
        {9, "FFMA + HFMA2 interleaved", 3, 2},
        {7, "MUFU.COS", 1, 1},
        {8, "3 x FFMA2 + MUFU.COS", 6, 4},
    };
    for (int warps : {16, 32}) {
        for (const M& m : modes) {
            launch_mode(m.mode, sms, warps * 32, dcyc, dsink);
            CK(cudaDeviceSynchronize());
            launch_mode(m.mode, sms, warps * 32, dcyc, dsink);
            CK(cudaDeviceSynchronize());
            long long c = 0;
            CK(cudaMemcpy(&c, dcyc, sizeof(c), cudaMemcpyDeviceToHost));
            const double groups = double(ITERS) * CH * warps;  // warp-level instruction groups per SM
            const double cyc_per_group_smsp = double(c) / (groups / 4.0);
            printf("%2d warps/SM  %-38s %6.2f cycles per warp-group per SMSP  (%4.2f instr/clk/SMSP, %6.1f FMA-or-op lanes/clk/SM)\n",
                   warps, m.name, cyc_per_group_smsp, m.instr_per_group / cyc_per_group_smsp,
                   m.fma_per_group * 32.0 * 4.0 / cyc_per_group_smsp);
        }
    }
    return 0;
}
