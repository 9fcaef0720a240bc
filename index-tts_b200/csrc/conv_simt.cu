// BVG_EXACT_F32 convolution: the same implicit-GEMM formulation as conv_umma.cu (channels-last, per-phase tap
// bases, fused bias/residual/scale/accum epilogue) evaluated with fp32 operands and fp32 FMA accumulation on the
// CUDA cores.  This is the parity mode (fp32 max-abs <= 1e-4, SNR >= 80 dB against the reference's fp32 CPU path);
// the throughput mode is the tcgen05 kernel.
#include "bvg_common.cuh"

namespace bvg {

constexpr int SIMT_BM = 128;  // time rows per CTA
constexpr int SIMT_BN = 64;   // output channels per CTA
constexpr int SIMT_BK = 16;   // input channels per smem chunk
constexpr int SIMT_THREADS = 256;
constexpr int SIMT_PAD = 4;

struct ConvSimtParams {
    const float* x;  // [B,T,Cin]
    const float* w;  // [taps][N][Cin]
    const float* bias;
    const float* cond;  // per-utterance bias [B, cout_phase] or null (v1 generator conditioning, models.py:236, 244-245)
    const float* residual;
    const float* accum;
    float* y_f32;
    __nv_bfloat16* y_bf16;
    int64_t T;
    int Cin, N, cout_phase, taps, dil, num_phases, tiles_per_phase;
    int tap_base[BVG_MAX_PHASES];
    float out_scale;
};

// grid = (m_tiles, num_phases * tiles_per_phase, B)
__global__ void __launch_bounds__(SIMT_THREADS)
conv_simt_kernel(const __grid_constant__ ConvSimtParams p) {
    __shared__ float Xs[SIMT_BK][SIMT_BM + SIMT_PAD];
    __shared__ float Ws[SIMT_BK][SIMT_BN + SIMT_PAD];
    const int tid = threadIdx.x;
    const int64_t b = blockIdx.z;
    const int phase = blockIdx.y / p.tiles_per_phase;
    const int c0 = (blockIdx.y % p.tiles_per_phase) * SIMT_BN;  // channel offset inside the phase
    const int n0 = phase * p.cout_phase + c0;
    const int64_t t0 = (int64_t)blockIdx.x * SIMT_BM;
    const int ty = tid / 16, tx = tid % 16;

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const float* xb = p.x + b * p.T * p.Cin;
    for (int j = 0; j < p.taps; ++j) {
        const int64_t row_base = t0 + p.tap_base[phase] + (int64_t)j * p.dil;
        const float* wj = p.w + ((int64_t)j * p.N) * p.Cin;
        for (int ci0 = 0; ci0 < p.Cin; ci0 += SIMT_BK) {
            // X chunk: 128 rows x 16 channels = 512 float4, 2 per thread
#pragma unroll
            for (int l = 0; l < 2; ++l) {
                const int idx = tid + l * SIMT_THREADS;
                const int r = idx / 4, g = idx % 4;
                const int64_t row = row_base + r;
                const int ci = ci0 + g * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (row >= 0 && row < p.T && ci < p.Cin) v = *reinterpret_cast<const float4*>(xb + row * p.Cin + ci);
                Xs[g * 4 + 0][r] = v.x; Xs[g * 4 + 1][r] = v.y; Xs[g * 4 + 2][r] = v.z; Xs[g * 4 + 3][r] = v.w;
            }
            // W chunk: 64 rows x 16 channels = 256 float4, 1 per thread
            {
                const int r = tid / 4, g = tid % 4;
                const int n = n0 + r;
                const int ci = ci0 + g * 4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (n < p.N && ci < p.Cin) v = __ldg(reinterpret_cast<const float4*>(wj + (int64_t)n * p.Cin + ci));
                Ws[g * 4 + 0][r] = v.x; Ws[g * 4 + 1][r] = v.y; Ws[g * 4 + 2][r] = v.z; Ws[g * 4 + 3][r] = v.w;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < SIMT_BK; ++kk) {
                const float4 a0 = *reinterpret_cast<const float4*>(&Xs[kk][ty * 8]);
                const float4 a1 = *reinterpret_cast<const float4*>(&Xs[kk][ty * 8 + 4]);
                const float4 bb = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
                const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                const float w4[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[i][q] = fmaf(a[i], w4[q], acc[i][q]);
            }
            __syncthreads();
        }
    }

    const int ch = c0 + tx * 4;  // channel inside the phase
    if (ch >= p.cout_phase) return;
    float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.bias) bias4 = __ldg(reinterpret_cast<const float4*>(p.bias + ch));
    float4 cond4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.cond) cond4 = __ldg(reinterpret_cast<const float4*>(p.cond + b * p.cout_phase + ch));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int64_t t = t0 + ty * 8 + i;
        if (t >= p.T) break;
        const int64_t off = (b * p.T + t) * p.N + phase * p.cout_phase + ch;
        float o[4] = {acc[i][0] + bias4.x, acc[i][1] + bias4.y, acc[i][2] + bias4.z, acc[i][3] + bias4.w};
        if (p.cond) {  // after the bias, like the reference's separate add of the 1x1 conv output
            o[0] += cond4.x; o[1] += cond4.y; o[2] += cond4.z; o[3] += cond4.w;
        }
        if (p.residual) {
            const float4 r = *reinterpret_cast<const float4*>(p.residual + off);
            o[0] += r.x; o[1] += r.y; o[2] += r.z; o[3] += r.w;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) o[q] *= p.out_scale;
        if (p.accum) {
            const float4 r = *reinterpret_cast<const float4*>(p.accum + off);
            o[0] += r.x; o[1] += r.y; o[2] += r.z; o[3] += r.w;
        }
        if (p.y_f32) *reinterpret_cast<float4*>(p.y_f32 + off) = make_float4(o[0], o[1], o[2], o[3]);
        if (p.y_bf16) {
            __nv_bfloat162 h0 = __floats2bfloat162_rn(o[0], o[1]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(o[2], o[3]);
            *reinterpret_cast<uint2*>(p.y_bf16 + off) =
                make_uint2(*reinterpret_cast<unsigned int*>(&h0), *reinterpret_cast<unsigned int*>(&h1));
        }
    }
}

int conv_simt_launch(const void* x, const void* w_packed, const float* bias, const float* residual, const float* accum,
                     float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t N, int64_t cout_phase,
                     int taps, int dil, const int* tap_base, int num_phases, float out_scale, cudaStream_t st,
                     const float* cond) {
    BVG_REQUIRE(x && w_packed && (y_f32 || y_bf16), BVG_ERR_INVALID, "conv_simt: null pointer");
    BVG_REQUIRE(Cin % 4 == 0 && cout_phase % 4 == 0 && N == cout_phase * num_phases, BVG_ERR_UNSUPPORTED,
                "conv_simt: Cin=%lld and Cout=%lld must be multiples of 4", (long long)Cin, (long long)cout_phase);
    BVG_REQUIRE(num_phases >= 1 && num_phases <= BVG_MAX_PHASES && taps >= 1, BVG_ERR_UNSUPPORTED, "conv_simt: bad taps/phases");
    BVG_REQUIRE(B < 65536, BVG_ERR_UNSUPPORTED, "conv_simt: B too large");
    if (B == 0 || T == 0) return BVG_OK;
    ConvSimtParams p{};
    p.x = static_cast<const float*>(x); p.w = static_cast<const float*>(w_packed);
    p.bias = bias; p.cond = cond; p.residual = residual; p.accum = accum; p.y_f32 = y_f32;
    p.y_bf16 = static_cast<__nv_bfloat16*>(y_bf16);
    p.T = T; p.Cin = (int)Cin; p.N = (int)N; p.cout_phase = (int)cout_phase; p.taps = taps; p.dil = dil;
    p.num_phases = num_phases; p.tiles_per_phase = (int)ceil_div(cout_phase, SIMT_BN); p.out_scale = out_scale;
    for (int r = 0; r < num_phases; ++r) p.tap_base[r] = tap_base[r];
    dim3 grid((unsigned)ceil_div(T, SIMT_BM), (unsigned)(num_phases * p.tiles_per_phase), (unsigned)B);
    conv_simt_kernel<<<grid, SIMT_THREADS, 0, st>>>(p);
    BVG_LAUNCH_CHECK("conv_simt_kernel");
    return BVG_OK;
}

}  // namespace bvg
