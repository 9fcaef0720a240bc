// Weight preparation (weight-norm fold, tap-major / polyphase packing), layout changes at the path's edges and the
// fused conv_post + clamp/tanh tail.
#include "bvg_common.cuh"

namespace bvg {

// ------------------------------------------------------------------------------------------------------------
// w[r,:] = g[r] * v[r,:] / ||v[r,:]||   (torch._weight_norm(v, g, 0); reference bigvgan.py:388-400)
// ------------------------------------------------------------------------------------------------------------
__global__ void fold_weight_norm_kernel(const float* __restrict__ v, const float* __restrict__ g, float* __restrict__ w,
                                        int64_t row_len) {
    const int64_t r = blockIdx.x;
    const float* vr = v + r * row_len;
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < row_len; i += blockDim.x) {
        const double a = vr[i];
        s += a * a;
    }
    __shared__ double red[32];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) red[0] = s;
    }
    __syncthreads();
    const float scale = g[r] / static_cast<float>(sqrt(red[0]));
    for (int64_t i = threadIdx.x; i < row_len; i += blockDim.x) w[r * row_len + i] = vr[i] * scale;
}

template <typename TOut>
__device__ __forceinline__ TOut cvt_out(float v);
template <>
__device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// out[j][co][ci] = w[co][ci][j]
template <typename TOut>
__global__ void pack_conv1d_kernel(const float* __restrict__ w, TOut* __restrict__ out, int64_t Cout, int64_t Cin,
                                   int64_t k) {
    const int64_t total = Cout * Cin * k;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ci = i % Cin;
        const int64_t co = (i / Cin) % Cout;
        const int64_t j = i / (Cin * Cout);
        out[i] = cvt_out<TOut>(w[(co * Cin + ci) * k + j]);
    }
}

// out[j][r*Cout + co][ci] = w[ci][co][(r+p)%u + u*(K/u-1-j)],  p = (K-u)/2
template <typename TOut>
__global__ void pack_convt_kernel(const float* __restrict__ w, TOut* __restrict__ out, int64_t Cin, int64_t Cout,
                                  int64_t K, int64_t u) {
    const int64_t taps = K / u, p = (K - u) / 2;
    const int64_t total = taps * u * Cout * Cin;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ci = i % Cin;
        const int64_t n = (i / Cin) % (u * Cout);
        const int64_t j = i / (Cin * u * Cout);
        const int64_t r = n / Cout, co = n % Cout;
        const int64_t kk = (r + p) % u + u * (taps - 1 - j);
        out[i] = cvt_out<TOut>(w[(ci * Cout + co) * K + kk]);
    }
}

// bf16x3 split forms of the two packings (bvg_common.cuh): out[j][n][t*Cin + ci] = term t of the same element
__global__ void pack_conv1d_split_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int64_t Cout,
                                         int64_t Cin, int64_t k) {
    const int64_t total = Cout * Cin * k;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ci = i % Cin;
        const int64_t co = (i / Cin) % Cout;
        const int64_t j = i / (Cin * Cout);
        const Bf16x3 s = split_bf16x3(w[(co * Cin + ci) * k + j]);
        __nv_bfloat16* o = out + (j * Cout + co) * (BVG_SPLIT_TERMS * Cin) + ci;
#pragma unroll
        for (int t = 0; t < BVG_SPLIT_TERMS; ++t) o[t * Cin] = split_w_term(s, t);
    }
}

__global__ void pack_convt_split_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int64_t Cin,
                                        int64_t Cout, int64_t K, int64_t u) {
    const int64_t taps = K / u, p = (K - u) / 2;
    const int64_t total = taps * u * Cout * Cin;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ci = i % Cin;
        const int64_t n = (i / Cin) % (u * Cout);
        const int64_t j = i / (Cin * u * Cout);
        const int64_t r = n / Cout, co = n % Cout;
        const int64_t kk = (r + p) % u + u * (taps - 1 - j);
        const Bf16x3 s = split_bf16x3(w[(ci * Cout + co) * K + kk]);
        __nv_bfloat16* o = out + (j * u * Cout + n) * (BVG_SPLIT_TERMS * Cin) + ci;
#pragma unroll
        for (int t = 0; t < BVG_SPLIT_TERMS; ++t) o[t * Cin] = split_w_term(s, t);
    }
}

// fp32 [rows, C] -> bf16 [rows, 6*C] activation split (the tensors a conv reads that no Activation1d produced: the mel,
// conv_pre's output, the stage sums).  One thread per channel quad; 16-byte loads, 8-byte stores.
__global__ void split_rows_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t rows, int C) {
    const int64_t quads = rows * (C / 4);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < quads; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / (C / 4);
        const int c = static_cast<int>(i - r * (C / 4)) * 4;
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + r * C + c));
        const Bf16x3 s0 = split_bf16x3(v.x), s1 = split_bf16x3(v.y), s2 = split_bf16x3(v.z), s3 = split_bf16x3(v.w);
        __nv_bfloat16* o = y + r * (int64_t)(BVG_SPLIT_TERMS * C) + c;
#pragma unroll
        for (int t = 0; t < BVG_SPLIT_TERMS; ++t) {
            __nv_bfloat162 a = __halves2bfloat162(split_act_term(s0, t), split_act_term(s1, t));
            __nv_bfloat162 b = __halves2bfloat162(split_act_term(s2, t), split_act_term(s3, t));
            uint2 pk;
            pk.x = *reinterpret_cast<unsigned int*>(&a);
            pk.y = *reinterpret_cast<unsigned int*>(&b);
            *reinterpret_cast<uint2*>(o + (int64_t)t * C) = pk;
        }
    }
}

int split_rows(const float* x, void* y, int64_t rows, int64_t C, cudaStream_t st) {
    BVG_REQUIRE(x && y, BVG_ERR_INVALID, "split_bf16x3: null pointer");
    BVG_REQUIRE(rows >= 0 && C > 0 && C % 4 == 0 && C < (int64_t(1) << 28), BVG_ERR_INVALID,
                "split_bf16x3: bad shape rows=%lld C=%lld (C %% 4 == 0)", (long long)rows, (long long)C);
    BVG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 8 == 0, BVG_ERR_INVALID,
                "split_bf16x3: x must be 16-byte and y 8-byte aligned");
    if (rows == 0) return BVG_OK;
    const int64_t quads = rows * (C / 4);
    const unsigned grid = (unsigned)(ceil_div(quads, 256) < 148 * 16 ? ceil_div(quads, 256) : 148 * 16);
    split_rows_kernel<<<grid, 256, 0, st>>>(x, static_cast<__nv_bfloat16*>(y), rows, (int)C);
    BVG_LAUNCH_CHECK("split_rows_kernel");
    return BVG_OK;
}

// ------------------------------------------------------------------------------------------------------------
// [B,C,T] <-> [B,T,C] with dtype conversion, 32x32 smem tiles
// ------------------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float ld_as_float(const T* p);
template <>
__device__ __forceinline__ float ld_as_float<float>(const float* p) { return *p; }
template <>
__device__ __forceinline__ float ld_as_float<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// src[b][r][c] (row stride src_sr, batch stride src_sb, c contiguous) -> dst[b][c][r] dense.  R rows, Cc cols.
template <typename TIn, typename TOut>
__global__ void transpose_kernel(const TIn* __restrict__ src, TOut* __restrict__ dst, int64_t R, int64_t Cc,
                                 int64_t src_sb, int64_t src_sr) {
    __shared__ float tile[32][33];
    const int64_t b = blockIdx.y;
    const int64_t tiles_c = (Cc + 31) / 32;
    const int64_t c0 = ((int64_t)blockIdx.x % tiles_c) * 32, r0 = ((int64_t)blockIdx.x / tiles_c) * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int64_t r = r0 + i, c = c0 + threadIdx.x;
        if (r < R && c < Cc) tile[i][threadIdx.x] = ld_as_float<TIn>(src + b * src_sb + r * src_sr + c);
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int64_t c = c0 + i, r = r0 + threadIdx.x;
        if (r < R && c < Cc) dst[(b * Cc + c) * R + r] = cvt_out<TOut>(tile[threadIdx.x][i]);
    }
}

template <typename TIn, typename TOut>
static int launch_transpose(const void* src, void* dst, int64_t B, int64_t R, int64_t Cc, int64_t sb, int64_t sr,
                            cudaStream_t st) {
    const int64_t tiles = ceil_div(Cc, 32) * ceil_div(R, 32);
    BVG_REQUIRE(tiles < (int64_t(1) << 31) && B < 65536, BVG_ERR_UNSUPPORTED, "transpose: shape too large");
    dim3 grid((unsigned)tiles, (unsigned)B);
    transpose_kernel<TIn, TOut><<<grid, dim3(32, 8), 0, st>>>(static_cast<const TIn*>(src), static_cast<TOut*>(dst), R, Cc,
                                                              sb, sr);
    BVG_LAUNCH_CHECK("transpose_kernel");
    return BVG_OK;
}

int transpose_any(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t B, int64_t R, int64_t Cc, int64_t sb,
                  int64_t sr, cudaStream_t st) {
    if (B == 0 || R == 0 || Cc == 0) return BVG_OK;
    if (src_dtype == BVG_F32 && dst_dtype == BVG_F32) return launch_transpose<float, float>(src, dst, B, R, Cc, sb, sr, st);
    if (src_dtype == BVG_F32 && dst_dtype == BVG_BF16)
        return launch_transpose<float, __nv_bfloat16>(src, dst, B, R, Cc, sb, sr, st);
    if (src_dtype == BVG_BF16 && dst_dtype == BVG_F32)
        return launch_transpose<__nv_bfloat16, float>(src, dst, B, R, Cc, sb, sr, st);
    if (src_dtype == BVG_BF16 && dst_dtype == BVG_BF16)
        return launch_transpose<__nv_bfloat16, __nv_bfloat16>(src, dst, B, R, Cc, sb, sr, st);
    return set_error(BVG_ERR_UNSUPPORTED, "transpose: unsupported dtypes %d -> %d", src_dtype, dst_dtype);
}

// ------------------------------------------------------------------------------------------------------------
// conv_post (Cin -> 1, k taps, zero padding (k-1)/2) + clamp(-1,1) | tanh     (reference bigvgan.py:379-384)
// ------------------------------------------------------------------------------------------------------------
constexpr int POST_MAX_W = 1024;  // Cin * k

// TOut = float: the module's output.  TOut = int16_t: the caller's PCM conversion fused in (reference infer_v2.py:866,
// 912: clamp(32767 * wav, -32767, 32767).type(int16), i.e. truncation toward zero).
template <typename TOut>
__device__ __forceinline__ TOut post_convert(float o);
template <>
__device__ __forceinline__ float post_convert<float>(float o) { return o; }
template <>
__device__ __forceinline__ int16_t post_convert<int16_t>(float o) {
    const float p = fminf(fmaxf(32767.0f * o, -32767.0f), 32767.0f);
    return static_cast<int16_t>(__float2int_rz(p));
}

template <typename TOut>
__global__ void __launch_bounds__(256)
conv_post_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                 TOut* __restrict__ y, int64_t T, int Cin, int k, int use_tanh) {
    __shared__ float ws[POST_MAX_W];  // ws[j*Cin + c] = w[c*k + j]
    for (int i = threadIdx.x; i < Cin * k; i += blockDim.x) {
        const int j = i / Cin, c = i % Cin;
        ws[i] = w[c * k + j];
    }
    __syncthreads();
    const int64_t b = blockIdx.y;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    const int pad = (k - 1) / 2;
    const float* xb = x + b * T * Cin;
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
    for (int j = 0; j < k; ++j) {
        const int64_t r = t + j - pad;
        if (r < 0 || r >= T) continue;
        const float4* xr = reinterpret_cast<const float4*>(xb + r * Cin);
        const float4* wr = reinterpret_cast<const float4*>(ws + j * Cin);
        for (int c = 0; c < Cin / 4; ++c) {
            const float4 a = __ldg(xr + c);
            const float4 ww = wr[c];
            acc0 = fmaf(a.x, ww.x, acc0);
            acc1 = fmaf(a.y, ww.y, acc1);
            acc2 = fmaf(a.z, ww.z, acc2);
            acc3 = fmaf(a.w, ww.w, acc3);
        }
    }
    float o = (acc0 + acc1) + (acc2 + acc3);
    if (bias) o += bias[0];
    o = use_tanh ? tanhf(o) : fminf(fmaxf(o, -1.0f), 1.0f);
    y[b * T + t] = post_convert<TOut>(o);
}

// Tiled version for the generator's own tail (Cin = 24, k = 7 in bigvgan_v2_22khz_80band_256x).  The one-thread-per-
// sample kernel above reads 7 rows x 96 B per thread with a 96-byte lane stride: 24 L1 wavefronts per load instruction,
// 230 us at cfg2 against a 52 us HBM floor.  Here a CTA stages NT*R + K - 1 channels-last rows in shared memory with
// coalesced 16-byte loads (row stride CIN + 1 words: with R odd the lane stride R*(CIN+1) is odd, so scalar LDS are
// conflict free), every thread produces R consecutive samples from a sliding window of R + K - 1 rows (weights of one
// channel quad live in registers across the window), and the results leave through shared memory as coalesced stores.
// TOut = float: the module's output.  TOut = int16_t: the caller's PCM conversion (reference infer_v2.py:866, 912:
// clamp(32767 * wav, -32767, 32767).type(int16), truncation toward zero) fused in.

template <int CIN, int K, int R, int NT, typename TOut>
__global__ void __launch_bounds__(NT)
conv_post_tile_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                      TOut* __restrict__ y, int64_t T, int use_tanh) {
    constexpr int TILE = NT * R, PAD = (K - 1) / 2, ROWS = TILE + K - 1, S = CIN + 1, Q = CIN / 4;
    constexpr int XS_WORDS = (ROWS * S + 3) / 4 * 4;
    extern __shared__ __align__(16) float post_smem[];
    float* xs = post_smem;             // [ROWS][S]
    float* wsm = post_smem + XS_WORDS;  // [K][CIN]: wsm[j*CIN + c] = w[c*K + j]
    const int tid = threadIdx.x;
    const int64_t b = blockIdx.y;
    const int64_t t0 = (int64_t)blockIdx.x * TILE;
    for (int i = tid; i < CIN * K; i += NT) wsm[i] = w[(i % CIN) * K + i / CIN];
    {
        const float* xb = x + b * T * CIN;
        const int64_t row_lo = t0 - PAD;
        constexpr int NV = ROWS * Q, U = 4;  // U independent 16-byte loads in flight per thread before the first store
        for (int i0 = tid; i0 < NV; i0 += U * NT) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * NT;
                const int row = i / Q, q = i - row * Q;
                const int64_t gr = row_lo + row;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);  // zero padding of the conv outside [0, T)
                if (i < NV && gr >= 0 && gr < T) v[u] = __ldg(reinterpret_cast<const float4*>(xb + gr * CIN) + q);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + u * NT;
                if (i < NV) {
                    const int row = i / Q, q = i - row * Q;
                    float* d = xs + row * S + 4 * q;
                    d[0] = v[u].x; d[1] = v[u].y; d[2] = v[u].z; d[3] = v[u].w;
                }
            }
        }
    }
    __syncthreads();
    float acc[R];
#pragma unroll
    for (int o = 0; o < R; ++o) acc[o] = 0.f;
    const float* xr = xs + (R * tid) * S;
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        float4 wv[K];
#pragma unroll
        for (int j = 0; j < K; ++j) wv[j] = *reinterpret_cast<const float4*>(wsm + j * CIN + 4 * q);
#pragma unroll
        for (int rr = 0; rr < R + K - 1; ++rr) {
            const float a0 = xr[rr * S + 4 * q], a1 = xr[rr * S + 4 * q + 1], a2 = xr[rr * S + 4 * q + 2],
                        a3 = xr[rr * S + 4 * q + 3];
#pragma unroll
            for (int j = 0; j < K; ++j) {
                const int o = rr - j;  // window row rr is tap j of output o
                if (o >= 0 && o < R) acc[o] = fmaf(a3, wv[j].w, fmaf(a2, wv[j].z, fmaf(a1, wv[j].y, fmaf(a0, wv[j].x, acc[o]))));
            }
        }
    }
    __syncthreads();  // everyone is done reading xs: reuse it as the output staging row
    const float bv = bias ? bias[0] : 0.f;
#pragma unroll
    for (int o = 0; o < R; ++o) {
        float v = acc[o] + bv;
        v = use_tanh ? tanhf(v) : fminf(fmaxf(v, -1.0f), 1.0f);
        xs[R * tid + o] = v;
    }
    __syncthreads();
    TOut* yb = y + b * T + t0;
    for (int i = tid; i < TILE; i += NT)
        if (t0 + i < T) yb[i] = post_convert<TOut>(xs[i]);
}

template <typename TOut>
static int launch_conv_post(const float* x, const float* w, const float* bias, TOut* y, int64_t B, int64_t T, int64_t Cin,
                            int k, int use_tanh, cudaStream_t st) {
    if (Cin == 24 && k == 7) {
        constexpr int CIN = 24, K = 7, R = 3, NT = 128;  // 39 KB of shared memory per CTA: 5 CTAs per SM
        constexpr int XS_WORDS = ((NT * R + K - 1) * (CIN + 1) + 3) / 4 * 4;
        constexpr int smem = (XS_WORDS + K * CIN) * 4;
        auto kern = conv_post_tile_kernel<CIN, K, R, NT, TOut>;
        static bool configured = false;  // idempotent attribute; a race only repeats the call
        if (!configured) {
            BVG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            configured = true;
        }
        dim3 grid((unsigned)ceil_div(T, (int64_t)NT * R), (unsigned)B);
        kern<<<grid, NT, smem, st>>>(x, w, bias, y, T, use_tanh);
        BVG_LAUNCH_CHECK("conv_post_tile_kernel");
        return BVG_OK;
    }
    dim3 grid((unsigned)ceil_div(T, 256), (unsigned)B);
    conv_post_kernel<TOut><<<grid, 256, 0, st>>>(x, w, bias, y, T, (int)Cin, k, use_tanh);
    BVG_LAUNCH_CHECK("conv_post_kernel");
    return BVG_OK;
}

}  // namespace bvg

using namespace bvg;

extern "C" int bvg_fold_weight_norm(const float* v, const float* g, float* w, int64_t rows, int64_t row_len,
                                    void* stream) {
    BVG_REQUIRE(v && g && w, BVG_ERR_INVALID, "bvg_fold_weight_norm: null pointer");
    BVG_REQUIRE(rows >= 0 && row_len > 0 && rows < (int64_t(1) << 31), BVG_ERR_INVALID, "bvg_fold_weight_norm: bad shape");
    if (rows == 0) return BVG_OK;
    fold_weight_norm_kernel<<<(unsigned)rows, 256, 0, static_cast<cudaStream_t>(stream)>>>(v, g, w, row_len);
    BVG_LAUNCH_CHECK("fold_weight_norm_kernel");
    return BVG_OK;
}

extern "C" int bvg_pack_conv1d_weight(const float* w, void* out, int64_t Cout, int64_t Cin, int64_t k, int dtype,
                                      void* stream) {
    BVG_REQUIRE(w && out, BVG_ERR_INVALID, "bvg_pack_conv1d_weight: null pointer");
    BVG_REQUIRE(Cout > 0 && Cin > 0 && k > 0, BVG_ERR_INVALID, "bvg_pack_conv1d_weight: bad shape");
    const int64_t total = Cout * Cin * k;
    const unsigned grid = (unsigned)(ceil_div(total, 256) < 4096 ? ceil_div(total, 256) : 4096);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == BVG_BF16)
        pack_conv1d_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(w, static_cast<__nv_bfloat16*>(out), Cout, Cin, k);
    else if (dtype == BVG_F32)
        pack_conv1d_kernel<float><<<grid, 256, 0, st>>>(w, static_cast<float*>(out), Cout, Cin, k);
    else if (dtype == BVG_BF16X3)
        pack_conv1d_split_kernel<<<grid, 256, 0, st>>>(w, static_cast<__nv_bfloat16*>(out), Cout, Cin, k);
    else
        return set_error(BVG_ERR_UNSUPPORTED, "bvg_pack_conv1d_weight: dtype %d", dtype);
    BVG_LAUNCH_CHECK("pack_conv1d_kernel");
    return BVG_OK;
}

extern "C" int bvg_pack_convtranspose1d_weight(const float* w, void* out, int64_t Cin, int64_t Cout, int64_t K,
                                               int64_t u, int dtype, void* stream) {
    BVG_REQUIRE(w && out, BVG_ERR_INVALID, "bvg_pack_convtranspose1d_weight: null pointer");
    BVG_REQUIRE(Cout > 0 && Cin > 0 && K > 0 && u > 0, BVG_ERR_INVALID, "bvg_pack_convtranspose1d_weight: bad shape");
    BVG_REQUIRE(K % u == 0 && (K - u) % 2 == 0 && u <= BVG_MAX_PHASES, BVG_ERR_UNSUPPORTED,
                "bvg_pack_convtranspose1d_weight: need K %% u == 0, (K-u) even, u <= %d (got K=%lld u=%lld)",
                BVG_MAX_PHASES, (long long)K, (long long)u);
    const int64_t total = Cout * Cin * K;
    const unsigned grid = (unsigned)(ceil_div(total, 256) < 4096 ? ceil_div(total, 256) : 4096);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == BVG_BF16)
        pack_convt_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(w, static_cast<__nv_bfloat16*>(out), Cin, Cout, K, u);
    else if (dtype == BVG_F32)
        pack_convt_kernel<float><<<grid, 256, 0, st>>>(w, static_cast<float*>(out), Cin, Cout, K, u);
    else if (dtype == BVG_BF16X3)
        pack_convt_split_kernel<<<grid, 256, 0, st>>>(w, static_cast<__nv_bfloat16*>(out), Cin, Cout, K, u);
    else
        return set_error(BVG_ERR_UNSUPPORTED, "bvg_pack_convtranspose1d_weight: dtype %d", dtype);
    BVG_LAUNCH_CHECK("pack_convt_kernel");
    return BVG_OK;
}

extern "C" int bvg_split_bf16x3_nlc(const float* x, void* y, int64_t rows, int64_t C, void* stream) {
    return split_rows(x, y, rows, C, static_cast<cudaStream_t>(stream));
}

extern "C" int bvg_ncl_to_nlc(const void* x, int x_dtype, int64_t x_stride_b, int64_t x_stride_c, void* y, int y_dtype,
                              int64_t B, int64_t C, int64_t T, void* stream) {
    BVG_REQUIRE(x && y, BVG_ERR_INVALID, "bvg_ncl_to_nlc: null pointer");
    // src rows = channels (stride x_stride_c), cols = time; dst [B][T][C]
    return transpose_any(x, x_dtype, y, y_dtype, B, C, T, x_stride_b, x_stride_c, static_cast<cudaStream_t>(stream));
}

extern "C" int bvg_nlc_to_ncl(const void* x, int x_dtype, void* y, int y_dtype, int64_t B, int64_t C, int64_t T,
                              void* stream) {
    BVG_REQUIRE(x && y, BVG_ERR_INVALID, "bvg_nlc_to_ncl: null pointer");
    // src rows = time (stride C), cols = channels; dst [B][C][T]
    return transpose_any(x, x_dtype, y, y_dtype, B, T, C, T * C, C, static_cast<cudaStream_t>(stream));
}

static int conv_post_checks(const void* x, const void* w, const void* y, int64_t B, int64_t Cin, int k, const char* who) {
    BVG_REQUIRE(x && w && y, BVG_ERR_INVALID, "%s: null pointer", who);
    BVG_REQUIRE(Cin % 4 == 0 && Cin * k <= POST_MAX_W && k % 2 == 1, BVG_ERR_UNSUPPORTED, "%s: Cin=%lld k=%d unsupported", who,
                (long long)Cin, k);
    BVG_REQUIRE(B < 65536, BVG_ERR_UNSUPPORTED, "%s: B too large", who);
    return BVG_OK;
}

extern "C" int bvg_conv_post_fwd(const float* x, const float* w, const float* bias, float* y, int64_t B, int64_t T,
                                 int64_t Cin, int k, int use_tanh, void* stream) {
    const int rc = conv_post_checks(x, w, y, B, Cin, k, "bvg_conv_post_fwd");
    if (rc != BVG_OK) return rc;
    if (B == 0 || T == 0) return BVG_OK;
    return launch_conv_post<float>(x, w, bias, y, B, T, Cin, k, use_tanh, static_cast<cudaStream_t>(stream));
}

extern "C" int bvg_conv_post_pcm16_fwd(const float* x, const float* w, const float* bias, int16_t* pcm, int64_t B, int64_t T,
                                       int64_t Cin, int k, int use_tanh, void* stream) {
    const int rc = conv_post_checks(x, w, pcm, B, Cin, k, "bvg_conv_post_pcm16_fwd");
    if (rc != BVG_OK) return rc;
    if (B == 0 || T == 0) return BVG_OK;
    return launch_conv_post<int16_t>(x, w, bias, pcm, B, T, Cin, k, use_tanh, static_cast<cudaStream_t>(stream));
}
