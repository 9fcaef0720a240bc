// Per-utterance conditioning of the v1 generator (reference indextts/BigVGAN/models.py:197-202 cond_layer / conds[i],
// :236 `x = x + self.cond_layer(speaker_embedding)`, :244-245 `x = x + self.conds[i](speaker_embedding)`): the 1x1
// Conv1d over a length-1 sequence is a [C, D] x [D] product per utterance, and the add broadcasts it over time.
#include "bvg_common.cuh"

namespace bvg {

// y[b][c] = bias[c] + sum_d w[c][d] * e[b][d]; one warp per (b, c), lanes stride over d (coalesced rows of w).
__global__ void __launch_bounds__(256)
cond_linear_kernel(const float* __restrict__ e, const float* __restrict__ w, const float* __restrict__ bias,
                   float* __restrict__ y, int64_t B, int64_t D, int64_t C) {
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= B * C) return;
    const int64_t b = warp / C, c = warp % C;
    const float* wr = w + c * D;
    const float* er = e + b * D;
    float acc = 0.f;
    for (int64_t d = lane; d < D; d += 32) acc = fmaf(__ldg(wr + d), __ldg(er + d), acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) y[warp] = acc + (bias ? bias[c] : 0.f);
}

template <typename T>
struct Quad;
template <>
struct Quad<float> {
    static __device__ __forceinline__ float4 load(const float* p) { return *reinterpret_cast<const float4*>(p); }
    static __device__ __forceinline__ void store(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
};
template <>
struct Quad<__nv_bfloat16> {
    static __device__ __forceinline__ float4 load(const __nv_bfloat16* p) {
        const uint2 r = *reinterpret_cast<const uint2*>(p);
        const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&r.x);
        const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&r.y);
        const float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
        return make_float4(fa.x, fa.y, fb.x, fb.y);
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, float4 v) {
        const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        uint2 r;
        r.x = *reinterpret_cast<const uint32_t*>(&a);
        r.y = *reinterpret_cast<const uint32_t*>(&b);
        *reinterpret_cast<uint2*>(p) = r;
    }
};

// y[b][t][c] = x[b][t][c] + v[b][c]  (v == nullptr: dtype conversion only).  One thread per channel quad, rows of one
// utterance walked with a grid stride; x and y may alias when the dtypes match.
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256)
add_channel_vector_kernel(const TIn* x, const float* __restrict__ v, TOut* y, int64_t TC4, int C4) {
    const int64_t b = blockIdx.y;
    const TIn* xb = x + b * TC4 * 4;
    TOut* yb = y + b * TC4 * 4;
    const float* vb = v ? v + b * (int64_t)C4 * 4 : nullptr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < TC4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 a = Quad<TIn>::load(xb + i * 4);
        if (vb) {
            const float4 c = __ldg(reinterpret_cast<const float4*>(vb) + (i % C4));
            a.x += c.x; a.y += c.y; a.z += c.z; a.w += c.w;
        }
        Quad<TOut>::store(yb + i * 4, a);
    }
}

template <typename TIn, typename TOut>
static int launch_add(const void* x, const float* v, void* y, int64_t B, int64_t T, int64_t C, cudaStream_t st) {
    const int64_t tc4 = T * C / 4;
    int64_t gx = ceil_div(tc4, 256 * 4);  // four quads per thread
    const int64_t cap = 148 * 16;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid((unsigned)gx, (unsigned)B);
    add_channel_vector_kernel<TIn, TOut><<<grid, 256, 0, st>>>(static_cast<const TIn*>(x), v, static_cast<TOut*>(y), tc4,
                                                               (int)(C / 4));
    BVG_LAUNCH_CHECK("add_channel_vector_kernel");
    return BVG_OK;
}

int add_channel_vector(const void* x, int x_dtype, const float* v, void* y, int y_dtype, int64_t B, int64_t T, int64_t C,
                       cudaStream_t st) {
    BVG_REQUIRE(B >= 0 && T >= 0 && C > 0 && C % 4 == 0 && B < 65536, BVG_ERR_UNSUPPORTED,
                "add_channel_vector: need C %% 4 == 0 and B < 65536 (B=%lld C=%lld)", (long long)B, (long long)C);
    if (B == 0 || T == 0) return BVG_OK;
    BVG_REQUIRE(x && y, BVG_ERR_INVALID, "add_channel_vector: null pointer");
    if (x_dtype == BVG_F32 && y_dtype == BVG_F32) return launch_add<float, float>(x, v, y, B, T, C, st);
    if (x_dtype == BVG_F32 && y_dtype == BVG_BF16) return launch_add<float, __nv_bfloat16>(x, v, y, B, T, C, st);
    if (x_dtype == BVG_BF16 && y_dtype == BVG_BF16) return launch_add<__nv_bfloat16, __nv_bfloat16>(x, v, y, B, T, C, st);
    if (x_dtype == BVG_BF16 && y_dtype == BVG_F32) return launch_add<__nv_bfloat16, float>(x, v, y, B, T, C, st);
    return set_error(BVG_ERR_UNSUPPORTED, "add_channel_vector: unsupported dtypes %d -> %d", x_dtype, y_dtype);
}

}  // namespace bvg

using namespace bvg;

extern "C" int bvg_cond_linear_fwd(const float* e, const float* w, const float* bias, float* y, int64_t B, int64_t D,
                                   int64_t C, void* stream) {
    BVG_REQUIRE(B >= 0 && D > 0 && C > 0 && B * C < (int64_t(1) << 31) / 32, BVG_ERR_INVALID, "bvg_cond_linear_fwd: bad shape");
    if (B == 0) return BVG_OK;  // an empty batch has no storage behind its pointers
    BVG_REQUIRE(e && w && y, BVG_ERR_INVALID, "bvg_cond_linear_fwd: null pointer");
    const unsigned grid = (unsigned)ceil_div(B * C * 32, 256);
    cond_linear_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(e, w, bias, y, B, D, C);
    BVG_LAUNCH_CHECK("cond_linear_kernel");
    return BVG_OK;
}

extern "C" int bvg_add_channel_vector_nlc(const void* x, int x_dtype, const float* v, void* y, int y_dtype, int64_t B,
                                          int64_t T, int64_t C, void* stream) {
    return add_channel_vector(x, x_dtype, v, y, y_dtype, B, T, C, static_cast<cudaStream_t>(stream));
}
