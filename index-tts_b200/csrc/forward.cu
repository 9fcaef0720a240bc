// C-ABI entry points for the convolutions and the whole-generator launch plan (reference bigvgan.py:360-386).
#include "bvg_common.cuh"

#include <cstdlib>
#include <atomic>
#include <vector>

namespace bvg {
int conv_umma_launch(const void* x, const void* w_packed, const float* bias, const float* residual, const float* accum,
                     float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t N, int64_t cout_phase,
                     int taps, int dil, const int* tap_base, int num_phases, float out_scale, cudaStream_t st,
                     const float* cond, unsigned int* tile_ctr);
int conv_simt_launch(const void* x, const void* w_packed, const float* bias, const float* residual, const float* accum,
                     float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t N, int64_t cout_phase,
                     int taps, int dil, const int* tap_base, int num_phases, float out_scale, cudaStream_t st,
                     const float* cond);
int activation1d_nlc(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha, const float* beta,
                     int64_t B, int64_t T, int64_t C, int precise, cudaStream_t st);
int activation1d_nlc_multi(const void* x, int x_dtype, void* const* ys, int y_dtype, const float* const* alphas,
                           const float* const* betas, int n_heads, int64_t B, int64_t T, int64_t C, int precise,
                           cudaStream_t st);
int transpose_any(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t B, int64_t R, int64_t Cc, int64_t sb,
                  int64_t sr, cudaStream_t st);
int add_channel_vector(const void* x, int x_dtype, const float* v, void* y, int y_dtype, int64_t B, int64_t T, int64_t C,
                       cudaStream_t st);
int split_rows(const float* x, void* y, int64_t rows, int64_t C, cudaStream_t st);

// Tile counters for the conv kernel's dynamic scheduler: forward_impl carves an array of zeroed counter pairs out of the
// caller's workspace and every tensor-core conv launch of that call takes the next one (the stand-alone conv entry points
// have no workspace and keep static tile striding).  The cursor is per host thread and only set while a forward runs.
constexpr int BVG_MAX_TILE_COUNTERS = 1024;
struct TileCounterCursor {
    unsigned int* base = nullptr;
    int next = 0;
};
static thread_local TileCounterCursor t_tile_ctr;
static unsigned int* take_tile_counter() {
    if (t_tile_ctr.base == nullptr || t_tile_ctr.next >= BVG_MAX_TILE_COUNTERS) return nullptr;
    return t_tile_ctr.base + 2 * (t_tile_ctr.next++);
}
struct TileCounterScope {  // clears the cursor on every exit path of forward_impl
    ~TileCounterScope() { t_tile_ctr = TileCounterCursor(); }
};

static int conv_dispatch(int precision, const void* x, const void* w, const float* bias, const float* residual,
                         const float* accum, float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t N,
                         int64_t cout_phase, int taps, int dil, const int* tap_base, int num_phases, float out_scale,
                         cudaStream_t st, const float* cond = nullptr) {
    if (precision == BVG_BF16_TC)
        return conv_umma_launch(x, w, bias, residual, accum, y_f32, y_bf16, B, T, Cin, N, cout_phase, taps, dil, tap_base,
                                num_phases, out_scale, st, cond, take_tile_counter());
    if (precision == BVG_F32_SPLIT_TC) {
        // fp32 parity mode: x and w are bf16x3 split tensors (six term slots of Cin channels each on the K axis of the
        // same kernel, bvg_common.cuh); fp32 accumulation in TMEM, fp32 outputs only
        BVG_REQUIRE(y_bf16 == nullptr && y_f32 != nullptr, BVG_ERR_INVALID, "split-operand conv: the output is fp32");
        return conv_umma_launch(x, w, bias, residual, accum, y_f32, nullptr, B, T, Cin * BVG_SPLIT_TERMS, N, cout_phase, taps,
                                dil, tap_base, num_phases, out_scale, st, cond, take_tile_counter());
    }
    if (precision == BVG_EXACT_F32)
        return conv_simt_launch(x, w, bias, residual, accum, y_f32, y_bf16, B, T, Cin, N, cout_phase, taps, dil, tap_base,
                                num_phases, out_scale, st, cond);
    return set_error(BVG_ERR_UNSUPPORTED, "unknown precision mode %d", precision);
}

static int conv1d(int precision, const void* x, const void* w, const float* bias, const float* residual,
                  const float* accum, float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t Cout, int k,
                  int dil, float out_scale, cudaStream_t st, const float* cond = nullptr) {
    BVG_REQUIRE(k % 2 == 1, BVG_ERR_UNSUPPORTED, "conv1d: kernel size %d must be odd ('same' padding)", k);
    const int base = -((k - 1) / 2) * dil;
    return conv_dispatch(precision, x, w, bias, residual, accum, y_f32, y_bf16, B, T, Cin, Cout, Cout, k, dil, &base, 1,
                         out_scale, st, cond);
}

// AMP-block conv through its descriptor: plain, or time-folded (bvg.h, bvg_conv_desc::fold) in tensor-core mode.
static int conv1d_desc(int precision, const bvg_conv_desc& c, const void* x, const float* residual, const float* accum,
                       float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t C, float out_scale, cudaStream_t st) {
    const int r = c.fold;
    if (r <= 1 || precision != BVG_BF16_TC)
        return conv1d(precision, x, c.w, c.bias, residual, accum, y_f32, y_bf16, B, T, C, C, c.k, c.dilation, out_scale, st);
    BVG_REQUIRE(T % r == 0, BVG_ERR_INVALID, "conv1d: time folding by %d needs T %% %d == 0 (T = %lld)", r, r, (long long)T);
    const int reach = (c.k - 1) / 2 * c.dilation;
    const int taps = 2 * ((reach + r - 1) / r) + 1;
    return conv1d(precision, x, c.w, c.bias, residual, accum, y_f32, y_bf16, B, T / r, C * r, C * r, taps, 1, out_scale, st);
}

static int convtranspose1d(int precision, const void* x, const void* w, const float* bias, float* y_f32, void* y_bf16,
                           int64_t B, int64_t Tin, int64_t Cin, int64_t Cout, int K, int u, cudaStream_t st,
                           const float* cond = nullptr) {
    BVG_REQUIRE(u >= 1 && u <= BVG_MAX_PHASES && K % u == 0 && (K - u) % 2 == 0, BVG_ERR_UNSUPPORTED,
                "convtranspose1d: need K %% u == 0, (K-u) even, u <= %d (got K=%d u=%d)", BVG_MAX_PHASES, K, u);
    const int taps = K / u, p = (K - u) / 2;
    int base[BVG_MAX_PHASES];
    for (int r = 0; r < u; ++r) base[r] = (r + p) / u - (taps - 1);
    return conv_dispatch(precision, x, w, bias, nullptr, nullptr, y_f32, y_bf16, B, Tin, Cin, (int64_t)u * Cout, Cout, taps,
                         1, base, u, 1.0f, st, cond);
}

// ------------------------------------------------------------------------------------------------------------
// whole generator
// ------------------------------------------------------------------------------------------------------------
struct Plan {
    int64_t B, T_mel, E;  // E = max over stages of B*T_i*C_i
    size_t op_es;         // bytes per MMA-operand element
    size_t off_mel, off_pre, off_x, off_sum, off_sumop, off_ctr, total;
    size_t off_r[BVG_MAX_KERNELS], off_a[BVG_MAX_KERNELS], off_y1[BVG_MAX_KERNELS];  // per AMP block (they run concurrently)
};

static size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

// ---- optional per-launch timing (bvg_forward_profile) -------------------------------------------------------
struct ProfEntry {
    bvg_launch_record rec;
    cudaEvent_t e0, e1;
};
static thread_local std::vector<ProfEntry>* g_prof = nullptr;

struct ProfScope {
    cudaStream_t st;
    bool on;
    ProfScope(int kind, int stage, int cin, int cout, int k, int dil, int64_t rows, double flops, double bytes,
              cudaStream_t s)
        : st(s), on(g_prof != nullptr) {
        if (!on) return;
        ProfEntry e{};
        e.rec.kind = kind; e.rec.stage = stage; e.rec.cin = cin; e.rec.cout = cout; e.rec.k = k; e.rec.dilation = dil;
        e.rec.rows = rows; e.rec.flops = flops; e.rec.bytes = bytes; e.rec.ms = 0.f;
        cudaEventCreate(&e.e0);
        cudaEventCreate(&e.e1);
        cudaEventRecord(e.e0, st);
        g_prof->push_back(e);
    }
    ~ProfScope() {
        if (on) cudaEventRecord(g_prof->back().e1, st);
    }
};

static int make_plan(const bvg_model_desc* m, int64_t B, int64_t T_mel, Plan* pl) {
    BVG_REQUIRE(m != nullptr, BVG_ERR_INVALID, "null model descriptor");
    BVG_REQUIRE(m->num_stages >= 1 && m->num_stages <= BVG_MAX_STAGES && m->num_kernels >= 1 &&
                    m->num_kernels <= BVG_MAX_KERNELS && m->num_dilations >= 1 && m->num_dilations <= BVG_MAX_DILATIONS,
                BVG_ERR_UNSUPPORTED, "model descriptor: stages/kernels/dilations out of range");
    BVG_REQUIRE(m->precision == BVG_EXACT_F32 || m->precision == BVG_BF16_TC || m->precision == BVG_F32_SPLIT_TC,
                BVG_ERR_UNSUPPORTED, "model descriptor: unknown precision %d", m->precision);
    BVG_REQUIRE(m->num_mels % 8 == 0 && (m->upsample_initial_channel >> m->num_stages) % 8 == 0, BVG_ERR_UNSUPPORTED,
                "channel counts must be multiples of 8 (num_mels=%d, last stage %d)", m->num_mels,
                m->upsample_initial_channel >> m->num_stages);
    pl->B = B;
    pl->T_mel = T_mel;
    // bytes per conv-operand element: bf16, bf16x3 split (six term slots), fp32
    pl->op_es = m->precision == BVG_BF16_TC ? 2 : (m->precision == BVG_F32_SPLIT_TC ? 2 * BVG_SPLIT_TERMS : 4);
    const size_t y1_es = m->precision == BVG_BF16_TC ? 2 : 4;  // conv1 -> activation intermediate
    BVG_REQUIRE(m->resblock >= 0 && m->resblock <= 2, BVG_ERR_UNSUPPORTED, "model descriptor: unknown resblock type %d",
                m->resblock);
    int64_t T = T_mel, E = B * T_mel * m->upsample_initial_channel;
    for (int i = 0; i < m->num_stages; ++i) {
        T *= m->upsample_rates[i];
        const int64_t C = m->upsample_initial_channel >> (i + 1);
        E = E > B * T * C ? E : B * T * C;
    }
    pl->E = E;
    size_t o = 0;
    pl->off_mel = o;   o += align256((size_t)B * T_mel * m->num_mels * pl->op_es);
    pl->off_pre = o;   o += align256((size_t)B * T_mel * m->upsample_initial_channel * pl->op_es);
    pl->off_x = o;     o += align256((size_t)E * 4);
    pl->off_sum = o;   o += align256((size_t)E * 4);
    pl->off_sumop = o; o += align256((size_t)E * pl->op_es);
    for (int j = 0; j < m->num_kernels; ++j) {
        pl->off_r[j] = o;  o += align256((size_t)E * 4);
        pl->off_a[j] = o;  o += align256((size_t)E * pl->op_es);
        pl->off_y1[j] = o; o += align256((size_t)E * y1_es);
    }
    pl->off_ctr = o;   o += align256((size_t)BVG_MAX_TILE_COUNTERS * 2 * sizeof(unsigned int));
    pl->total = o;
    return BVG_OK;
}

// ---- concurrent AMP blocks -----------------------------------------------------------------------------------
// The num_kernels AMP blocks of a stage read the same input and are independent until their last conv, so they are
// launched on separate streams: the tensor-pipe-bound convs of one block overlap the FP32-pipe-bound Activation1d
// kernels of another.  The auxiliary streams / events are host-side handles cached per host thread and device (no
// device memory, no cross-thread sharing); everything is forked from and joined back into the caller's stream, so
// the caller still sees ordinary stream semantics (and stream capture works).
struct AuxStreams {
    int dev = -1;
    cudaStream_t s[BVG_MAX_KERNELS - 1] = {};
    cudaEvent_t fork = nullptr, last[BVG_MAX_KERNELS] = {};
    bool ok = false;
};

// One entry per (host thread, device), looked up by the exact device ordinal (a process may drive any number of GPUs).
static AuxStreams* aux_streams() {
    static thread_local std::vector<AuxStreams*> cache;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    for (AuxStreams* a : cache)
        if (a->dev == dev) return a->ok ? a : nullptr;
    AuxStreams* a = new AuxStreams();
    a->dev = dev;
    bool good = cudaEventCreateWithFlags(&a->fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; good && i < BVG_MAX_KERNELS - 1; ++i)
        good = cudaStreamCreateWithFlags(&a->s[i], cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; good && i < BVG_MAX_KERNELS; ++i)
        good = cudaEventCreateWithFlags(&a->last[i], cudaEventDisableTiming) == cudaSuccess;
    a->ok = good;
    cache.push_back(a);
    return good ? a : nullptr;
}

// Joins the auxiliary streams back into the caller's stream on EVERY exit path of a stage: an error return between the
// fork and the regular join must not leave kernels of the aux streams running behind the caller's back (the caller frees
// the workspace in stream order on `st`) nor a stream capture with dangling forked streams.
struct AuxJoin {
    AuxStreams* aux;
    cudaStream_t st;
    int n;          // aux streams forked (num_kernels - 1), 0 = nothing to join
    AuxJoin(AuxStreams* a, cudaStream_t s) : aux(a), st(s), n(0) {}
    void forked(int num_kernels) { n = num_kernels - 1; }
    void joined() { n = 0; }
    ~AuxJoin() {
        if (aux == nullptr) return;
        for (int j = 0; j < n; ++j) {
            // a fresh record on each aux stream covers whatever was queued there before the error
            if (cudaEventRecord(aux->last[j + 1], aux->s[j]) == cudaSuccess) cudaStreamWaitEvent(st, aux->last[j + 1], 0);
        }
    }
};

// BVG_ACT_MULTI=0: one Activation1d launch per AMP block at the top of a stage instead of the shared-up-sampler kernel
static int act_multi_requested() {
    static const int v = []() {
        const char* e = getenv("BVG_ACT_MULTI");
        return (e != nullptr && e[0] == '0' && e[1] == 0) ? 0 : 1;
    }();
    return v;
}

// BVG_COND_FUSED=0: the v1 conditioning vectors are added by separate kernels instead of in the conv epilogues (A/B)
static int cond_fused_requested() {
    static const int v = []() {
        const char* e = getenv("BVG_COND_FUSED");
        return (e != nullptr && e[0] == '0' && e[1] == 0) ? 0 : 1;
    }();
    return v;
}

static int concurrency_requested() {
    static const int v = []() {
        const char* e = getenv("BVG_STREAMS");
        return (e != nullptr && e[0] == '1' && e[1] == 0) ? 0 : 1;
    }();
    return v;
}

}  // namespace bvg

using namespace bvg;

extern "C" int bvg_conv1d_nlc_fwd(const void* x, const void* w_packed, const float* bias, const float* residual,
                                  const float* accum, float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin,
                                  int64_t Cout, int k, int dilation, float out_scale, int precision, void* stream) {
    return conv1d(precision, x, w_packed, bias, residual, accum, y_f32, y_bf16, B, T, Cin, Cout, k, dilation, out_scale,
                  static_cast<cudaStream_t>(stream));
}

extern "C" int bvg_convtranspose1d_nlc_fwd(const void* x, const void* w_packed, const float* bias, float* y_f32,
                                           void* y_bf16, int64_t B, int64_t Tin, int64_t Cin, int64_t Cout, int K, int u,
                                           int precision, void* stream) {
    return convtranspose1d(precision, x, w_packed, bias, y_f32, y_bf16, B, Tin, Cin, Cout, K, u,
                           static_cast<cudaStream_t>(stream));
}

namespace bvg {
int conv_umma_init();
int activation1d_init();
}  // namespace bvg

// Pre-creates everything the library would otherwise create lazily inside the first call on this (host thread, device):
// the auxiliary streams / events of the forward plan and the kernels' shared-memory attributes.  Optional -- every entry
// point initialises on demand -- but a caller whose FIRST call would happen under a global-mode stream capture (where
// stream / event creation is illegal) calls this once beforehand.
extern "C" int bvg_init(int device) {
    int prev = 0;
    BVG_CUDA(cudaGetDevice(&prev));
    if (device < 0) device = prev;
    int rc = bvg_device_check(device);
    if (rc != BVG_OK) return rc;
    if (device != prev) BVG_CUDA(cudaSetDevice(device));
    rc = conv_umma_init();
    if (rc == BVG_OK) rc = activation1d_init();
    if (rc == BVG_OK && aux_streams() == nullptr)
        rc = set_error(BVG_ERR_CUDA, "bvg_init: could not create the auxiliary streams / events on device %d", device);
    if (device != prev) cudaSetDevice(prev);
    return rc;
}

extern "C" size_t bvg_workspace_bytes(const bvg_model_desc* m, int64_t B, int64_t T_mel) {
    Plan pl;
    if (B <= 0 || T_mel <= 0) return 256;
    if (make_plan(m, B, T_mel, &pl) != BVG_OK) return 0;
    return pl.total + 256;
}

// wav_dtype: BVG_F32 (the module's output) or BVG_I16 (the caller's PCM conversion fused into the tail kernel)
// mel_layout 0: ncl with element strides; 1: dense nlc.  cond: nullptr or num_stages + 1 device pointers (v1 generator).
static int forward_impl(const bvg_model_desc* m, const float* mel, int mel_layout, int64_t mel_stride_b,
                        int64_t mel_stride_c, const float* const* cond, void* wav, int wav_dtype, int64_t B, int64_t T_mel,
                        void* workspace, size_t workspace_bytes, void* stream) {
    BVG_REQUIRE(m && mel && wav, BVG_ERR_INVALID, "bvg_forward: null pointer");
    BVG_REQUIRE(B >= 0 && T_mel >= 0, BVG_ERR_INVALID, "bvg_forward: negative shape");
    if (B == 0 || T_mel == 0) return BVG_OK;
    Plan pl;
    int rc = make_plan(m, B, T_mel, &pl);
    if (rc != BVG_OK) return rc;
    BVG_REQUIRE(workspace != nullptr && workspace_bytes >= pl.total, BVG_ERR_WORKSPACE,
                "bvg_forward: workspace of %zu bytes is smaller than the %zu required", workspace_bytes, pl.total);
    BVG_REQUIRE(reinterpret_cast<uintptr_t>(workspace) % 256 == 0, BVG_ERR_INVALID, "bvg_forward: workspace must be 256-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int prec = m->precision;
    const bool tc = prec == BVG_BF16_TC;
    const bool split = prec == BVG_F32_SPLIT_TC;  // fp32 parity mode on the tensor cores (bf16x3 split operands)
    const int op_dt = tc ? BVG_BF16 : (split ? BVG_BF16X3 : BVG_F32);
    const int y1_dt = tc ? BVG_BF16 : BVG_F32;
    const int precise = tc ? 0 : 1;
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    // tile counters of this call's conv launches, zeroed in stream order before the first kernel (a memset node under capture)
    TileCounterScope counter_scope;
    static const bool dyn_env = []() { const char* e = getenv("BVG_CONV_DYNAMIC"); return e != nullptr && atoi(e) != 0; }();
    if (dyn_env && prec != BVG_EXACT_F32) {
        BVG_CUDA(cudaMemsetAsync(ws + pl.off_ctr, 0, (size_t)BVG_MAX_TILE_COUNTERS * 2 * sizeof(unsigned int), st));
        t_tile_ctr.base = reinterpret_cast<unsigned int*>(ws + pl.off_ctr);
        t_tile_ctr.next = 0;
    }
    void* mel_nlc = ws + pl.off_mel;
    void* pre = ws + pl.off_pre;
    float* X = reinterpret_cast<float*>(ws + pl.off_x);
    // profiling keeps the plan on one stream so that every event pair brackets exactly one kernel
    AuxStreams* aux = (g_prof == nullptr && concurrency_requested() && m->num_kernels > 1) ? aux_streams() : nullptr;
    float* SUM = reinterpret_cast<float*>(ws + pl.off_sum);
    void* SUMOP = ws + pl.off_sumop;

    const double es = tc ? 2.0 : 4.0;  // SURVEY.md 8(d) element size of the mode
    auto conv_flops = [](double cin, double cout, double k, double rows) { return 2.0 * cin * cout * k * rows; };
    auto conv_bytes = [&](double cin, double cout, double k, double rows, bool res) {
        return es * rows * (cin + cout) + es * cin * cout * k + 4.0 * cout + (res ? es * rows * cout : 0.0);
    };
    BVG_REQUIRE(mel_layout == 0 || mel_layout == 1, BVG_ERR_INVALID, "bvg_forward: unknown input layout %d", mel_layout);
    const bool amp2 = m->resblock == 2;
    const bool fused_cond = cond_fused_requested() != 0;
    // mel [B, num_mels, T] -> [B, T, num_mels] in the operand dtype (an nlc input only changes dtype)
    {
    ProfScope ps(BVG_K_LAYOUT, -1, m->num_mels, m->num_mels, 0, 0, B * T_mel, 0.0, 0.0, st);
    if (split) {
        // fp32 channels-last copy in SUM (free until the first stage ends), then the operand split
        const float* mel32 = mel;
        if (mel_layout == 0) {
            rc = transpose_any(mel, BVG_F32, SUM, BVG_F32, B, m->num_mels, T_mel, mel_stride_b, mel_stride_c, st);
            mel32 = SUM;
        }
        if (rc == BVG_OK) rc = split_rows(mel32, mel_nlc, B * T_mel, m->num_mels, st);
    } else if (mel_layout == 0)
        rc = transpose_any(mel, BVG_F32, mel_nlc, op_dt, B, m->num_mels, T_mel, mel_stride_b, mel_stride_c, st);
    else if (tc)
        rc = add_channel_vector(mel, BVG_F32, nullptr, mel_nlc, op_dt, B, T_mel, m->num_mels, st);
    else
        mel_nlc = const_cast<float*>(mel);  // fp32 mode reads the caller's tensor in place
    }
    if (rc != BVG_OK) return rc;
    // conv_pre (bigvgan.py:362): output feeds the first transposed conv directly -> operand dtype
    const bool cond_pre = cond != nullptr && cond[0] != nullptr;
    {
        const bvg_conv_desc& c = m->conv_pre;
        ProfScope ps(BVG_K_CONV1D, -1, c.cin, c.cout, c.k, 1, B * T_mel, conv_flops(c.cin, c.cout, c.k, (double)B * T_mel),
                     conv_bytes(c.cin, c.cout, c.k, (double)B * T_mel, false), st);
        BVG_REQUIRE(c.cin == m->num_mels && c.cout == m->upsample_initial_channel, BVG_ERR_INVALID, "conv_pre: channel mismatch");
        // conditioned (v1, indextts/BigVGAN/models.py:236): the per-utterance vector rides in the conv epilogue as a second
        // bias (one rounding to the operand dtype).  BVG_COND_FUSED=0: fp32 result into X and a separate add kernel.
        // Split mode: the fp32 result goes to X and is split into the operand tensor afterwards.
        float* pre_f32 = (split || (cond_pre && !fused_cond)) ? X : (tc ? nullptr : reinterpret_cast<float*>(pre));
        rc = conv1d(prec, mel_nlc, c.w, c.bias, nullptr, nullptr, pre_f32, pre_f32 ? nullptr : pre, B, T_mel, c.cin, c.cout,
                    c.k, 1, 1.0f, st, (cond_pre && fused_cond) ? cond[0] : nullptr);
        if (rc != BVG_OK) return rc;
    }
    if (cond_pre && !fused_cond) {
        // x = x + cond_layer(speaker_embedding)   (indextts/BigVGAN/models.py:236)
        const double n = (double)B * T_mel * m->upsample_initial_channel;
        ProfScope ps(BVG_K_COND, -1, m->upsample_initial_channel, m->upsample_initial_channel, 0, 0, B * T_mel, n, (4.0 + es) * n, st);
        rc = split ? add_channel_vector(X, BVG_F32, cond[0], X, BVG_F32, B, T_mel, m->upsample_initial_channel, st)
                   : add_channel_vector(X, BVG_F32, cond[0], pre, op_dt, B, T_mel, m->upsample_initial_channel, st);
        if (rc != BVG_OK) return rc;
    }
    if (split) {
        ProfScope ps(BVG_K_LAYOUT, -1, m->upsample_initial_channel, m->upsample_initial_channel, 0, 0, B * T_mel, 0.0, 0.0, st);
        rc = split_rows(X, pre, B * T_mel, m->upsample_initial_channel, st);
        if (rc != BVG_OK) return rc;
    }
    const void* stage_in = pre;
    int64_t T = T_mel;
    int C = m->upsample_initial_channel;
    const float inv_nk = 1.0f / static_cast<float>(m->num_kernels);
    for (int i = 0; i < m->num_stages; ++i) {
        const bvg_conv_desc& up = m->ups[i];
        const int u = m->upsample_rates[i];
        BVG_REQUIRE(up.cin == C && up.cout == C / 2, BVG_ERR_INVALID, "ups[%d]: channel mismatch", i);
        // upsample (bigvgan.py:366-367): fp32 residual stream X
        {
            const double rin = (double)B * T;
            ProfScope ps(BVG_K_CONVT, i, C, C / 2, up.k, 1, B * T * u, 2.0 * C * (C / 2) * up.k * rin,
                         es * rin * (C + (C / 2) * (double)u) + es * (double)C * (C / 2) * up.k, st);
            // v1 (models.py:244-245): x = x + conds[i](speaker_embedding), fused into the epilogue of every phase
            rc = convtranspose1d(prec, stage_in, up.w, up.bias, X, nullptr, B, T, C, C / 2, up.k, u, st,
                                 (cond != nullptr && fused_cond) ? cond[i + 1] : nullptr);
        }
        if (rc != BVG_OK) return rc;
        T *= u;
        C /= 2;
        if (cond != nullptr && cond[i + 1] != nullptr && !fused_cond) {
            // x = x + conds[i](speaker_embedding)   (indextts/BigVGAN/models.py:244-245), in place on the residual stream
            const double n = (double)B * T * C;
            ProfScope ps(BVG_K_COND, i, C, C, 0, 0, B * T, n, 8.0 * n, st);
            rc = add_channel_vector(X, BVG_F32, cond[i + 1], X, BVG_F32, B, T, C, st);
            if (rc != BVG_OK) return rc;
        }
        const bool last_stage = (i + 1 == m->num_stages);
        // The AMP blocks all begin with an Activation1d of X (bigvgan.py:133-135): one launch computes the 2x up-sampling
        // of X once and applies the blocks' snake parameters to it (activation1d_nlc_multi), before the streams fork.
        const bool heads = act_multi_requested() && m->num_kernels == 3 && m->num_dilations >= 1;
        if (heads) {
            void* ys[BVG_MAX_KERNELS];
            const float* als[BVG_MAX_KERNELS];
            const float* bes[BVG_MAX_KERNELS];
            for (int j = 0; j < m->num_kernels; ++j) {
                ys[j] = ws + pl.off_a[j];
                als[j] = m->acts[i][j][0].alpha_log;
                bes[j] = m->acts[i][j][0].beta_log;
            }
            const double rows = (double)B * T;
            // algorithmic work = the sum of the three reference layers this launch implements
            ProfScope ps(BVG_K_ACT, i, C, C, 0, 0, B * T, m->num_kernels * 62.0 * rows * C,
                         m->num_kernels * (2.0 * es * rows * C + 8.0 * C), st);
            rc = activation1d_nlc_multi(X, BVG_F32, ys, op_dt, als, bes, m->num_kernels, B, T, C, precise, st);
            if (rc != BVG_OK) return rc;
        }
        AuxJoin join_guard(aux, st);
        if (aux) {
            BVG_CUDA(cudaEventRecord(aux->fork, st));
            for (int j = 1; j < m->num_kernels; ++j) BVG_CUDA(cudaStreamWaitEvent(aux->s[j - 1], aux->fork, 0));
            join_guard.forked(m->num_kernels);
        }
        for (int j = 0; j < m->num_kernels; ++j) {
            cudaStream_t sj = (aux && j > 0) ? aux->s[j - 1] : st;
            float* R = reinterpret_cast<float*>(ws + pl.off_r[j]);
            void* A = ws + pl.off_a[j];
            void* Y1 = ws + pl.off_y1[j];
            const float* xr = X;  // residual stream of this block (bigvgan.py:132-141)
            for (int l = 0; l < m->num_dilations; ++l) {
                const bvg_conv_desc& c1 = m->convs1[i][j][l];
                const bvg_conv_desc& c2_ = m->convs2[i][j][l];
                const bvg_act_desc& a1 = m->acts[i][j][amp2 ? l : 2 * l];
                const bvg_act_desc& a2 = m->acts[i][j][amp2 ? l : 2 * l + 1];
                const double rows = (double)B * T;
                const double act_flops = 62.0 * rows * C, act_bytes = 2.0 * es * rows * C + 8.0 * C;
                if (!(heads && l == 0)) {
                    ProfScope ps(BVG_K_ACT, i, C, C, 0, 0, B * T, act_flops, act_bytes, sj);
                    rc = activation1d_nlc(xr, BVG_F32, A, op_dt, a1.alpha_log, a1.beta_log, B, T, C, precise, sj);
                }
                if (rc != BVG_OK) return rc;
                // xt between c1 and a2 is an intermediate (not the residual stream): operand dtype (bf16 in TC mode;
                // costs ~1.2 dB of the 43.8 dB budget, tools/emulate_bf16_plan.py)
                if (!amp2) {
                    ProfScope ps(BVG_K_CONV1D, i, C, C, c1.k, c1.dilation, B * T, conv_flops(C, C, c1.k, rows),
                                 conv_bytes(C, C, c1.k, rows, false), sj);
                    rc = conv1d_desc(prec, c1, A, nullptr, nullptr, tc ? nullptr : reinterpret_cast<float*>(Y1),
                                     tc ? Y1 : nullptr, B, T, C, 1.0f, sj);
                }
                if (rc != BVG_OK) return rc;
                if (!amp2) {
                    ProfScope ps(BVG_K_ACT, i, C, C, 0, 0, B * T, act_flops, act_bytes, sj);
                    rc = activation1d_nlc(Y1, y1_dt, A, op_dt, a2.alpha_log, a2.beta_log, B, T, C, precise, sj);
                }
                if (rc != BVG_OK) return rc;
                // AMPBlock2: the block's only conv of this dilation carries the residual add (the conv the code below
                // calls c2 is convs1[l], dilated)
                const bvg_conv_desc& c2 = amp2 ? c1 : c2_;
                ProfScope ps(BVG_K_CONV1D, i, C, C, c2.k, c2.dilation, B * T, conv_flops(C, C, c2.k, rows),
                             conv_bytes(C, C, c2.k, rows, true), sj);
                const bool last = (l + 1 == m->num_dilations);
                if (!last) {
                    // x = xt + x
                    rc = conv1d_desc(prec, c2, A, xr, nullptr, R, nullptr, B, T, C, 1.0f, sj);
                    xr = R;
                } else {
                    // xs (+)= block output; x = xs / num_kernels (bigvgan.py:369-375) folded into the epilogue: the
                    // blocks' last convs accumulate into SUM one after the other (event chain).  The last block of a
                    // stage also emits the operand-dtype copy the next transposed conv reads.
                    if (aux && j > 0) BVG_CUDA(cudaStreamWaitEvent(sj, aux->last[j - 1], 0));
                    const bool emit_op = tc && !last_stage && (j + 1 == m->num_kernels);
                    rc = conv1d_desc(prec, c2, A, xr, j > 0 ? SUM : nullptr, SUM, emit_op ? SUMOP : nullptr, B, T, C, inv_nk,
                                     sj);
                    if (aux && rc == BVG_OK) BVG_CUDA(cudaEventRecord(aux->last[j], sj));
                }
                if (rc != BVG_OK) return rc;
            }
        }
        if (aux) {  // join: the last block's final conv waited for every other block's (event chain)
            BVG_CUDA(cudaStreamWaitEvent(st, aux->last[m->num_kernels - 1], 0));
            join_guard.joined();
        }
        if (split && !last_stage) {  // operand split of the stage output for the next transposed conv
            ProfScope ps(BVG_K_LAYOUT, i, C, C, 0, 0, B * T, 0.0, 0.0, st);
            rc = split_rows(SUM, SUMOP, B * T, C, st);
            if (rc != BVG_OK) return rc;
        }
        stage_in = (tc || split) ? SUMOP : static_cast<const void*>(SUM);
    }
    // activation_post -> conv_post -> clamp / tanh (bigvgan.py:378-384), fp32 in both modes
    {
        const double rows = (double)B * T;
        ProfScope ps(BVG_K_ACT, -1, C, C, 0, 0, B * T, 62.0 * rows * C, 2.0 * es * rows * C + 8.0 * C, st);
        float* R = reinterpret_cast<float*>(ws + pl.off_r[0]);
        // same sine mode as the rest of the plan.  MUFU cos without range reduction was probed for this layer's worst case
        // (tools/act_range_probe.py, profiles/r01_v8_act_range_probe.log: log-parameters up to +-4, arguments up to
        // 4000 rad): within 4 dB of the range-reduced sinf path, both limited by the fp32 rounding of a*u that the
        // reference shares; the precise path costs 2.7x on this launch (197 vs 526 G elements/s at cfg3).
        rc = activation1d_nlc(SUM, BVG_F32, R, BVG_F32, m->act_post.alpha_log, m->act_post.beta_log, B, T, C, precise, st);
    }
    if (rc != BVG_OK) return rc;
    const bvg_conv_desc& cp = m->conv_post;
    BVG_REQUIRE(cp.cout == 1 && cp.cin == C, BVG_ERR_UNSUPPORTED, "conv_post must map %d channels to 1", C);
    ProfScope ps(BVG_K_POST, -1, C, 1, cp.k, 1, B * T, 2.0 * C * cp.k * (double)B * T, es * (double)B * T * (C + 1.0), st);
    const float* post_in = reinterpret_cast<const float*>(ws + pl.off_r[0]);
    if (wav_dtype == BVG_I16)
        return bvg_conv_post_pcm16_fwd(post_in, static_cast<const float*>(cp.w), cp.bias, static_cast<int16_t*>(wav), B, T, C,
                                       cp.k, m->use_tanh_at_final, stream);
    return bvg_conv_post_fwd(post_in, static_cast<const float*>(cp.w), cp.bias, static_cast<float*>(wav), B, T, C, cp.k,
                             m->use_tanh_at_final, stream);
}

extern "C" int bvg_forward(const bvg_model_desc* m, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                           float* wav, int64_t B, int64_t T_mel, void* workspace, size_t workspace_bytes, void* stream) {
    return forward_impl(m, mel, 0, mel_stride_b, mel_stride_c, nullptr, wav, BVG_F32, B, T_mel, workspace, workspace_bytes,
                        stream);
}

extern "C" int bvg_forward_cond(const bvg_model_desc* m, const float* x, int x_layout, int64_t x_stride_b,
                                int64_t x_stride_c, const float* const* cond_host, float* wav, int64_t B, int64_t T,
                                void* workspace, size_t workspace_bytes, void* stream) {
    return forward_impl(m, x, x_layout, x_stride_b, x_stride_c, cond_host, wav, BVG_F32, B, T, workspace, workspace_bytes,
                        stream);
}

extern "C" int bvg_forward_pcm16(const bvg_model_desc* m, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                                 int16_t* pcm, int64_t B, int64_t T_mel, void* workspace, size_t workspace_bytes,
                                 void* stream) {
    return forward_impl(m, mel, 0, mel_stride_b, mel_stride_c, nullptr, pcm, BVG_I16, B, T_mel, workspace, workspace_bytes,
                        stream);
}

// ---- bvg_forward without a caller workspace ---------------------------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the plan's
// workspace lives in a file-scope device array of BVG_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library.
// The array's name is only a host-side stand-in here; its device address comes from cudaGetSymbolAddress, looked up once
// per device with that device current.
static __device__ __align__(256) unsigned char g_noalloc_scratch[BVG_NOALLOC_SCRATCH_BYTES];

namespace bvg {
static int noalloc_scratch_ptr(void** out) {
    constexpr int kMaxDevices = 64;
    static std::atomic<void*> cache[kMaxDevices];
    int dev = 0;
    BVG_CUDA(cudaGetDevice(&dev));
    BVG_REQUIRE(dev >= 0 && dev < kMaxDevices, BVG_ERR_UNSUPPORTED, "bvg_forward_noalloc: device ordinal %d >= %d", dev,
                kMaxDevices);
    void* p = cache[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            return set_error(BVG_ERR_CUDA, "bvg_forward_noalloc: cudaGetSymbolAddress on device %d: %s", dev,
                             cudaGetErrorString(e));
        cache[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return BVG_OK;
}
}  // namespace bvg

extern "C" int bvg_forward_noalloc(const bvg_model_desc* m, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                                   float* wav, int64_t B, int64_t T_mel, void* stream) {
    BVG_REQUIRE(m && mel && wav, BVG_ERR_INVALID, "bvg_forward_noalloc: null pointer");
    BVG_REQUIRE(B >= 0 && T_mel >= 0, BVG_ERR_INVALID, "bvg_forward_noalloc: negative shape");
    if (B == 0 || T_mel == 0) return BVG_OK;
    Plan pl;
    int rc = make_plan(m, B, T_mel, &pl);
    if (rc != BVG_OK) return rc;
    BVG_REQUIRE(pl.total <= (size_t)BVG_NOALLOC_SCRATCH_BYTES, BVG_ERR_WORKSPACE,
                "bvg_forward_noalloc: B=%lld, T_mel=%lld needs %zu bytes of scratch, the built-in array holds %zu "
                "(use bvg_forward with a caller workspace for larger inputs)",
                (long long)B, (long long)T_mel, pl.total, (size_t)BVG_NOALLOC_SCRATCH_BYTES);
    void* scratch = nullptr;
    rc = noalloc_scratch_ptr(&scratch);
    if (rc != BVG_OK) return rc;
    // Nothing has to start at zero: every buffer of the plan is written in full by its producer before any consumer reads
    // it, and rows outside [0, T) come from TMA zero fill / clamped indices, never from memory -- what the previous call
    // left in the array is not read.
    return forward_impl(m, mel, 0, mel_stride_b, mel_stride_c, nullptr, wav, BVG_F32, B, T_mel, scratch,
                        (size_t)BVG_NOALLOC_SCRATCH_BYTES, stream);
}

extern "C" int bvg_forward_profile(const bvg_model_desc* m, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                                   float* wav, int64_t B, int64_t T_mel, void* workspace, size_t workspace_bytes,
                                   void* stream, bvg_launch_record* records_host, int max_records, int* n_records) {
    return bvg_forward_cond_profile(m, mel, 0, mel_stride_b, mel_stride_c, nullptr, wav, B, T_mel, workspace, workspace_bytes,
                                    stream, records_host, max_records, n_records);
}

extern "C" int bvg_forward_cond_profile(const bvg_model_desc* m, const float* mel, int x_layout, int64_t mel_stride_b,
                                        int64_t mel_stride_c, const float* const* cond_host, float* wav, int64_t B,
                                        int64_t T_mel, void* workspace, size_t workspace_bytes, void* stream,
                                        bvg_launch_record* records_host, int max_records, int* n_records) {
    BVG_REQUIRE(records_host && n_records && max_records >= 0, BVG_ERR_INVALID, "bvg_forward_profile: null record buffer");
    std::vector<ProfEntry> entries;
    entries.reserve(512);
    g_prof = &entries;
    const int rc = forward_impl(m, mel, x_layout, mel_stride_b, mel_stride_c, cond_host, wav, BVG_F32, B, T_mel, workspace,
                                workspace_bytes, stream);
    g_prof = nullptr;
    cudaError_t e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
    int n = 0;
    for (auto& en : entries) {
        float ms = 0.f;
        if (e == cudaSuccess) cudaEventElapsedTime(&ms, en.e0, en.e1);
        en.rec.ms = ms;
        if (n < max_records) records_host[n] = en.rec;
        ++n;
        cudaEventDestroy(en.e0);
        cudaEventDestroy(en.e1);
    }
    *n_records = n;
    if (rc != BVG_OK) return rc;
    if (e != cudaSuccess) return set_error(BVG_ERR_CUDA, "bvg_forward_profile: %s", cudaGetErrorString(e));
    return BVG_OK;
}
