/* bvg.h -- C ABI of libbvg.so: the B200 (sm_100a) BigVGAN-v2 vocoder forward path.
 *
 * This is the native-op boundary that replaces the reference's torch extension
 *   /root/reference/indextts/s2mel/modules/bigvgan/alias_free_activation/cuda/   (BV/AF/cuda below)
 *       anti_alias_activation.cpp:19-23      pybind `forward(input, up_ftr, down_ftr, alpha, beta)`
 *       anti_alias_activation_cuda.cu:212-246 `fwd_cuda` (callee-allocated torch::Tensor, current stream)
 * and the ATen/cuDNN calls the reference makes for the rest of the generator
 *   /root/reference/indextts/s2mel/modules/bigvgan/bigvgan.py                    (BV/bigvgan.py below)
 * with a real C interface: plain pointers and sizes, no torch types, no C++ exceptions.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name ends in _host; the caller owns all memory (inputs,
 *    outputs, workspace) and keeps it alive until `stream` has been synchronised;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); the library never
 *    synchronises, never allocates device memory and keeps no global mutable state: concurrent calls from
 *    several host threads on different streams are safe (the reference's server does that, serve_tars.py:66);
 *  - every function returns BVG_OK (0) or a negative bvg_status; the message is available from
 *    bvg_last_error() (thread-local).  Unsupported device / dtype / shape is an ERROR, never a fallback;
 *  - 64-bit indexing throughout (the reference kernel overflows int32 at B*C*T >= 2^31, .cu:68-74);
 *  - layouts: "ncl" = [B, C, T] time innermost (the reference's public layout),
 *             "nlc" = [B, T, C] channel innermost (the library's internal layout for the implicit-GEMM convs).
 */
#ifndef BVG_H_
#define BVG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BVG_VERSION 100

#if defined(__GNUC__)
#define BVG_API __attribute__((visibility("default")))
#else
#define BVG_API
#endif

typedef enum {
    BVG_OK = 0,
    BVG_ERR_INVALID = -1,     /* bad argument (null pointer, shape, alignment) */
    BVG_ERR_UNSUPPORTED = -2, /* dtype / shape / hyper-parameter outside what the kernels implement */
    BVG_ERR_DEVICE = -3,      /* not an sm_100 device */
    BVG_ERR_CUDA = -4,        /* a CUDA runtime / driver call failed */
    BVG_ERR_WORKSPACE = -5    /* workspace too small */
} bvg_status;

typedef enum {
    BVG_F32 = 0,
    BVG_BF16 = 1,
    BVG_I16 = 2,   /* PCM output of the tail kernel only */
    BVG_BF16X3 = 3 /* an fp32 tensor as three bf16 pieces v = h + m + l, laid out for the split-operand tensor-core mode:
                    * activations [.., 6*C] = [l | m | m | h | h | h], packed weights [.., 6*Cin] = [h | m | h | l | m | h]
                    * (BVG_F32_SPLIT_TC below).  Output of bvg_split_bf16x3_nlc / bvg_activation1d_nlc_fwd / bvg_pack_*. */
} bvg_dtype;

/* bvg_precision: arithmetic mode of the convolutions.
 *   BVG_EXACT_F32    : fp32 operands, fp32 FMA accumulation on the CUDA cores, accurate sin.  The cross-check
 *                      implementation (a plain tiled kernel); not a throughput path.
 *   BVG_BF16_TC      : bf16 operands on tcgen05 tensor cores with fp32 TMEM accumulation; fp32 residual stream and
 *                      fp32 Activation1d arithmetic                                  -> the throughput mode
 *   BVG_F32_SPLIT_TC : the fp32 PARITY mode on the tensor cores (what the reference's fp32 vocoder call,
 *                      infer_v2.py:830-831, 860, is compared with): every fp32 operand is split into three bf16 pieces
 *                      (BVG_BF16X3) and the six significant partial products accumulate in one fp32 TMEM accumulator --
 *                      the same tcgen05 implicit-GEMM kernel with a six times longer K axis (small terms first); what
 *                      is dropped is below 2^-24 of a product, what remains is the tensor core's truncating fp32
 *                      accumulation (~1e-6..1e-5 of a layer's output for the deepest K).  fp32 residual stream, accurate sin.  In the standalone conv entry points the
 *                      operands are BVG_BF16X3 tensors (x: [B, T, 6*Cin], w_packed from bvg_pack_* with dtype BVG_BF16X3). */
typedef enum { BVG_EXACT_F32 = 0, BVG_BF16_TC = 1, BVG_F32_SPLIT_TC = 2 } bvg_precision;

#define BVG_MAX_STAGES 8
#define BVG_MAX_KERNELS 4
#define BVG_MAX_DILATIONS 4
#define BVG_MAX_PHASES 8

/* ---------------------------------------------------------------------------------------------------------- */
/* diagnostics                                                                                                 */
BVG_API const char* bvg_last_error(void);
BVG_API int bvg_version(void);
/* BVG_OK iff `device` (or the current device when < 0) has compute capability 10.x.  Replaces the reference's
 * silent torch fallback at infer_v2.py:176-182 with a hard error. */
BVG_API int bvg_device_check(int device);
/* number of kernel launches issued by this library on the calling thread since the last reset (bench.py's
 * gpu_launches claim) */
BVG_API int64_t bvg_launch_count(void);
BVG_API void bvg_reset_launch_count(void);
/* Optional eager initialisation for `device` (< 0: the current device) on the calling host thread: creates the forward
 * plan's auxiliary streams / events and sets the kernels' shared-memory attributes, all of which are otherwise created on
 * first use.  A caller whose first bvg_forward would run under a global-mode stream capture (where stream and event
 * creation is illegal) calls this once before capturing.  No reference counterpart (torch initialises lazily too). */
BVG_API int bvg_init(int device);

/* ---------------------------------------------------------------------------------------------------------- */
/* (a) fused anti-aliased activation: y = down2(snake_beta(up2(x)))
 * Replaces BV/AF/cuda/anti_alias_activation_cuda.cu:43-246 and the torch path BV/AF/torch/act.py:25-30
 * (resample.py:30-38, activations.py:107-120, filter.py:94-101).  The 12-tap Kaiser-sinc filter
 * (filter.py:30-62, cutoff 0.25, half-width 0.3) is a compile-time constant.
 * alpha_log / beta_log: per-channel fp32 parameters in LOG scale (the kernel applies exp, like the reference
 * extension, activation1d.py:68-72).
 * precise != 0: range-reduced polynomial sin (fp32 parity mode); 0: MUFU approximations.                     */

/* ncl layout.  x element strides: x[b][c][t] = x[b*x_stride_b + c*x_stride_c + t] (time stride 1; honours the
 * non-contiguous time slice infer_v2.py:856 produces).  y is dense [B, C, T].  dtype applies to x and y. */
BVG_API int bvg_activation1d_fwd(const void* x, void* y, const float* alpha_log, const float* beta_log, int64_t B, int64_t C,
                         int64_t T, int64_t x_stride_b, int64_t x_stride_c, int dtype, int precise, void* stream);

/* nlc layout, dense.  C % 8 == 0.  in/out dtypes independent (fp32 residual stream in, bf16 MMA operand out); with an fp32
 * input y_dtype may also be BVG_BF16X3: y is then the split operand tensor [B, T, 6*C] of BVG_F32_SPLIT_TC. */
BVG_API int bvg_activation1d_nlc_fwd(const void* x, int x_dtype, void* y, int y_dtype, const float* alpha_log,
                             const float* beta_log, int64_t B, int64_t T, int64_t C, int precise, void* stream);

/* n_heads Activation1d of the SAME input with different snake parameters: ys[h] = Activation1d(x; alphas_log[h],
 * betas_log[h]).  The three AMP blocks of a stage start with exactly this (BV/bigvgan.py:369-374 applies
 * resblocks[3i+j] to one x; AMPBlock1.forward :132-141 begins with a1(x)).  Groups of three heads with fp32 input run as
 * one kernel that computes the (parameter independent) 2x up-sampler once; ys, alphas_log, betas_log are HOST arrays of
 * n_heads device pointers.  Results are bit-identical to n_heads bvg_activation1d_nlc_fwd calls. */
BVG_API int bvg_activation1d_nlc_multi_fwd(const void* x, int x_dtype, void* const* ys, int y_dtype,
                                   const float* const* alphas_log, const float* const* betas_log, int n_heads, int64_t B,
                                   int64_t T, int64_t C, int precise, void* stream);

/* ---------------------------------------------------------------------------------------------------------- */
/* weight preparation (BV/bigvgan.py:388-400 remove_weight_norm is where the host calls these)                 */

/* w[r, :] = g[r] * v[r, :] / ||v[r, :]||_2  (torch._weight_norm(v, g, dim=0)); rows = dim 0, row_len = rest. */
BVG_API int bvg_fold_weight_norm(const float* v, const float* g, float* w, int64_t rows, int64_t row_len, void* stream);

/* Conv1d weight [Cout, Cin, k] fp32 -> tap-major K-major [k][Cout][Cin] in `dtype` (bf16 for BVG_BF16_TC, fp32 for
 * BVG_EXACT_F32; BVG_BF16X3 for BVG_F32_SPLIT_TC: [k][Cout][6*Cin] bf16, weight term order). */
BVG_API int bvg_pack_conv1d_weight(const float* w, void* out, int64_t Cout, int64_t Cin, int64_t k, int dtype, void* stream);

/* ConvTranspose1d weight [Cin, Cout, K] fp32 -> polyphase [K/u][u*Cout][Cin]: row n = r*Cout + co of tap j holds
 * w[:, co, (r+p)%u + u*(K/u-1-j)], p = (K-u)/2; tap j of phase r reads input row q + floor((r+p)/u)-(K/u-1)+j.
 * No zero-stuffed tensor is ever formed (north_star (c)).  K % u == 0. */
BVG_API int bvg_pack_convtranspose1d_weight(const float* w, void* out, int64_t Cin, int64_t Cout, int64_t K, int64_t u,
                                    int dtype, void* stream);

/* fp32 [rows, C] -> BVG_BF16X3 [rows, 6*C] (activation term order), C % 4 == 0: the operand split of BVG_F32_SPLIT_TC for
 * tensors no Activation1d produced (the mel, conv_pre's output, the stage sums feeding the transposed convs). */
BVG_API int bvg_split_bf16x3_nlc(const float* x, void* y_bf16x3, int64_t rows, int64_t C, void* stream);

/* layout changes at the path's edges */
BVG_API int bvg_ncl_to_nlc(const void* x, int x_dtype, int64_t x_stride_b, int64_t x_stride_c, void* y, int y_dtype, int64_t B,
                   int64_t C, int64_t T, void* stream);
BVG_API int bvg_nlc_to_ncl(const void* x, int x_dtype, void* y, int y_dtype, int64_t B, int64_t C, int64_t T, void* stream);

/* ---------------------------------------------------------------------------------------------------------- */
/* (b) dilated Conv1d as implicit GEMM, nlc layout.  Replaces nn.Conv1d / cuDNN at BV/bigvgan.py:58-66,75-83,
 * 285-287.
 *   y[b,t,co] = out_scale * ( sum_{j<k} sum_ci w[j][co][ci] * x[b, t + (j-(k-1)/2)*dilation, ci] + bias[co]
 *                             + residual[b,t,co] ) + accum[b,t,co]
 * x: [B,T,Cin] (bf16 for BVG_BF16_TC, fp32 for BVG_EXACT_F32), zero outside [0,T);
 * w_packed from bvg_pack_conv1d_weight in the matching dtype; bias fp32 or NULL; residual / accum fp32 [B,T,Cout]
 * or NULL (accum may alias y_f32); outputs: y_f32 and/or y_bf16 (either may be NULL, not both).
 * Cin % 8 == 0, Cout % 8 == 0, k odd <= 11.                                                                   */
BVG_API int bvg_conv1d_nlc_fwd(const void* x, const void* w_packed, const float* bias, const float* residual,
                       const float* accum, float* y_f32, void* y_bf16, int64_t B, int64_t T, int64_t Cin, int64_t Cout,
                       int k, int dilation, float out_scale, int precision, void* stream);

/* (c) ConvTranspose1d(Cin, Cout, K, stride u, padding (K-u)/2) in polyphase form, nlc layout.  Replaces
 * nn.ConvTranspose1d at BV/bigvgan.py:305-313.  x: [B,Tin,Cin]; y: [B, u*Tin, Cout].                         */
BVG_API int bvg_convtranspose1d_nlc_fwd(const void* x, const void* w_packed, const float* bias, float* y_f32, void* y_bf16,
                                int64_t B, int64_t Tin, int64_t Cin, int64_t Cout, int K, int u, int precision,
                                void* stream);

/* activation_post output -> conv_post (Cin -> 1, k taps, optional bias) -> clamp(-1,1) or tanh, fp32.
 * Replaces BV/bigvgan.py:379-384.  x: fp32 nlc [B,T,Cin]; w: fp32 [Cin][k] (the module's [1,Cin,k] weight);
 * y: fp32 [B,1,T]. */
BVG_API int bvg_conv_post_fwd(const float* x, const float* w, const float* bias, float* y, int64_t B, int64_t T, int64_t Cin,
                      int k, int use_tanh, void* stream);
/* Same layer with the caller's PCM conversion fused into the epilogue: replaces BV/bigvgan.py:379-384 followed by
 * infer_v2.py:866 (clamp(32767 * wav, -32767, 32767)) and infer_v2.py:912 (.type(torch.int16), truncation toward zero).
 * pcm: int16 [B,1,T]. */
BVG_API int bvg_conv_post_pcm16_fwd(const float* x, const float* w, const float* bias, int16_t* pcm, int64_t B, int64_t T,
                            int64_t Cin, int k, int use_tanh, void* stream);

/* ---------------------------------------------------------------------------------------------------------- */
/* whole generator (BV/bigvgan.py:360-386 BigVGAN.forward)                                                     */
typedef struct {
    const void* w;     /* packed weight (bvg_pack_*), dtype per precision */
    const float* bias; /* [Cout] or NULL */
    int32_t cin, cout, k, dilation;
    /* Time folding (AMP-block convs of small-channel stages, BVG_BF16_TC only; 0 or 1 = off).  A [B, T, C] tensor is
     * bit-identical in memory to [B, T/fold, fold*C], and a C -> C conv with k taps and dilation d over the first view
     * equals a (fold*C) -> (fold*C) conv with 2*ceil((k-1)*d/2 / fold) + 1 taps and dilation 1 over the second one, with a
     * block-structured weight.  When fold > 1, `w` holds that folded weight packed like any Conv1d
     * ([taps'][fold*Cout][fold*Cin]) and `bias` the bias repeated fold times; cin / cout / k / dilation keep describing
     * the reference layer.  It turns 48-byte rows and N = 24 MMAs into 192-byte rows and N = 96 MMAs with 4x fewer tiles. */
    int32_t fold, reserved;
} bvg_conv_desc;

typedef struct {
    const float* alpha_log; /* [C] */
    const float* beta_log;  /* [C] */
} bvg_act_desc;

typedef struct {
    int32_t num_mels;
    int32_t upsample_initial_channel;
    int32_t num_stages;                         /* len(upsample_rates) */
    int32_t num_kernels;                        /* len(resblock_kernel_sizes) */
    int32_t num_dilations;                      /* len(resblock_dilation_sizes[j]) */
    int32_t upsample_rates[BVG_MAX_STAGES];
    int32_t upsample_kernel_sizes[BVG_MAX_STAGES];
    int32_t use_tanh_at_final;
    int32_t precision;                          /* bvg_precision */
    int32_t resblock;                           /* 1 (or 0) = AMPBlock1: act, convs1[l], act, convs2[l], +x per dilation
                                                 * (BV/bigvgan.py:132-141); 2 = AMPBlock2: act, convs1[l], +x (:231-236 of
                                                 * indextts/BigVGAN/bigvgan.py; convs2 and acts[..][l >= num_dilations]
                                                 * unused).  Occupies what used to be struct padding. */
    bvg_conv_desc conv_pre;
    bvg_conv_desc ups[BVG_MAX_STAGES];          /* k = transposed-conv kernel size K, dilation unused */
    bvg_conv_desc convs1[BVG_MAX_STAGES][BVG_MAX_KERNELS][BVG_MAX_DILATIONS];
    bvg_conv_desc convs2[BVG_MAX_STAGES][BVG_MAX_KERNELS][BVG_MAX_DILATIONS];
    bvg_act_desc acts[BVG_MAX_STAGES][BVG_MAX_KERNELS][2 * BVG_MAX_DILATIONS];
    bvg_act_desc act_post;
    bvg_conv_desc conv_post;                    /* w: fp32 [1][Cin][k] as stored by the module */
} bvg_model_desc;

/* bytes of scratch bvg_forward needs for a [B, num_mels, T_mel] input */
BVG_API size_t bvg_workspace_bytes(const bvg_model_desc* m_host, int64_t B, int64_t T_mel);

/* mel: fp32 [B, num_mels, T_mel] with element strides (mel_stride_b, mel_stride_c, 1);
 * wav: fp32 [B, 1, T_mel * prod(upsample_rates)], values in [-1, 1].  m_host is a HOST struct of device pointers. */
BVG_API int bvg_forward(const bvg_model_desc* m_host, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c, float* wav,
                int64_t B, int64_t T_mel, void* workspace, size_t workspace_bytes, void* stream);

/* bvg_forward with 16-bit PCM output (bvg_conv_post_pcm16_fwd as the tail): pcm int16 [B, 1, T_mel * prod(rates)] holds
 * exactly what infer_v2.py:860-866, 912 compute from the fp32 waveform. */
BVG_API int bvg_forward_pcm16(const bvg_model_desc* m_host, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                      int16_t* pcm, int64_t B, int64_t T_mel, void* workspace, size_t workspace_bytes, void* stream);

/* bvg_forward for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and result as bvg_forward, minus the workspace: the plan's scratch lives in a static device array of
 * BVG_NOALLOC_SCRATCH_BYTES inside the library (every GPU that loads libbvg.so holds it until the process exits).  The
 * call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); call bvg_init() beforehand if even the
 * host-side stream / event handles of the plan must exist before the first call.
 *  - Limit: bvg_workspace_bytes(m, B, T_mel) <= BVG_NOALLOC_SCRATCH_BYTES, else BVG_ERR_WORKSPACE with both numbers in
 *    bvg_last_error() (22 kHz generator: about 300 mel frames per call in BVG_BF16_TC, 130 in BVG_F32_SPLIT_TC).
 *  - All calls on a device share the one array.  Nothing in it has to start at zero (each plan buffer is written in
 *    full before it is read), so contents left by the previous call are harmless -- but two calls whose GPU work
 *    OVERLAPS (different streams, or graphs replayed concurrently) race on it: serialise them on one stream. */
#define BVG_NOALLOC_SCRATCH_BYTES (64u << 20)
BVG_API int bvg_forward_noalloc(const bvg_model_desc* m_host, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                        float* wav, int64_t B, int64_t T_mel, void* stream);

/* ---------------------------------------------------------------------------------------------------------- */
/* v1 generator (indextts/BigVGAN/models.py:130-258 BigVGAN.forward after the speaker encoder): the same stack with a
 * per-utterance conditioning vector added after conv_pre (:236) and after every transposed conv (:244-245), and a
 * channels-last input (the GPT latent [B, T, gpt_dim], :231 `x.transpose(1, 2)`).                               */

/* y[b, :] = w[C, D] . e[b, :] + bias: cond_layer / conds[i], a 1x1 Conv1d applied to the length-1 speaker embedding
 * (models.py:197-202).  e: fp32 [B, D]; w: fp32 [C, D] (the module's [C, D, 1] weight); bias [C] or NULL; y: fp32 [B, C]. */
BVG_API int bvg_cond_linear_fwd(const float* e, const float* w, const float* bias, float* y, int64_t B, int64_t D, int64_t C,
                        void* stream);

/* y[b, t, c] = x[b, t, c] + v[b, c] (nlc, dense; v fp32 [B, C] or NULL = dtype conversion only; x may alias y when the
 * dtypes match).  C % 4 == 0. */
BVG_API int bvg_add_channel_vector_nlc(const void* x, int x_dtype, const float* v, void* y, int y_dtype, int64_t B, int64_t T,
                               int64_t C, void* stream);

/* bvg_forward for the v1 generator.  x: fp32, either ncl [B, num_mels, T] with element strides (x_stride_b, x_stride_c, 1)
 * (x_layout = 0, exactly bvg_forward's input) or dense nlc [B, T, num_mels] (x_layout = 1; strides ignored).
 * cond_host: HOST array of num_stages + 1 device pointers (or NULL = no conditioning at all): cond_host[0] is the fp32
 * [B, upsample_initial_channel] vector added to conv_pre's output, cond_host[i + 1] the [B, C_i] vector added to
 * ups[i]'s output; NULL entries are skipped (models.py:244 `cond_in_each_up_layer` false).  m_host->num_mels is conv_pre's
 * input width (gpt_dim for the v1 model). */
BVG_API int bvg_forward_cond(const bvg_model_desc* m_host, const float* x, int x_layout, int64_t x_stride_b, int64_t x_stride_c,
                     const float* const* cond_host, float* wav, int64_t B, int64_t T, void* workspace,
                     size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------------- */
/* measurement: the same launch plan as bvg_forward with a CUDA event pair around every kernel launch (recorded on
 * `stream`).  Synchronises `stream` before returning -- a profiling entry point, not a product path.  Fills up to
 * max_records host records in launch order; *n_records = number of launches.  flops / bytes are the ALGORITHMIC
 * work of the reference layer the launch implements (bf16 IO element size in BVG_BF16_TC mode,
 * fp32 in BVG_EXACT_F32 mode; intermediates at the 2x rate and padding are not counted).                        */
typedef enum { BVG_K_LAYOUT = 0, BVG_K_CONV1D = 1, BVG_K_CONVT = 2, BVG_K_ACT = 3, BVG_K_POST = 4, BVG_K_COND = 5 } bvg_launch_kind;
typedef struct {
    int32_t kind;          /* bvg_launch_kind */
    int32_t stage;         /* -1 = pre / post, else upsample stage */
    int32_t cin, cout, k, dilation;
    int64_t rows;          /* B * T of the layer's output */
    double flops, bytes;   /* algorithmic */
    float ms;              /* measured with CUDA events on `stream` */
} bvg_launch_record;

BVG_API int bvg_forward_profile(const bvg_model_desc* m_host, const float* mel, int64_t mel_stride_b, int64_t mel_stride_c,
                        float* wav, int64_t B, int64_t T_mel, void* workspace, size_t workspace_bytes, void* stream,
                        bvg_launch_record* records_host, int max_records, int* n_records);

/* the same for bvg_forward_cond (v1 generator) */
BVG_API int bvg_forward_cond_profile(const bvg_model_desc* m_host, const float* x, int x_layout, int64_t x_stride_b,
                             int64_t x_stride_c, const float* const* cond_host, float* wav, int64_t B, int64_t T,
                             void* workspace, size_t workspace_bytes, void* stream, bvg_launch_record* records_host,
                             int max_records, int* n_records);

/* Developer instrumentation (no reference counterpart), live only in a library built with -DBVG_CONV_TRACE
 * (tools/build_variant.sh): when `buf` is a device buffer of 16 * n_tiles uint64 the next tcgen05 conv launches record SM
 * cycle counters of CTA 0's pipeline roles (TMA producer / UMMA issuer / epilogue) for their first n_tiles tiles
 * (tools/conv_timeline.py).  Pass NULL to switch it off.  The shipped build keeps NO trace state: it returns
 * BVG_ERR_UNSUPPORTED for a non-NULL buffer. */
BVG_API int bvg_debug_conv_trace(void* buf, int n_tiles);

#ifdef __cplusplus
}
#endif
#endif /* BVG_H_ */
