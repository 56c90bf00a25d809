/*
 * mstk.h -- C ABI of libmstk.so: the sm_100a kernels behind the spectrogram-latent
 * denoising hot path (UNet eps-prediction, VAE encode/decode, DDIM loop, losses,
 * mel front end).
 *
 * The reference (PrioteasaAndrei/music-style-transfer-ldm) has no FFI of its own: every
 * device op is a torch.nn / ATen call made from models/model.py, models/loss.py and
 * data/audio_processor.py.  Each entry point below therefore cites the reference
 * call site (file:line under the reference tree) whose ATen/cuDNN/cuBLAS work it
 * replaces.  The Python host layer (music_style_transfer_ldm_b200/) binds these with
 * ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless said
 *     otherwise; `stream` is a cudaStream_t / CUstream passed as void*.
 *   - every call returns 0 (MSTK_OK) or a negative error code; the message is
 *     available (thread-local) from mstk_last_error().  Nothing throws across the ABI.
 *   - no allocation, no stream synchronisation, no host-visible side effects:
 *     all work is enqueued on `stream` and is CUDA-graph capturable.
 *   - activations are NHWC ("channels last"), dtype MSTK_F32 or MSTK_BF16;
 *     parameters are packed once by the mstk_pack_* calls.
 *   - alignment: tensor base addresses should be 16-byte aligned (any cudaMalloc / framework allocation is; a view that
 *     starts part-way into one may not be).  The tensor-core, convolution, weight-gradient, attention, split, BatchNorm
 *     backward, feature-loss and optimizer entry points move data as 128-bit vectors or TMA boxes and REFUSE a misaligned
 *     tensor (MSTK_EINVAL, message naming the alignment); the remaining element-wise / reduction / layout calls accept
 *     any naturally aligned pointer and any element count (scalar path).  No call faults on a misaligned pointer
 *     (tests/test_gpu_ops_fuzz.py).
 */
#ifndef MSTK_H_
#define MSTK_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSTK_ABI_VERSION 1

enum { MSTK_OK = 0, MSTK_EINVAL = -1, MSTK_ECUDA = -2, MSTK_EUNSUPPORTED = -3 };
enum { MSTK_F32 = 0, MSTK_BF16 = 1 };
enum { MSTK_ACT_NONE = 0, MSTK_ACT_RELU = 1, MSTK_ACT_TANH = 2 };
/* kernel family used for a contraction.  AUTO = tcgen05 implicit GEMM whenever the
 * shape allows (bf16, Cin % 32 == 0, Cout % 16 == 0), else the CUDA-core kernel. */
enum { MSTK_BACKEND_AUTO = 0, MSTK_BACKEND_SIMT = 1, MSTK_BACKEND_TCGEN05 = 2 };

int mstk_version(void);
const char* mstk_last_error(void);
/* 1 when the tcgen05/TMA path could be initialised (driver entry point for
 * cuTensorMapEncodeTiled resolved); it is an error, not a fallback, if it is 0. */
int mstk_tcgen05_available(void);

/* ------------------------------------------------------------------------------
 * Fused DDIM update, applied in the epilogue of the last UNet conv (dec1) or
 * stand-alone.  Replaces model.py:441-458 / 535-552 (7 elementwise ATen launches):
 *   x0   = (x - s1m_t*eps) / sab_t
 *   x'   = sab_n*x0 + s1m_n*eps + eta*(s1m_n*eps - s1m_t*eps)
 * with sab = sqrt(alpha_bar), s1m = sqrt(1-alpha_bar) (host computes them in fp32
 * exactly as torch does).  NOTE the reference's eta only rescales the predicted
 * noise; no fresh noise is injected.
 * ---------------------------------------------------------------------------- */
typedef struct mstk_ddim_args {
    const float* x_cur;   /* [N,H,W,C] fp32 current latent                       */
    float* x_next;        /* [N,H,W,C] fp32 (may alias x_cur)                    */
    void* x_next_lp;      /* optional bf16 copy of x_next (next step's conv input) */
    float* x0_pred;       /* optional fp32 log of x0 (sampling_logs['pred_x0'])  */
    float sab_t, s1m_t, sab_n, s1m_n, eta;
} mstk_ddim_args;

/* ------------------------------------------------------------------------------
 * Convolution family: Conv2d k3 p1 (s1/s2), k1 (= nn.Linear over tokens),
 * ConvTranspose2d k3 s2 p1 op1 and k4 s2 p1, all as one tap-gather implicit GEMM.
 * Replaces F.conv2d / F.conv_transpose2d / F.linear + the elementwise ops that
 * follow them in the reference:
 *   UNet.forward            model.py:205-229 (bias, ReLU, +t_emb, +skip)
 *   SpectrogramEncoder      model.py:16-25   (conv; BatchNorm statistics gathered here)
 *   SpectrogramDecoder      model.py:37-46   (convT; tanh; (y+1)/2 of model.py:371)
 *   StyleEncoder            model.py:70-79
 *   MultiheadAttention in/out projections  model.py:155 (F.linear inside nn.MultiheadAttention)
 *   VGGish features         thirdparty/torchvggish/vggish.py:108-118 driven by loss.py:78-90
 * Epilogue order:  v = acc + bias ; stats += (v, v*v) ; v = act(v) ;
 *                  v = v*out_scale + out_shift ; v += add_nc[n,c] ; v += add[n,h,w,c] ;
 *                  store y (and/or DDIM update).
 * ---------------------------------------------------------------------------- */
typedef struct mstk_conv_args {
    const void* x;        /* [n,h,w,cin] NHWC, `dtype`                              */
    const void* wpacked;  /* [k*k][cout][cin], `dtype` (mstk_pack_conv_weight)      */
    const float* bias;    /* [cout] fp32 or NULL                                    */
    void* y;              /* [n,ho,wo,cout] NHWC, `out_dtype`; NULL allowed with ddim */
    const void* add;      /* optional NHWC tensor shaped like y, `dtype`            */
    const float* add_nc;  /* optional [n,cout] fp32 (time embedding, model.py:206)  */
    double* stats;        /* optional [2*cout]: sum, sum of squares of v (BatchNorm
                             batch statistics, model.py:18,21,24,39,42); accumulated
                             with atomics, caller zeroes it                         */
    const mstk_ddim_args* ddim; /* optional (HOST pointer)                          */
    void* workspace;      /* split-K scratch, zero-initialised once by the caller;
                             kernels leave it zeroed.  NULL = no split-K            */
    size_t workspace_bytes;
    int32_t n, h, w, cin, cout;
    int32_t ksize;        /* 1, 3 or 4                                              */
    int32_t stride;       /* 1 or 2                                                 */
    int32_t transposed;   /* 0 conv (pad = ksize/2 for k3, 0 for k1);
                             1 conv-transpose (k3 s2 p1 op1, or k4 s2 p1)           */
    int32_t act;          /* MSTK_ACT_*                                             */
    float out_scale, out_shift; /* applied after act; use 1,0 for identity          */
    int32_t dtype;        /* dtype of x, wpacked, add                               */
    int32_t out_dtype;    /* dtype of y                                             */
    int32_t backend;      /* MSTK_BACKEND_*                                         */
} mstk_conv_args;

int mstk_conv2d_fwd(const mstk_conv_args* a, void* stream);
/* bytes of split-K workspace that lets the tcgen05 path spread this problem over
 * the whole GPU (0 if split-K would not be used) */
size_t mstk_conv2d_workspace_bytes(const mstk_conv_args* a);

/* nn.Conv2d weight [cout,cin,k,k] fp32 -> [k*k][cout][cin] in `dtype`, optionally
 * scaled per output channel (eval-mode BatchNorm folding, model.py:18-24). */
int mstk_pack_conv_weight(const float* w_oihw, const float* scale_cout, void* dst,
                          int cout, int cin, int k, int dtype, void* stream);
/* nn.ConvTranspose2d weight [cin,cout,k,k] fp32 -> [k*k][cout][cin] in `dtype`. */
int mstk_pack_convt_weight(const float* w_iohw, const float* scale_cout, void* dst,
                           int cin, int cout, int k, int dtype, void* stream);

/* layout changes at the nn.Module boundary (reference tensors are NCHW fp32) */
int mstk_nchw_to_nhwc(const float* src, void* dst, int n, int c, int h, int w,
                      int dst_dtype, void* stream);
int mstk_nhwc_to_nchw(const void* src, float* dst, int n, int c, int h, int w,
                      int src_dtype, void* stream);
/* dtype cast of a flat buffer */
int mstk_cast(const void* src, int src_dtype, void* dst, int dst_dtype, size_t count, void* stream);

/* ------------------------------------------------------------------------------
 * BatchNorm2d: finalise batch statistics gathered by the producing conv and apply
 * normalise + affine (+ReLU) in one 128-bit vectorised streaming pass.  Replaces
 * nn.BatchNorm2d + nn.ReLU at model.py:18-19, 21-22, 24, 39-40, 42-43 (train mode:
 * biased batch variance; running buffers updated with momentum and the unbiased
 * variance).  Eval mode never reaches this call: it is folded into the conv.
 * ---------------------------------------------------------------------------- */
typedef struct mstk_bn_args {
    const void* x;          /* [count, c] NHWC rows, `dtype`                        */
    void* y;                /* same shape, `out_dtype` (may alias x if same dtype)  */
    const double* stats;    /* [2*c] sum, sumsq over `count` rows                   */
    const float* gamma;     /* [c]                                                  */
    const float* beta;      /* [c]                                                  */
    float* running_mean;    /* optional [c], updated in place                       */
    float* running_var;     /* optional [c]                                         */
    float* save_mean;       /* optional [c]: batch mean (for backward)              */
    float* save_invstd;     /* optional [c]: 1/sqrt(var+eps)                        */
    int64_t count;          /* statistics count = N*H*W over ALL ranks (what `stats` sums over) */
    int64_t rows;           /* rows of x / y held by THIS rank (0 = same as count)      */
    int32_t c;
    float eps, momentum;
    int32_t act;            /* MSTK_ACT_NONE or MSTK_ACT_RELU                       */
    int32_t dtype, out_dtype;
} mstk_bn_args;
int mstk_bn_apply(const mstk_bn_args* a, void* stream);

/* ------------------------------------------------------------------------------
 * UNet.time_mlp: sinusoidal embedding -> Linear -> GELU(erf) -> Linear.
 * Replaces model.py:170-175, 203, 239-246.  t is fp32 (the reference accepts int64
 * or float timesteps, models/tests.py:447); out is [b,128] fp32.
 * ---------------------------------------------------------------------------- */
int mstk_time_embed(const float* t, int b, const float* w1, const float* b1,
                    const float* w2, const float* b2, float* out, int dim, void* stream);

/* ------------------------------------------------------------------------------
 * Cross-attention core for nn.MultiheadAttention(E, heads) (model.py:155,
 * SURVEY A.3) after the projections: per (sample, head) softmax(q k^T / sqrt(d)) v.
 * q: [b, l, e]; k, v: [b, s, e] with row strides given in elements (k and v may be
 * slices of one fused [b, s, 2e] projection); out: [b, l, e].  All `dtype`.
 * ---------------------------------------------------------------------------- */
typedef struct mstk_attn_args {
    const void* q; const void* k; const void* v; void* out;
    int32_t b, l, s, e, heads;
    int32_t q_stride, k_stride, v_stride, out_stride;   /* elements between tokens */
    int32_t dtype;
} mstk_attn_args;
int mstk_attention_fwd(const mstk_attn_args* a, void* stream);

/* ------------------------------------------------------------------------------
 * fp32 mode on the tensor cores (csrc/split_f32.cu): [rows, c] fp32 -> [rows, 3c] bf16 with every value carried as
 * hi + lo; weights == 0: [hi | lo | hi] (activations, NHWC rows = pixels), weights != 0: [hi | hi | lo] (packed
 * conv weights, rows = taps * cout).  A bf16 mstk_conv2d_fwd over the 3c channels with fp32 output then equals the
 * fp32 contraction to ~1e-5 relative.  c % 8 == 0.
 * ---------------------------------------------------------------------------- */
int mstk_split3(const float* src, void* dst, size_t rows, int c, int weights, void* stream);

/* ------------------------------------------------------------------------------
 * Style cross-attention with the projections composed into the style side (same layer:
 * CrossAttention.forward, models/model.py:126-160; algebra in csrc/xattn.cu).  The style map
 * is projected ONCE per sampling loop by a k1 mstk_conv2d_fwd with the composed weight
 *   mstk_xattn_compose: in_proj_weight [3e,e], in_proj_bias [3e], out_proj.weight [e,e] (fp32)
 *                       -> w_packed [ctot, e] (`dtype`), bias [ctot] fp32,  ctot = mstk_xattn_ctot(e, heads)
 * and every evaluation of the layer is one launch:
 *   x: [b, l, e] features; kvp: [b, s, ctot] projected style tokens; bias: out_proj.bias [e] fp32; out: [b, l, e].
 *   kv_static != 0: kvp was produced at least two launches earlier on the stream (it may then be staged before the
 *   programmatic grid dependency resolves).  e in {256, 512}.
 * ---------------------------------------------------------------------------- */
typedef struct mstk_xattn_args {
    const void* x; const void* kvp; const float* bias; void* out;
    int32_t b, l, s, e, heads, ctot;
    int32_t kv_static;
    int32_t dtype;
} mstk_xattn_args;
int mstk_xattn_ctot(int e, int heads);
int mstk_xattn_compose(const float* in_proj_weight, const float* in_proj_bias, const float* out_proj_weight, void* w_packed,
                       float* bias, int e, int heads, int dtype, void* stream);
int mstk_xattn_fwd(const mstk_xattn_args* a, void* stream);

/* ------------------------------------------------------------------------------
 * Scheduler arithmetic (ForwardDiffusion, model.py:102-124), NHWC or NCHW agnostic:
 * `inner` = elements per sample.
 *   q_sample:   z_t = sqrt(ab[t_b])*x0 + sqrt(1-ab[t_b])*eps
 *   predict_x0: x0  = (z_t - sqrt(1-ab[t_b])*eps_hat) / sqrt(ab[t_b])
 * `z_lp` / `x0_lp`: optional bf16 copy for the next conv.
 * ---------------------------------------------------------------------------- */
int mstk_q_sample(const float* x0, const float* eps, const float* alpha_bar, const int64_t* t,
                  float* z_t, void* z_lp, int b, int64_t inner, void* stream);
int mstk_predict_x0(const float* z_t, const float* eps_hat, const float* alpha_bar, const int64_t* t,
                    float* x0, void* x0_lp, int b, int64_t inner, void* stream);
/* stand-alone DDIM update over `count` elements (same math as the fused epilogue);
 * eps is fp32 */
int mstk_ddim_step(const mstk_ddim_args* a, const float* eps, size_t count, void* stream);

/* ------------------------------------------------------------------------------
 * Loss reductions (models/loss.py).  Each writes ONE fp32 scalar to `out` (device)
 * using a double accumulator `scratch` ([4] doubles + [1] uint32 counter, zeroed by
 * the caller once; kernels leave it zeroed).
 *   mse:  mean((a-b)^2)                      loss.py:36,49
 *   kl:   mean(0.5(z^2 - 1 - log(z^2+1e-8))) loss.py:31-32
 * ---------------------------------------------------------------------------- */
int mstk_mse(const void* a, const void* b, int dtype, size_t count, double* scratch, float* out, void* stream);
int mstk_kl(const float* z, size_t count, double* scratch, float* out, void* stream);
/* VGGish tap term (loss.py:94-99): per sample i, f/(std_unbiased(f_i)+1e-8) vs
 * g/(std_unbiased(g_i)+1e-8), mean squared difference over all b*inner elements.
 * tmp: [b*4 + 3] doubles scratch (zeroed by caller; left zeroed). Accumulates
 * `weight * term` into out[0] (so the six taps can be summed with weight 1/6). */
int mstk_feature_std_mse(const void* f, const void* g, int dtype, int b, int64_t inner,
                         double* tmp, float weight, float* out, void* stream);
/* The same term for the training step, which evaluates it on the pairs (f0,f1) and (f1,f2)
 * (compression_loss's perceptual term loss.py:40 and style_loss loss.py:104-106 share the
 * reconstruction's features): per-sample statistics in one pass, both pair terms in one pass.
 *   mstk_sample_stats : stats[s*2 + {0,1}] += sum, sum of squares of f[s, 0..inner)
 *   mstk_feature_mse3 : out_a += weight * mean((f0/(std0+1e-8) - f1/(std1+1e-8))^2),
 *                       out_c += weight * mean((f1/(std1+1e-8) - f2/(std2+1e-8))^2)  (f2 may be NULL)
 * st0/st1/st2: [2*b] doubles from mstk_sample_stats; scratch: 3 doubles.  Statistics and scratch
 * are zeroed by the caller once and left zeroed.  inner % 8 == 0 (bf16) / 4 (fp32). */
int mstk_sample_stats(const void* f, int dtype, int samples, int64_t inner, double* stats, void* stream);
int mstk_feature_mse3(const void* f0, const void* f1, const void* f2, int dtype, int b, int64_t inner,
                      double* st0, double* st1, double* st2, double* scratch, float weight,
                      float* out_a, float* out_c, void* stream);
/* One-pass form of the same two terms for a whole tap: f = [3*b, h, w, c] (x, reconstruction, style stacked);
 * per-sample sums and cross terms in one read of f, the terms follow algebraically
 * (sum (x r_x - y r_y)^2 = r_x^2 Sxx - 2 r_x r_y Sxy + r_y^2 Syy); `pooled` (optional, [3*b, h/2, w/2, c])
 * receives the 2x2/2 max pool of f (the next VGGish conv's input, vggish.py:113) from the same read.
 * acc: [8*b] doubles, scratch: 1 double; zeroed by the caller once, left zeroed. */
int mstk_feature_fused3(const void* f, void* pooled, int dtype, int b, int h, int w, int c, double* acc,
                        double* scratch, float weight, float* out_a, float* out_c, void* stream);
/* The FIRST tap of the same training-step form without materialising it: img [3b, h, w] fp32 single-channel images
 * (x; recon; style), wpacked [9][64] / bias [64] fp32 = features.0 (vggish.py:108-118, Conv2d(1, 64, 3, padding=1) + ReLU).
 * Accumulates the tap's two loss terms into out_a / out_c exactly as mstk_feature_fused3 does and writes ONLY the 2x2
 * max-pooled map [3b, h/2, w/2, 64] (`pooled_dtype`).  acc / scratch: as mstk_feature_fused3. */
int mstk_feature_first3(const float* img, const float* wpacked, const float* bias, void* pooled, int pooled_dtype, int b, int h, int w,
                        double* acc, double* scratch, float weight, float* out_a, float* out_c, void* stream);
/* 2x2/2 max pool on NHWC (vggish.py:113) */
int mstk_maxpool2x2(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream);

/* ------------------------------------------------------------------------------
 * Packed uint8 spectrogram cache -> batch (models/dataset.py:28-53, 206-260: PNG -> crop (0,0,128,128) -> grayscale
 * -> ToTensor): pool [n, h, w] uint8 images as stored (128 x 130 for the reference's 3 s chunks), index [b] int64
 * image numbers; out [b, 1, crop_h, crop_w] fp32 = top-left crop / 255.  Indices are not range-checked.
 * ---------------------------------------------------------------------------- */
int mstk_u8_gather_crop(const uint8_t* pool, const int64_t* index, float* out, int b, int h, int w, int crop_h, int crop_w,
                        void* stream);

/* ------------------------------------------------------------------------------
 * Mel front end (data/audio_processor.py:44-73 -> librosa.feature.melspectrogram +
 * power_to_db(ref=np.max) + the uint8 image quantisation):
 * wave [b, len] fp32 -> frames (n_fft 2048, hop 512, periodic Hann, centre zero pad)
 * -> |rFFT|^2 -> Slaney mel (n_mels) -> dB re per-clip max, floor -80 ->
 * img = trunc(clip((S+80)*255/80,0,255)+0.5)/255 in [0,1], cropped to `crop_frames`.
 * out: [b, n_mels, crop_frames] fp32 (row 0 = lowest band).  db_out optional
 * [b, n_mels, frames] (the un-quantised dB values).  work: see mstk_mel_workspace_bytes.
 * ---------------------------------------------------------------------------- */
typedef struct mstk_mel_args {
    const float* wave; float* out; float* db_out;
    const float* mel_fb;     /* [n_mels, 1025] filterbank (host builds it, fp32)   */
    void* work; size_t work_bytes;
    int32_t b, len, n_mels, crop_frames;
    const int32_t* band_ranges;  /* optional [n_mels][2]: first / one-past-last non-zero bin of every band, from
                                  * mstk_mel_band_ranges (once per filterbank); NULL = derived on every call */
} mstk_mel_args;
size_t mstk_mel_workspace_bytes(int b, int len, int n_mels);
int mstk_mel_band_ranges(const float* mel_fb, int n_mels, int32_t* ranges, void* stream);
int mstk_mel_frontend(const mstk_mel_args* a, void* stream);
/* Allocation-free variant of mstk_mel_frontend.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results as mstk_mel_frontend; a->work / a->work_bytes are IGNORED: the scratch (mel power,
 * clip maxima, band ranges) is a file-scope `static __device__` array of mstk_mel_static_scratch_bytes() = 8 MiB that
 * every GPU loading the library holds until exit.  Problems with mstk_mel_workspace_bytes(b, len, n_mels) above that
 * limit are rejected with MSTK_EINVAL and a message (64 clips x 128 bands x 3 s at 22.05 kHz need 4.1 MiB).  Calls no
 * cudaMalloc / cudaFree / allocating library routine.  All calls on a device share the one array: calls that overlap on
 * the GPU (different streams, no ordering) race; calls on one stream, or ordered by events, are safe -- everything read
 * is cleared or overwritten first by the call itself. */
size_t mstk_mel_static_scratch_bytes(void);
int mstk_mel_frontend_static(const mstk_mel_args* a, void* stream);

/* ------------------------------------------------------------------------------
 * Inverse audio path (data/audio_processor.py:81-100: librosa.db_to_power + librosa.feature.inverse.mel_to_audio;
 * :137-162: librosa.db_to_amplitude + librosa.griffinlim).  SURVEY 8(f) rank 2.  Every spectrogram-shaped array on this
 * side of the ABI is FRAME-MAJOR: [b, frames, 1025].
 *
 * mstk_mel_to_stft = audio_processor.py:93-97 + librosa.feature.inverse.mel_to_stft: image [b, n_mels, frames] uint8
 * (row 0 = lowest band) -> dB = u8 * max_db / 255 - max_db -> power 10^(dB/10) -> x = clip(pinv(mel_fb) @ power, 0)
 * (librosa.util.nnls's starting point, and its result whenever L-BFGS-B's projected-gradient test passes there: always
 * for a dB image, whose power is <= 1) -> mag = sqrt(x).  pinv: [1025, n_mels] fp32 (host: numpy.linalg.pinv of the
 * filterbank, a constant of (sr, n_mels)).  pg_norm (optional, [b]) receives the L-BFGS-B projected-gradient norm
 * max |proj(A^T (A x - B) / B.size)| of every clip's result: the caller compares it with scipy's pgtol = 1e-5.  pg_iters > 0
 * runs that many projected-gradient steps x <- max(x - pg_step * A^T (A x - B), 0) after the initialisation (a stand-in
 * for the L-BFGS-B iterations the reference would run on an input outside the dB-image domain; not the same iterates).
 *
 * mstk_db_image_to_mag = audio_processor.py:150-155: image [b, bins, frames] uint8 -> mag [b, frames, bins] amplitude.
 *
 * mstk_griffinlim = librosa.griffinlim(S, n_iter, hop 512, Hann, center, pad_mode constant, momentum, init random):
 * mag [b, frames, 1025]; phase_u [b, frames, 1025] uniform draws in [0,1) (the caller's RNG: initial phase 2 pi u);
 * audio [b, 512 * (frames - 1)] fp32.  n_fft = 2048 only.  2 launches per iteration (inverse STFT; forward STFT fused
 * with the overlap-add read, the momentum update and the magnitude projection).
 * ---------------------------------------------------------------------------- */
typedef struct mstk_mel_to_stft_args {
    const uint8_t* image; const float* pinv; const float* mel_fb; const int32_t* band_ranges;
    float* mag; float* pg_norm;
    int32_t b, n_mels, frames;
    float max_db;
    int32_t pg_iters; float pg_step;
} mstk_mel_to_stft_args;
int mstk_mel_to_stft(const mstk_mel_to_stft_args* a, void* stream);
int mstk_db_image_to_mag(const uint8_t* image, float* mag, int b, int bins, int frames, float max_db, void* stream);
typedef struct mstk_griffinlim_args {
    const float* mag; const float* phase_u; float* audio;
    void* work; size_t work_bytes;
    int32_t b, frames, n_iter;
    float momentum;
} mstk_griffinlim_args;
size_t mstk_griffinlim_workspace_bytes(int b, int frames);
int mstk_griffinlim(const mstk_griffinlim_args* a, void* stream);

/* ==============================================================================
 * Backward pass and optimizer of the training step (models/train.py:163-208).  The
 * reference gets these from autograd + torch.optim.Adam; here every gradient op is an
 * explicit kernel so that the whole step is one graph-capturable stream of launches.
 * Data gradients of the contractions re-use mstk_conv2d_fwd on a re-packed weight
 * (conv s2 <-> conv-transpose s2; stride-1 convs and nn.Linear through the dgrad pack).
 * ============================================================================== */

/* Conv2d / Linear weight [cout,cin,k,k] fp32 -> [k*k][cin][cout] in `dtype`, taps flipped:
 * the packed weight of the stride-1 convolution (k3 p1, or k1) that maps dL/dy to dL/dx. */
int mstk_pack_conv_weight_dgrad(const float* w_oihw, void* dst, int cout, int cin, int k, int dtype, void* stream);

/* One-shot all-reduce (sum) of a small fp64 vector over peer-mapped memory, for the BatchNorm statistic exchanges of the
 * data-parallel step (models/train.py:163-208 run under DDP with global-batch statistics; the reference is
 * single-process and has no exchange).  One process per GPU: every rank allocates one exchange buffer
 * (mstk_peer_alloc -> device pointer + 64-byte CUDA IPC handle), gathers the world's handles with its own plumbing,
 * maps the others (mstk_peer_open) and fills mstk_peer_ctx.  mstk_peer_allreduce_f64 then replaces data[0..n) by the sum
 * over ranks with ONE single-CTA launch per rank: push to every peer, system fence, epoch flag, spin on the own flags,
 * sum in rank order (replicas stay bit-identical).  Every rank must issue the same sequence of (slot, n) calls; a slot may
 * be reused once per step (32 slots, n <= 512).  Graph-capturable (the epoch lives in device memory).  A peer that never
 * arrives makes the launch fail after ~2 s instead of hanging. */
#define MSTK_PEER_MAX_WORLD 8
#define MSTK_PEER_HANDLE_BYTES 64
typedef struct mstk_peer_ctx {
    void* bufs[MSTK_PEER_MAX_WORLD];   /* exchange buffer of every rank as mapped in THIS process (bufs[rank] = own) */
    int32_t rank, world;
} mstk_peer_ctx;
size_t mstk_peer_buffer_bytes(void);
int mstk_peer_alloc(void** dev_ptr, void* ipc_handle_out);
int mstk_peer_open(const void* ipc_handle, void** dev_ptr);
int mstk_peer_close(void* dev_ptr);
int mstk_peer_free(void* dev_ptr);
int mstk_peer_allreduce_f64(double* data, int n, int slot, const mstk_peer_ctx* ctx, void* stream);

/* All packed copies of a step's trainable weights in ONE launch: entry i re-does mstk_pack_conv_weight (mode 0),
 * mstk_pack_convt_weight (mode 1) or mstk_pack_conv_weight_dgrad (mode 2) without a scale, kk = k * k.  The table lives in
 * DEVICE memory; first_block = running sum of mstk_pack_blocks(cout, cin, kk) over the entries before it, total_blocks the
 * sum over all.  (The reference re-reads nn.Parameter storage in every cuDNN call; here the trainer refreshes the packed
 * copies once per optimizer step, models/train.py:200.) */
typedef struct mstk_pack_entry {
    const float* src; void* dst;
    int32_t cout, cin, kk, mode, dtype, first_block;
} mstk_pack_entry;
int mstk_pack_blocks(int cout, int cin, int kk);
int mstk_pack_multi(const mstk_pack_entry* entries_dev, int n_entries, int total_blocks, void* stream);

/* Weight gradient of the convolution family: for every kernel tap,
 *   dW[tap][co][ci] += sum over pixels of dy[out pixel, co] * x[in pixel shifted by tap, ci],
 * accumulated (atomically, fp32) straight into the gradient tensor in the PARAMETER's own
 * layout: [cout][cin][k][k] for Conv2d / Linear, [cin][cout][k][k] for ConvTranspose2d.
 * Geometry arguments are those of the FORWARD call (mstk_conv_args). */
typedef struct mstk_wgrad_args {
    const void* x;        /* [n,h,w,cin] forward input, x_dtype                          */
    const void* dy;       /* [n,ho,wo,cout] gradient w.r.t. the conv output, dy_dtype    */
    float* dw;            /* fp32 parameter-layout gradient, caller zeroes it per step   */
    int32_t n, h, w, cin, cout, ksize, stride, transposed;
    int32_t x_dtype, dy_dtype;
} mstk_wgrad_args;
int mstk_conv2d_wgrad(const mstk_wgrad_args* a, void* stream);

/* Activation backward + bias gradient in one streaming pass over [rows, c]:
 *   g = (dy [+ dy2]) * act'(.) * y_scale ;  dbias[c] += column sums of g
 * act' comes from the saved forward OUTPUT y: ReLU -> (y > 0); tanh -> 1 - t^2 with
 * t = (y - y_shift) / y_scale (the forward's fused out_scale/out_shift); NONE -> 1. */
typedef struct mstk_act_bwd_args {
    const void* dy; const void* dy2;   /* dy2 optional (skip-connection gradient), same dtype */
    const void* y;                      /* forward output (ignored for MSTK_ACT_NONE)          */
    void* g;                            /* may be NULL when only dbias is wanted               */
    float* dbias;                       /* optional [c], accumulated                           */
    int64_t rows; int32_t c;
    int32_t act; float y_scale, y_shift;
    int32_t dy_dtype, y_dtype, g_dtype;
} mstk_act_bwd_args;
int mstk_act_bwd(const mstk_act_bwd_args* a, void* stream);

/* BatchNorm2d (training mode) backward, two passes over v = the conv output that fed the
 * norm: (1) sums[0:c] = sum g, sums[c:2c] = sum g*xhat with g = dy * (relu mask of the
 * recomputed output); all-reduce `sums` across ranks here under data parallelism;
 * (2) dv = gamma*invstd*(g - sums0/count - xhat*sums1/count), dgamma += sums1, dbeta += sums0. */
typedef struct mstk_bn_bwd_args {
    const void* dy; const void* v; void* dv;
    const float* mean; const float* invstd; const float* gamma; const float* beta;
    double* sums;            /* [2*c], zeroed by the caller before pass 1 */
    float* dgamma; float* dbeta;
    int64_t rows;            /* local rows */
    int64_t count;           /* global rows (= rows * world size) */
    int32_t c, act, dy_dtype, v_dtype, dv_dtype;
} mstk_bn_bwd_args;
int mstk_bn_bwd_reduce(const mstk_bn_bwd_args* a, void* stream);
int mstk_bn_bwd_apply(const mstk_bn_bwd_args* a, void* stream);

/* Cross-attention core backward (recomputes the probabilities): given d_out [b,l,e],
 * writes dq [b,l,e], dk and dv [b,s,e] (strides as in mstk_attn_args; dk/dv may be the two
 * halves of one [b,s,2e] buffer). */
typedef struct mstk_attn_bwd_args {
    const void* q; const void* k; const void* v; const void* d_out;
    void* dq; void* dk; void* dv;
    int32_t b, l, s, e, heads;
    int32_t q_stride, k_stride, v_stride, do_stride, dq_stride, dk_stride, dv_stride;
    int32_t dtype;
} mstk_attn_bwd_args;
int mstk_attention_bwd(const mstk_attn_bwd_args* a, void* stream);

/* time MLP with the intermediates the backward needs (e = sinusoid, pre = Linear1 output, h = GELU(pre)) */
int mstk_time_embed_train(const float* t, int b, const float* w1, const float* b1, const float* w2, const float* b2,
                          float* out, float* e_out, float* pre_out, float* h_out, int dim, void* stream);
/* dpre = dh * gelu'(pre) (exact erf GELU), fp32 */
int mstk_gelu_bwd(const float* dh, const float* pre, float* dpre, size_t count, void* stream);
/* out[b, c] = sum over hw of (a[b,hw,c] + a2[b,hw,c]) ; a2 optional; fp32 out */
int mstk_sum_hw(const void* a, const void* a2, int dtype, int b, int64_t hw, int c, float* out, void* stream);
/* out = y + add (same shape, optional) + add_nc[n, c] (optional): the un-fused form of the conv epilogue adds,
 * used in training where the pre-add ReLU output must be kept for the mask */
int mstk_add_bcast(const void* y, const void* add, const float* add_nc, void* out, int n, int64_t hw, int c, int dtype, void* stream);
/* x[i] *= *factor (device scalar): chains a loss gradient with the upstream scalar of its autograd node */
int mstk_scale_dev(float* x, const float* factor, size_t count, void* stream);
/* g = scale * (a - b), a and b fp32, g in g_dtype: gradient of mean((a-b)^2) with scale = 2*w/count */
int mstk_mse_grad(const float* a, const float* b, float scale, void* g, int g_dtype, size_t count, void* stream);
/* gradient w.r.t. eps_hat of  w_d*mean((eps_hat-eps)^2) + L(x0_pred(eps_hat)):
 *   g = scale_d*(eps_hat - eps) - sqrt(1-ab[t])/sqrt(ab[t]) * dx0      (model.py:117-124, loss.py:48) */
int mstk_eps_grad(const void* dx0, int dx0_dtype, const float* eps_hat, const float* eps, const float* alpha_bar, const int64_t* t,
                  float scale_d, void* g, int g_dtype, int b, int64_t inner, void* stream);
/* torch.optim.Adam (no amsgrad, no weight decay) over a flat fp32 arena; `step` is the 1-based step count */
int mstk_adam_step(float* p, const float* g, float* m, float* v, size_t count, float lr, float beta1, float beta2, float eps,
                   int64_t step, float grad_scale, void* stream);
/* The same update with the step count (>= 1) read from device memory, so that a CUDA graph of the whole training
 * step can be replayed (the caller increments *step_dev on the stream before this call). */
int mstk_adam_step_dev(float* p, const float* g, float* m, float* v, size_t count, float lr, float beta1, float beta2, float eps,
                       const int64_t* step_dev, float grad_scale, void* stream);
/* torch.optim.AdamW (decoupled weight decay: p *= 1 - lr*weight_decay before the Adam update), the optimizer of the
 * reference's autoencoder stage (models/train.py:47, default weight_decay 0.01); step count in device memory. */
int mstk_adamw_step_dev(float* p, const float* g, float* m, float* v, size_t count, float lr, float beta1, float beta2, float eps,
                        float weight_decay, const int64_t* step_dev, float grad_scale, void* stream);
/* g = dz (optional, dz_dtype) + scale * (z - z/(z^2 + 1e-8)): gradient of scale*count * kl_regularization_loss(z)
 * (models/loss.py:31-32) added to the gradient arriving from the decoder; z and g fp32 */
int mstk_kl_grad(const float* z, const void* dz, int dz_dtype, float scale, float* g, size_t count, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MSTK_H_ */
