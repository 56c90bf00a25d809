// Inverse audio path: uint8 dB mel image -> mel power -> STFT magnitude (NNLS initialisation = clipped pseudo-inverse,
// which is where librosa's L-BFGS-B stops for a dB image) -> 32 Griffin-Lim iterations -> waveform.  Replaces
// librosa.db_to_power + librosa.feature.inverse.mel_to_audio of data/audio_processor.py:81-100 (and db_to_amplitude +
// librosa.griffinlim of :137-162).  SURVEY 8(f) rank 2.
//
// Layouts: every spectrogram-shaped array is FRAME-MAJOR, [clip][frame][1025], so that the warp that owns a frame
// reads and writes 256-byte runs.  One warp per frame, 1024-point complex FFT in registers (fft_warp.cuh):
//   gl_istft_kernel : X[k] -> Z[k] = E[k] + i O[k] (inverse of the real-spectrum split) -> IFFT as conj(FFT(conj Z))
//                     / 1024 -> * Hann -> windowed frame [2048] (overlap-add is deferred to the reader)
//   gl_stft_kernel  : sample m of the padded signal = (sum of the <= 4 windowed frames covering m) / window sum-square,
//                     zero outside [1024, 1024 + len) (librosa trims the centre padding, then pads with zeros again)
//                     -> * Hann -> FFT -> split -> rebuilt X[k]; then the Griffin-Lim update in the same pass:
//                     a = X - momentum / (1 + momentum) * tprev;  angles = a / (|a| + tiny) * S;  tprev = X
//   gl_ola_kernel   : the final overlap-add + normalisation + trim -> audio [512 * (frames - 1)]
// State per iteration: angles + tprev (complex64) + S + windowed frames = 28.7 KB per frame; 64 clips x 128 frames =
// 235 MB, i.e. HBM-streamed (not L2-resident); the FFT work is ~3 k warp instructions per frame and direction.
#include "common.cuh"
#include "fft_warp.cuh"

namespace mstk {

constexpr int GL_WARPS = 8, GL_FPW = 2, GL_FPC = GL_WARPS * GL_FPW;       // frames per CTA
constexpr float GL_TINY = 1.17549435e-38f;                                   // numpy.finfo(float32).tiny

// ---------------------------------------------------------------------------------------------------------------
// mel image -> STFT magnitude
// ---------------------------------------------------------------------------------------------------------------
constexpr int M2S_FR = 16;          // frames per CTA
constexpr int M2S_THREADS = 256;

// image [b][n_mels][frames] uint8 -> mag [b][frames][1025] = sqrt(max(pinv @ power, 0)) (+ optional projected-gradient
// iterations on 0.5 |A x - B|^2, x >= 0) and the L-BFGS-B projected-gradient norm of the result per clip.
__global__ void __launch_bounds__(M2S_THREADS) mel_to_stft_kernel(const uint8_t* __restrict__ image, const float* __restrict__ pinv,
                                                                 const float* __restrict__ fb, const int2* __restrict__ ranges, float* __restrict__ mag,
                                                                 unsigned int* __restrict__ pg_norm, int n_mels, int frames, float max_db,
                                                                 int pg_iters, float pg_step) {
    extern __shared__ __align__(16) uint8_t m2s_smem[];
    float* xs = reinterpret_cast<float*>(m2s_smem);                 // [NBINS][M2S_FR]
    float* Mc = xs + NBINS * M2S_FR;                                // [n_mels][M2S_FR]  mel power (B)
    float* rs = Mc + n_mels * M2S_FR;                               // [n_mels][M2S_FR]  residual A x - B
    const int tid = threadIdx.x, clip = blockIdx.y, f0 = blockIdx.x * M2S_FR;
    const int nf = min(M2S_FR, frames - f0);
    for (int i = tid; i < n_mels * M2S_FR; i += M2S_THREADS) {
        const int m = i / M2S_FR, j = i - m * M2S_FR;
        float p = 0.f;
        if (j < nf) {
            const float db = (float)image[((size_t)clip * n_mels + m) * frames + f0 + j] * (max_db / 255.0f) - max_db;
            p = exp10f(0.1f * db);                                   // librosa.db_to_power
        }
        Mc[i] = p;
    }
    __syncthreads();
    // x_init = clip(pinv(A) @ B, 0, None)
    for (int k = tid; k < NBINS; k += M2S_THREADS) {
        float acc[M2S_FR];
#pragma unroll
        for (int j = 0; j < M2S_FR; ++j) acc[j] = 0.f;
        const float* pr = pinv + (size_t)k * n_mels;
        for (int m = 0; m < n_mels; ++m) {
            const float pv = __ldg(pr + m);
            const float4* mc = reinterpret_cast<const float4*>(Mc + m * M2S_FR);
#pragma unroll
            for (int j4 = 0; j4 < M2S_FR / 4; ++j4) {
                const float4 t = mc[j4];
                acc[4 * j4] = fmaf(pv, t.x, acc[4 * j4]); acc[4 * j4 + 1] = fmaf(pv, t.y, acc[4 * j4 + 1]);
                acc[4 * j4 + 2] = fmaf(pv, t.z, acc[4 * j4 + 2]); acc[4 * j4 + 3] = fmaf(pv, t.w, acc[4 * j4 + 3]);
            }
        }
#pragma unroll
        for (int j = 0; j < M2S_FR; ++j) xs[k * M2S_FR + j] = fmaxf(acc[j], 0.f);
    }
    __syncthreads();
    const float inv_size = 1.0f / ((float)n_mels * (float)frames);   // 1 / B.size of the clip's problem
    float pgmax = 0.f;
    for (int it = 0; it <= pg_iters; ++it) {
        // residual r = A x - B over each band's non-zero bin range
        for (int i = tid; i < n_mels * M2S_FR; i += M2S_THREADS) {
            const int m = i / M2S_FR, j = i - m * M2S_FR;
            const int2 rg = ranges[m];
            const float* fr = fb + (size_t)m * NBINS;
            float acc = 0.f;
            for (int k = rg.x; k < rg.y; ++k) acc = fmaf(__ldg(fr + k), xs[k * M2S_FR + j], acc);
            rs[i] = acc - Mc[i];
        }
        __syncthreads();
        const bool last = it == pg_iters;
        for (int k = tid; k < NBINS; k += M2S_THREADS) {
            float g[M2S_FR];
#pragma unroll
            for (int j = 0; j < M2S_FR; ++j) g[j] = 0.f;
            for (int m = 0; m < n_mels; ++m) {
                const int2 rg = ranges[m];
                if (k < rg.x || k >= rg.y) continue;
                const float a = __ldg(fb + (size_t)m * NBINS + k);
#pragma unroll
                for (int j = 0; j < M2S_FR; ++j) g[j] = fmaf(a, rs[m * M2S_FR + j], g[j]);
            }
            if (last) {
                // L-BFGS-B projgr for bounds (0, None): g >= 0 -> min(x - 0, g), g < 0 -> g; norm = max |.|
#pragma unroll
                for (int j = 0; j < M2S_FR; ++j) {
                    if (j >= nf) continue;
                    const float gs = g[j] * inv_size, x = xs[k * M2S_FR + j];
                    pgmax = fmaxf(pgmax, gs >= 0.f ? fminf(x, gs) : -gs);
                }
            } else {
#pragma unroll
                for (int j = 0; j < M2S_FR; ++j) xs[k * M2S_FR + j] = fmaxf(xs[k * M2S_FR + j] - pg_step * g[j], 0.f);
            }
        }
        __syncthreads();
    }
    if (pg_norm) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pgmax = fmaxf(pgmax, __shfl_xor_sync(0xffffffffu, pgmax, o));
        if ((tid & 31) == 0 && pgmax > 0.f) atomicMax(pg_norm + clip, __float_as_uint(pgmax));
    }
    for (int j = 0; j < nf; ++j)
        for (int k = tid; k < NBINS; k += M2S_THREADS) mag[((size_t)clip * frames + f0 + j) * NBINS + k] = sqrtf(xs[k * M2S_FR + j]);
}

// image [b][bins][frames] uint8 dB -> mag [b][frames][bins] amplitude (librosa.db_to_amplitude = db_to_power ** 0.5)
__global__ void db_image_to_mag_kernel(const uint8_t* __restrict__ image, float* __restrict__ mag, int b, int bins, int frames, float max_db) {
    const size_t total = (size_t)b * bins * frames;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % bins);
        const size_t r = i / bins;
        const int f = (int)(r % frames), clip = (int)(r / frames);
        const float db = (float)image[((size_t)clip * bins + k) * frames + f] * (max_db / 255.0f) - max_db;
        mag[i] = sqrtf(exp10f(0.1f * db));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Griffin-Lim
// ---------------------------------------------------------------------------------------------------------------
// window sum-square of the overlap-add (librosa.filters.window_sumsquare, float32 accumulation in frame order)
__global__ void gl_wss_kernel(float* __restrict__ inv_wss, int frames, int n) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n) return;
    float s = 0.f;
    const int fhi = min(frames - 1, m / HOP), flo = m >= NFFT ? (m - NFFT) / HOP + 1 : 0;
    for (int f = flo; f <= fhi; ++f) {
        const int j = m - f * HOP;
        if (j < 0 || j >= NFFT) continue;
        const double w = 0.5 - 0.5 * cospi(2.0 * (double)j / (double)NFFT);
        s = (float)((double)s + w * w);                  // float32 array += float64 window^2, as the reference
    }
    inv_wss[m] = s > GL_TINY ? 1.0f / s : 1.0f;
}

// angles = S * exp(2 pi i U)
__global__ void gl_init_kernel(const float* __restrict__ mag, const float* __restrict__ u, float2* __restrict__ spec, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        float sn, cs;
        sincospif(2.0f * u[i], &sn, &cs);
        const float s = mag[i];
        spec[i] = make_float2(s * cs, s * sn);
    }
}

__global__ void __launch_bounds__(GL_WARPS * 32, 2) gl_istft_kernel(const float2* __restrict__ spec, float* __restrict__ fr, int frames) {
    extern __shared__ __align__(16) uint8_t gl_smem[];
    float2* tw = reinterpret_cast<float2*>(gl_smem);                          // [FFT_TW] e^{-2 pi i m / 2048}
    float2* bufs = tw + FFT_TW;                                               // [GL_WARPS][FFT_BUF]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int clip = blockIdx.y, f0 = blockIdx.x * GL_FPC;
    fft_fill_twiddles(tw, tid, GL_WARPS * 32);
    __syncthreads();
    float2* buf = bufs + warp * FFT_BUF;
#pragma unroll 1
    for (int fi = 0; fi < GL_FPW; ++fi) {
        const int frame = f0 + warp * GL_FPW + fi;
        if (frame >= frames) break;
        const float2* X = spec + ((size_t)clip * frames + frame) * NBINS;
        float2 v[32];
#pragma unroll
        for (int a = 0; a < 32; ++a) {
            const int k = 32 * a + lane;
            const float2 xk = __ldg(X + k), xp = __ldg(X + NC - k);
            float2 z;
            if (k == 0) {                               // irfft ignores the imaginary parts of the DC and Nyquist bins
                z = make_float2(0.5f * (xk.x + xp.x), 0.5f * (xk.x - xp.x));
            } else {
                const float ex = 0.5f * (xk.x + xp.x), ey = 0.5f * (xk.y - xp.y);       // E = (X[k] + conj X[1024-k]) / 2
                const float dx = 0.5f * (xk.x - xp.x), dy = 0.5f * (xk.y + xp.y);       // D = (X[k] - conj X[1024-k]) / 2
                const float2 t = tw[tw_slot(k)];                                       // e^{-2 pi i k / 2048}; O = D * conj(t)
                const float ox = dx * t.x + dy * t.y, oy = dy * t.x - dx * t.y;
                z = make_float2(ex - oy, ey + ox);                                     // Z = E + i O
            }
            v[a] = make_float2(z.x, -z.y);                                             // conj: IFFT(Z) = conj(FFT(conj Z)) / 1024
        }
        fft1024_warp(v, tw, buf, lane);
        float2* out = reinterpret_cast<float2*>(fr + ((size_t)clip * frames + frame) * NFFT);
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const int n = lane + 32 * brev5(r);
            const float c0 = tw[tw_slot(2 * n)].x, c1 = tw[tw_slot(2 * n + 1)].x;
            const float x0 = v[r].x * (1.0f / 1024.0f), x1 = -v[r].y * (1.0f / 1024.0f);
            out[n] = make_float2(x0 * (0.5f - 0.5f * c0), x1 * (0.5f - 0.5f * c1));
        }
    }
}

// overlap-added, normalised, centre-trimmed sample pair (m, m + 1) of the PADDED signal (m even)
__device__ __forceinline__ float2 gl_sample_pair(const float* __restrict__ frc, const float* __restrict__ inv_wss, int frames, int len, int m) {
    if (m < NFFT / 2 || m >= NFFT / 2 + len) return make_float2(0.f, 0.f);
    const int fhi = min(frames - 1, m / HOP);
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 3; j >= 0; --j) {                      // ascending frame order, as the reference's overlap-add
        const int f = fhi - j;
        const int off = m - f * HOP;
        if (f < 0 || off >= NFFT) continue;
        const float2 t = *reinterpret_cast<const float2*>(frc + (size_t)f * NFFT + off);
        s.x += t.x; s.y += t.y;
    }
    const float2 w = *reinterpret_cast<const float2*>(inv_wss + m);
    return make_float2(s.x * w.x, s.y * w.y);
}

__global__ void __launch_bounds__(GL_WARPS * 32, 2) gl_stft_kernel(const float* __restrict__ fr, const float* __restrict__ inv_wss,
                                                                 const float* __restrict__ mag, float2* __restrict__ spec, float2* __restrict__ tprev,
                                                                 int frames, int first, float mom) {
    extern __shared__ __align__(16) uint8_t gl_smem[];
    float2* tw = reinterpret_cast<float2*>(gl_smem);                          // [FFT_TW] e^{-2 pi i m / 2048}
    float2* bufs = tw + FFT_TW;                                               // [GL_WARPS][FFT_BUF]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int clip = blockIdx.y, f0 = blockIdx.x * GL_FPC;
    fft_fill_twiddles(tw, tid, GL_WARPS * 32);
    __syncthreads();
    float2* buf = bufs + warp * FFT_BUF;
    const int len = HOP * (frames - 1);
    const float* frc = fr + (size_t)clip * frames * NFFT;
#pragma unroll 1
    for (int fi = 0; fi < GL_FPW; ++fi) {
        const int frame = f0 + warp * GL_FPW + fi;
        if (frame >= frames) break;
        // Sample m = frame * 512 + i of the padded signal, i = 2 (32 a + lane): its hop quarter q = i / 512 = a / 8 is the same for
        // the whole warp, so the covering frames f' = frame + q - j (j = 0..3), their validity and the centre-trim test
        // (2 <= frame + q <= frames) are warp-uniform and hoisted out of the 8 loads of a (q, j) pair (the per-sample
        // form with its predicates and 64-bit address arithmetic made this kernel 5x slower than the inverse transform).
        float2 v[32];
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
            const int fq = frame + q4;
            const bool inside = fq >= 2 && fq <= frames;
#pragma unroll
            for (int a8 = 0; a8 < 8; ++a8) v[8 * q4 + a8] = make_float2(0.f, 0.f);
            if (inside) {
#pragma unroll
                for (int j = 3; j >= 0; --j) {               // ascending frame order, as the reference's overlap-add
                    const int f = fq - j;
                    if (f < 0 || f >= frames) continue;
                    const float2* src = reinterpret_cast<const float2*>(frc + (size_t)f * NFFT + 512 * j) + lane;
#pragma unroll
                    for (int a8 = 0; a8 < 8; ++a8) {
                        const float2 t = src[32 * a8];
                        v[8 * q4 + a8].x += t.x; v[8 * q4 + a8].y += t.y;
                    }
                }
                const float2* wsp = reinterpret_cast<const float2*>(inv_wss + (size_t)fq * HOP) + lane;
#pragma unroll
                for (int a8 = 0; a8 < 8; ++a8) {
                    const float2 w = __ldg(wsp + 32 * a8);
                    v[8 * q4 + a8].x *= w.x; v[8 * q4 + a8].y *= w.y;
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 32; ++a) {
            const int n = 32 * a + lane;
            const float c0 = tw[tw_slot(2 * n)].x, c1 = tw[tw_slot(2 * n + 1)].x;
            v[a] = make_float2(v[a].x * (0.5f - 0.5f * c0), v[a].y * (0.5f - 0.5f * c1));
        }
        fft1024_warp(v, tw, buf, lane);                  // v[r] = Z[lane + 32 brev5(r)]
        const size_t base = ((size_t)clip * frames + frame) * NBINS;
        const int src = (32 - lane) & 31;
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const int k2 = brev5(r);
            const int k = lane + 32 * k2;
            const float2 zk = v[r];
            const float2 own = v[brev5((32 - k2) & 31)];
            float2 zc = make_float2(__shfl_sync(0xffffffffu, v[31 - r].x, src), __shfl_sync(0xffffffffu, v[31 - r].y, src));
            if (lane == 0) zc = own;
            const float ex = 0.5f * (zk.x + zc.x), ey = 0.5f * (zk.y - zc.y);
            const float ox = 0.5f * (zk.y + zc.y), oy = -0.5f * (zk.x - zc.x);
            const float2 t = tw[tw_slot(k)];
            float2 X = make_float2(ex + (ox * t.x - oy * t.y), ey + (ox * t.y + oy * t.x));
            if (k == 0) X = make_float2(zk.x + zk.y, 0.f);            // DC bin: real
            float2 a = X;
            if (!first) {
                const float2 tp = tprev[base + k];
                a.x -= mom * tp.x; a.y -= mom * tp.y;
            }
            const float sc = mag[base + k] / (sqrtf(a.x * a.x + a.y * a.y) + GL_TINY);
            spec[base + k] = make_float2(a.x * sc, a.y * sc);
            tprev[base + k] = X;
        }
        if (lane == 0) {                                               // Nyquist bin: Z[0] with e^{-i pi} = -1, real
            const float X = v[0].x - v[0].y;
            float a = X;
            if (!first) a -= mom * tprev[base + NC].x;
            const float sc = mag[base + NC] / (fabsf(a) + GL_TINY);
            spec[base + NC] = make_float2(a * sc, 0.f);
            tprev[base + NC] = make_float2(X, 0.f);
        }
        __syncwarp();
    }
}

__global__ void gl_ola_kernel(const float* __restrict__ fr, const float* __restrict__ inv_wss, float* __restrict__ audio, int frames, int len) {
    const int clip = blockIdx.y;
    const float* frc = fr + (size_t)clip * frames * NFFT;
    for (int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x); i < len; i += 2 * gridDim.x * blockDim.x) {
        const float2 t = gl_sample_pair(frc, inv_wss, frames, len, i + NFFT / 2);
        *reinterpret_cast<float2*>(audio + (size_t)clip * len + i) = t;
    }
}

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace mstk

using namespace mstk;

extern "C" {

int mstk_mel_to_stft(const mstk_mel_to_stft_args* a, void* stream) {
    MSTK_REQUIRE(a && a->image && a->pinv && a->mel_fb && a->band_ranges && a->mag, "mel_to_stft: null");
    MSTK_REQUIRE(a->b > 0 && a->b <= 65535 && a->n_mels > 0 && a->frames > 0, "mel_to_stft: bad dims");
    MSTK_REQUIRE(a->pg_iters >= 0, "mel_to_stft: pg_iters < 0");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = ((size_t)NBINS * M2S_FR + 2 * (size_t)a->n_mels * M2S_FR) * sizeof(float);
    MSTK_REQUIRE(smem <= 200 * 1024, "mel_to_stft: n_mels %d does not fit shared memory", a->n_mels);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(mel_to_stft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "mel_to_stft smem attr");
        attr_set = true;
    }
    if (a->pg_norm) {
        cudaError_t e = cudaMemsetAsync(a->pg_norm, 0, (size_t)a->b * sizeof(float), st);
        if (e != cudaSuccess) return cuda_fail(e, "mel_to_stft memset");
    }
    mel_to_stft_kernel<<<dim3((a->frames + M2S_FR - 1) / M2S_FR, a->b), M2S_THREADS, smem, st>>>(
        a->image, a->pinv, a->mel_fb, reinterpret_cast<const int2*>(a->band_ranges), a->mag, reinterpret_cast<unsigned int*>(a->pg_norm),
        a->n_mels, a->frames, a->max_db, a->pg_iters, a->pg_step);
    MSTK_CHECK_LAUNCH("mel_to_stft");
    return MSTK_OK;
}

int mstk_db_image_to_mag(const uint8_t* image, float* mag, int b, int bins, int frames, float max_db, void* stream) {
    MSTK_REQUIRE(image && mag && b > 0 && bins > 0 && frames > 0, "db_image_to_mag: bad args");
    const size_t total = (size_t)b * bins * frames;
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    db_image_to_mag_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(image, mag, b, bins, frames, max_db);
    MSTK_CHECK_LAUNCH("db_image_to_mag");
    return MSTK_OK;
}

size_t mstk_griffinlim_workspace_bytes(int b, int frames) {
    if (b <= 0 || frames <= 1) return 0;
    const size_t bf = (size_t)b * frames;
    return 2 * al256(bf * NBINS * sizeof(float2)) + al256(bf * NFFT * sizeof(float)) + al256(((size_t)NFFT + (size_t)HOP * (frames - 1)) * sizeof(float));
}

int mstk_griffinlim(const mstk_griffinlim_args* a, void* stream) {
    MSTK_REQUIRE(a && a->mag && a->phase_u && a->audio && a->work, "griffinlim: null");
    MSTK_REQUIRE(a->b > 0 && a->b <= 65535 && a->frames > 1, "griffinlim: bad dims (b %d, frames %d)", a ? a->b : 0, a ? a->frames : 0);
    MSTK_REQUIRE(a->n_iter >= 0 && a->momentum >= 0.f, "griffinlim: n_iter / momentum out of range");
    MSTK_REQUIRE(a->work_bytes >= mstk_griffinlim_workspace_bytes(a->b, a->frames), "griffinlim: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bf = (size_t)a->b * a->frames;
    uint8_t* w = reinterpret_cast<uint8_t*>(a->work);
    float2* spec = reinterpret_cast<float2*>(w);
    w += al256(bf * NBINS * sizeof(float2));
    float2* tprev = reinterpret_cast<float2*>(w);
    w += al256(bf * NBINS * sizeof(float2));
    float* fr = reinterpret_cast<float*>(w);
    w += al256(bf * NFFT * sizeof(float));
    float* inv_wss = reinterpret_cast<float*>(w);
    const int npad = NFFT + HOP * (a->frames - 1), len = HOP * (a->frames - 1);
    gl_wss_kernel<<<(npad + 255) / 256, 256, 0, st>>>(inv_wss, a->frames, npad);
    MSTK_CHECK_LAUNCH("gl_wss");
    const size_t total = bf * NBINS;
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    gl_init_kernel<<<(unsigned)blocks, 256, 0, st>>>(a->mag, a->phase_u, spec, total);
    MSTK_CHECK_LAUNCH("gl_init");
    const dim3 grid((a->frames + GL_FPC - 1) / GL_FPC, a->b);
    const size_t smem = (size_t)(FFT_TW + GL_WARPS * FFT_BUF) * sizeof(float2);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gl_istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(gl_stft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "griffinlim smem attr");
        attr_set = true;
    }
    const float mom = a->momentum / (1.0f + a->momentum);
    for (int it = 0; it < a->n_iter; ++it) {
        gl_istft_kernel<<<grid, GL_WARPS * 32, smem, st>>>(spec, fr, a->frames);
        MSTK_CHECK_LAUNCH("gl_istft");
        gl_stft_kernel<<<grid, GL_WARPS * 32, smem, st>>>(fr, inv_wss, a->mag, spec, tprev, a->frames, it == 0 ? 1 : 0, mom);
        MSTK_CHECK_LAUNCH("gl_stft");
    }
    gl_istft_kernel<<<grid, GL_WARPS * 32, smem, st>>>(spec, fr, a->frames);
    MSTK_CHECK_LAUNCH("gl_istft");
    int bx = (len / 2 + 255) / 256;
    if (bx > 148 * 4) bx = 148 * 4;
    gl_ola_kernel<<<dim3(bx, a->b), 256, 0, st>>>(fr, inv_wss, a->audio, a->frames, len);
    MSTK_CHECK_LAUNCH("gl_ola");
    return MSTK_OK;
}

}  // extern "C"
