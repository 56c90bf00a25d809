// Mel front end: wave -> framed Hann rFFT-2048 power -> Slaney mel -> dB re clip max ->
// uint8-quantised [0,1] image.  Replaces librosa.feature.melspectrogram + power_to_db +
// the image scaling of data/audio_processor.py:44-73; the filterbank matrix
// is passed in by the host (it is a constant of the configuration, not hot-path work).
//
// Kernel 1: one CTA per (frame, clip).  The 2048 real samples are packed as a 1024-point
// complex sequence, transformed with an in-shared-memory radix-2 FFT (256 threads), split
// back into the 1025-bin real spectrum, and reduced against each mel band's non-zero bin
// range by one warp per band.  The clip maximum is kept with an integer atomicMax on the
// (non-negative) float bit pattern.
// Kernel 2: dB, floor at -80 dB, quantise, crop; 1 thread per output pixel.
#include "common.cuh"

namespace mstk {

constexpr int NFFT = 2048, HOP = 512, NBINS = NFFT / 2 + 1, NC = NFFT / 2;

__global__ void mel_ranges_kernel(const float* __restrict__ fb, int n_mels, int2* ranges) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_mels) return;
    int lo = NBINS, hi = 0;
    for (int k = 0; k < NBINS; ++k)
        if (fb[(size_t)m * NBINS + k] != 0.f) { if (k < lo) lo = k; hi = k + 1; }
    if (lo > hi) lo = hi = 0;
    ranges[m] = make_int2(lo, hi);
}

__global__ void __launch_bounds__(256) mel_power_kernel(const float* __restrict__ wave, int len, int frames, const float* __restrict__ fb,
                                                       const int2* __restrict__ ranges, int n_mels, float* __restrict__ melp,
                                                       unsigned int* __restrict__ clipmax) {
    __shared__ float2 z[NC];
    __shared__ float pw[NBINS + 3];
    const int frame = blockIdx.x, clip = blockIdx.y, tid = threadIdx.x;
    const float* wv = wave + (size_t)clip * len;
    const int start = frame * HOP - NFFT / 2;
    // load, window (periodic Hann: 0.5 - 0.5 cos(2 pi n / N)), pack, bit-reverse
    for (int i = tid; i < NC; i += 256) {
        const int n0 = 2 * i, n1 = 2 * i + 1;
        const int s0 = start + n0, s1 = start + n1;
        const float x0 = (s0 >= 0 && s0 < len) ? wv[s0] : 0.f;
        const float x1 = (s1 >= 0 && s1 < len) ? wv[s1] : 0.f;
        const float w0 = 0.5f - 0.5f * cospif(2.0f * (float)n0 / (float)NFFT);
        const float w1 = 0.5f - 0.5f * cospif(2.0f * (float)n1 / (float)NFFT);
        const int r = (int)(__brev((unsigned)i) >> 22);   // 10-bit reversal
        z[r] = make_float2(x0 * w0, x1 * w1);
    }
    __syncthreads();
    // 10 radix-2 DIT stages, 512 butterflies each
    for (int s = 1; s <= 10; ++s) {
        const int half = 1 << (s - 1);
        for (int j = tid; j < NC / 2; j += 256) {
            const int grp = j / half, pos = j - grp * half;
            const int i0 = grp * 2 * half + pos, i1 = i0 + half;
            float sn, cs;
            sincospif(-(float)pos / (float)half, &sn, &cs);   // e^{-i pi pos/half}
            const float2 a = z[i0], b = z[i1];
            const float2 t = make_float2(b.x * cs - b.y * sn, b.x * sn + b.y * cs);
            z[i0] = make_float2(a.x + t.x, a.y + t.y);
            z[i1] = make_float2(a.x - t.x, a.y - t.y);
        }
        __syncthreads();
    }
    // split: X[k] = (Z[k] + conj(Z[NC-k]))/2 - (i/2) e^{-2 pi i k/NFFT} (Z[k] - conj(Z[NC-k]))
    for (int k = tid; k <= NC; k += 256) {
        const float2 zk = z[k & (NC - 1)];
        const float2 zc = z[(NC - k) & (NC - 1)];
        const float ex = 0.5f * (zk.x + zc.x), ey = 0.5f * (zk.y - zc.y);     // even part
        const float ox = 0.5f * (zk.y + zc.y), oy = -0.5f * (zk.x - zc.x);    // odd part = -i/2 (Z[k] - conj(Z[NC-k]))
        float sn, cs;
        sincospif(-2.0f * (float)k / (float)NFFT, &sn, &cs);
        const float re = ex + (ox * cs - oy * sn);
        const float im = ey + (ox * sn + oy * cs);
        pw[k] = re * re + im * im;
    }
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    for (int m = warp; m < n_mels; m += 8) {
        const int2 rg = ranges[m];
        float acc = 0.f;
        for (int k = rg.x + lane; k < rg.y; k += 32) acc = fmaf(fb[(size_t)m * NBINS + k], pw[k], acc);
        acc = warp_sum(acc);
        if (lane == 0) {
            melp[((size_t)clip * n_mels + m) * frames + frame] = acc;
            atomicMax(clipmax + clip, __float_as_uint(fmaxf(acc, 0.f)));
        }
    }
}

__global__ void mel_db_kernel(const float* __restrict__ melp, const unsigned int* __restrict__ clipmax, int b, int n_mels, int frames,
                              int crop, float* __restrict__ out, float* __restrict__ db_out) {
    const size_t total = (size_t)b * n_mels * frames;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int f = (int)(i % frames);
        const size_t r = i / frames;
        const int clip = (int)(r / n_mels);
        const float ref = fmaxf(__uint_as_float(clipmax[clip]), 1e-10f);
        // power_to_db: 10 log10(max(amin, S)) - 10 log10(max(amin, ref)), then max(., max - 80); the max is 0 dB
        float db = 10.f * log10f(fmaxf(melp[i], 1e-10f)) - 10.f * log10f(ref);
        db = fmaxf(db, -80.f);
        if (db_out) db_out[i] = db;
        if (f < crop) {
            float v = (db + 80.f) * (255.f / 80.f);
            v = fminf(fmaxf(v, 0.f), 255.f);
            const float q = truncf(v + 0.5f);
            out[r * (size_t)crop + f] = q / 255.f;
        }
    }
}

}  // namespace mstk

using namespace mstk;

extern "C" {

size_t mstk_mel_workspace_bytes(int b, int len, int n_mels) {
    if (b <= 0 || len <= 0 || n_mels <= 0) return 0;
    const size_t frames = 1 + (size_t)len / HOP;
    size_t bytes = (size_t)b * n_mels * frames * sizeof(float);     // mel power
    bytes = (bytes + 255) & ~(size_t)255;
    bytes += ((size_t)b * sizeof(unsigned int) + 255) & ~(size_t)255;   // clip max
    bytes += ((size_t)n_mels * sizeof(int2) + 255) & ~(size_t)255;      // band ranges
    return bytes;
}

int mstk_mel_frontend(const mstk_mel_args* a, void* stream) {
    MSTK_REQUIRE(a && a->wave && a->out && a->mel_fb && a->work, "mel: null");
    MSTK_REQUIRE(a->b > 0 && a->b <= 65535 && a->len > 0 && a->n_mels > 0, "mel: bad dims");
    const int frames = 1 + a->len / HOP;
    MSTK_REQUIRE(a->crop_frames > 0 && a->crop_frames <= frames, "mel: crop_frames %d out of range (frames %d)", a->crop_frames, frames);
    MSTK_REQUIRE(a->work_bytes >= mstk_mel_workspace_bytes(a->b, a->len, a->n_mels), "mel: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* w = reinterpret_cast<uint8_t*>(a->work);
    float* melp = reinterpret_cast<float*>(w);
    size_t off = (((size_t)a->b * a->n_mels * frames * sizeof(float)) + 255) & ~(size_t)255;
    unsigned int* clipmax = reinterpret_cast<unsigned int*>(w + off);
    off += ((size_t)a->b * sizeof(unsigned int) + 255) & ~(size_t)255;
    int2* ranges = reinterpret_cast<int2*>(w + off);
    cudaError_t e = cudaMemsetAsync(clipmax, 0, (size_t)a->b * sizeof(unsigned int), st);
    if (e != cudaSuccess) return cuda_fail(e, "mel memset");
    mel_ranges_kernel<<<(a->n_mels + 127) / 128, 128, 0, st>>>(a->mel_fb, a->n_mels, ranges);
    MSTK_CHECK_LAUNCH("mel_ranges");
    mel_power_kernel<<<dim3(frames, a->b), 256, 0, st>>>(a->wave, a->len, frames, a->mel_fb, ranges, a->n_mels, melp, clipmax);
    MSTK_CHECK_LAUNCH("mel_power");
    const size_t total = (size_t)a->b * a->n_mels * frames;
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    mel_db_kernel<<<(unsigned)blocks, 256, 0, st>>>(melp, clipmax, a->b, a->n_mels, frames, a->crop_frames, a->out, a->db_out);
    MSTK_CHECK_LAUNCH("mel_db");
    return MSTK_OK;
}

}  // extern "C"
