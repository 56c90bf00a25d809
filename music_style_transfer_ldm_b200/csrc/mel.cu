// Mel front end: wave -> framed Hann rFFT-2048 power -> Slaney mel -> dB re clip max ->
// uint8-quantised [0,1] image.  Replaces librosa.feature.melspectrogram + power_to_db +
// the image scaling of data/audio_processor.py:44-73; the filterbank matrix
// is passed in by the host (it is a constant of the configuration, not hot-path work).
//
// Kernel 1 (mel_power_kernel): ONE WARP PER FRAME, the FFT in registers.  The 2048 windowed real samples are packed
// as a 1024-point complex sequence z and transformed as 32 x 32 (n = 32a + b, k = k1 + 32 k2):
//     Y[b][k1] = W_1024^(b k1) * DFT32_a(z[32a + b])      lane = b, 32 points in registers, fully unrolled radix-2 DIF
//     Z[k1 + 32 k2] = DFT32_b(Y[b][k1])                   lane = k1, after one transposition through shared memory
// with compile-time 32-point twiddles and a shared table e^{-2 pi i m / 2048} (m < 2048) for the inter-pass
// twiddles, the Hann window (0.5 - 0.5 Re) and the real-spectrum split.  The partner Z[1024 - k] of the split lives
// in lane 32 - k1 at a statically known register: one shuffle pair per bin, no second pass through memory.  Power
// spectrum -> shared memory -> lane m of the warp reduces mel band m, m + 32, ... over its non-zero bin range
// (ranges computed once per filterbank, mstk_mel_band_ranges).  A CTA = 8 warps x 2 frames = 16 consecutive frames
// of one clip: the mel powers are staged as [band][16 frames] and written as 64-byte runs (the previous kernel wrote
// one 4-byte value per 32-byte sector, recomputed sincospi in every butterfly and ran 10 block-wide barriers per
// frame).  The clip maximum is kept with an integer atomicMax on the (non-negative) float bit pattern.
// Kernel 2 (mel_db_kernel): dB, floor at -80 dB, quantise, crop; 1 thread per output pixel, frame-contiguous.
#include "common.cuh"
#include "fft_warp.cuh"

#include <mutex>

namespace mstk {

constexpr int MEL_WARPS = 8, MEL_FPW = 2, MEL_FPC = MEL_WARPS * MEL_FPW;     // frames per CTA
constexpr int MEL_TW = FFT_TW;                                                // twiddle table, one pad slot per 32 entries
constexpr int MEL_BUF = 32 * 33;                                              // per warp: transposition buffer (float2), then reused
                                                                              // for the power spectrum (1025 floats)

__global__ void mel_ranges_kernel(const float* __restrict__ fb, int n_mels, int2* ranges) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_mels) return;
    int lo = NBINS, hi = 0;
    for (int k = 0; k < NBINS; ++k)
        if (fb[(size_t)m * NBINS + k] != 0.f) { if (k < lo) lo = k; hi = k + 1; }
    if (lo > hi) lo = hi = 0;
    ranges[m] = make_int2(lo, hi);
}

__global__ void __launch_bounds__(MEL_WARPS * 32, 2) mel_power_kernel(const float* __restrict__ wave, int len, int frames,
                                                                  const float* __restrict__ fb, const int2* __restrict__ ranges, int n_mels,
                                                                  float* __restrict__ melp, unsigned int* __restrict__ clipmax) {
    extern __shared__ __align__(16) uint8_t mel_smem[];
    float2* tw = reinterpret_cast<float2*>(mel_smem);                         // [MEL_TW] e^{-2 pi i m / 2048}
    float2* bufs = tw + MEL_TW;                                               // [MEL_WARPS][MEL_BUF]
    float* outs = reinterpret_cast<float*>(bufs + MEL_WARPS * MEL_BUF);       // [n_mels][MEL_FPC]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int clip = blockIdx.y, f0 = blockIdx.x * MEL_FPC;
    for (int m = tid; m < NFFT; m += MEL_WARPS * 32) {
        float sn, cs;
        sincospif(-(float)m * (2.0f / (float)NFFT), &sn, &cs);
        tw[tw_slot(m)] = make_float2(cs, sn);
    }
    __syncthreads();
    const float* wv = wave + (size_t)clip * len;
    const bool vec_ok = ((len & 1) == 0) && ((reinterpret_cast<uintptr_t>(wave) & 7) == 0);
    float2* buf = bufs + warp * MEL_BUF;
    float* pw = reinterpret_cast<float*>(buf);                                // the power spectrum overwrites the transposed tile
    float wmax = 0.f;
#pragma unroll 1
    for (int fi = 0; fi < MEL_FPW; ++fi) {
        const int fl = warp * MEL_FPW + fi, frame = f0 + fl;
        if (frame >= frames) break;                                           // warp-uniform
        const int start = frame * HOP - NFFT / 2;                             // even
        float2 v[32];
        // ---- pass 1: lane = b, v[a] = window * z[32a + b] ----
        if (vec_ok && start >= 0 && start + NFFT <= len) {                     // interior frame (warp-uniform): 8-byte loads, no predicates
#pragma unroll
            for (int a = 0; a < 32; ++a) {
                const float2 t = __ldg(reinterpret_cast<const float2*>(wv + start) + 32 * a + lane);
                v[a] = t;
            }
        } else {                                                               // the two frames at either end of a clip: zero padding
#pragma unroll
            for (int a = 0; a < 32; ++a) {
                const int s0 = start + 2 * (32 * a + lane);
                float x0 = 0.f, x1 = 0.f;
                if (s0 >= 0 && s0 < len) x0 = __ldg(wv + s0);
                if (s0 + 1 >= 0 && s0 + 1 < len) x1 = __ldg(wv + s0 + 1);
                v[a] = make_float2(x0, x1);
            }
        }
#pragma unroll
        for (int a = 0; a < 32; ++a) {
            // periodic Hann 0.5 - 0.5 cos(2 pi m / 2048), m = 2n and 2n + 1, n = 32a + lane
            const int n = 32 * a + lane;
            const float c0 = tw[tw_slot(2 * n)].x, c1 = tw[tw_slot(2 * n + 1)].x;
            v[a] = make_float2(v[a].x * (0.5f - 0.5f * c0), v[a].y * (0.5f - 0.5f * c1));
        }
        fft32(v);
        __syncwarp();                                                          // previous frame's reads of buf are done
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const int k1 = brev5(r);
            const float2 y = (k1 == 0) ? v[r] : cmul(v[r], tw[tw_slot(2 * lane * k1)]);   // W_1024^(b k1)
            buf[k1 * 33 + lane] = y;
        }
        __syncwarp();
        // ---- pass 2: lane = k1 ----
#pragma unroll
        for (int b = 0; b < 32; ++b) v[b] = buf[lane * 33 + b];
        __syncwarp();                                                          // the tile is dead: pw may overwrite it
        fft32(v);                                                              // v[r] = Z[lane + 32 brev5(r)]
        // ---- real-spectrum split: X[k] = E[k] + e^{-2 pi i k / 2048} O[k], partner Z[1024 - k] by shuffle ----
        const int src = (32 - lane) & 31;
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const int k2 = brev5(r);
            const int k = lane + 32 * k2;
            const float2 zk = v[r];
            const float2 own = v[brev5((32 - k2) & 31)];                       // lane 0: Z[32 (32 - k2) mod 1024]
            float2 zc = make_float2(__shfl_sync(0xffffffffu, v[31 - r].x, src), __shfl_sync(0xffffffffu, v[31 - r].y, src));
            if (lane == 0) zc = own;
            const float ex = 0.5f * (zk.x + zc.x), ey = 0.5f * (zk.y - zc.y);
            const float ox = 0.5f * (zk.y + zc.y), oy = -0.5f * (zk.x - zc.x);
            const float2 t = tw[tw_slot(k)];
            const float re = ex + (ox * t.x - oy * t.y);
            const float im = ey + (ox * t.y + oy * t.x);
            pw[k] = re * re + im * im;
        }
        if (lane == 0) { const float d = v[0].x - v[0].y; pw[NC] = d * d; }    // Nyquist bin: Z[0] with e^{-i pi} = -1
        __syncwarp();
        // ---- mel bands: lane m, m + 32, ... ----
        for (int m = lane; m < n_mels; m += 32) {
            const int2 rg = ranges[m];
            const float* fr = fb + (size_t)m * NBINS;
            float acc = 0.f;
            for (int k = rg.x; k < rg.y; ++k) acc = fmaf(__ldg(fr + k), pw[k], acc);
            outs[m * MEL_FPC + fl] = acc;
            wmax = fmaxf(wmax, acc);
        }
        __syncwarp();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = fmaxf(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0 && wmax > 0.f) atomicMax(clipmax + clip, __float_as_uint(wmax));
    __syncthreads();
    const int nf = min(MEL_FPC, frames - f0);
    for (int i = tid; i < n_mels * MEL_FPC; i += MEL_WARPS * 32) {
        const int m = i / MEL_FPC, j = i - m * MEL_FPC;
        if (j < nf) melp[((size_t)clip * n_mels + m) * frames + f0 + j] = outs[i];
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
    bytes += ((size_t)n_mels * sizeof(int2) + 255) & ~(size_t)255;      // band ranges (when the caller passes none)
    return bytes;
}

int mstk_mel_band_ranges(const float* mel_fb, int n_mels, int32_t* ranges, void* stream) {
    MSTK_REQUIRE(mel_fb && ranges && n_mels > 0, "mel_band_ranges: bad args");
    mel_ranges_kernel<<<(n_mels + 127) / 128, 128, 0, (cudaStream_t)stream>>>(mel_fb, n_mels, reinterpret_cast<int2*>(ranges));
    MSTK_CHECK_LAUNCH("mel_ranges");
    return MSTK_OK;
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
    const int2* ranges = reinterpret_cast<const int2*>(a->band_ranges);
    cudaError_t e = cudaMemsetAsync(clipmax, 0, (size_t)a->b * sizeof(unsigned int), st);
    if (e != cudaSuccess) return cuda_fail(e, "mel memset");
    if (!ranges) {      // one-off callers: derive the non-zero bin range of every band here (mstk_mel_band_ranges caches it)
        int2* r = reinterpret_cast<int2*>(w + off);
        mel_ranges_kernel<<<(a->n_mels + 127) / 128, 128, 0, st>>>(a->mel_fb, a->n_mels, r);
        MSTK_CHECK_LAUNCH("mel_ranges");
        ranges = r;
    }
    const size_t smem = (size_t)MEL_TW * sizeof(float2) + (size_t)MEL_WARPS * MEL_BUF * sizeof(float2) +
                        (size_t)a->n_mels * MEL_FPC * sizeof(float);
    MSTK_REQUIRE(smem <= 200 * 1024, "mel: n_mels %d does not fit the staging tile", a->n_mels);
    static size_t smem_set = 0;
    if (smem > smem_set) {
        e = cudaFuncSetAttribute(mel_power_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "mel smem attr");
        smem_set = 200 * 1024;
    }
    mel_power_kernel<<<dim3((frames + MEL_FPC - 1) / MEL_FPC, a->b), MEL_WARPS * 32, smem, st>>>(a->wave, a->len, frames, a->mel_fb, ranges,
                                                                                             a->n_mels, melp, clipmax);
    MSTK_CHECK_LAUNCH("mel_power");
    const size_t total = (size_t)a->b * a->n_mels * frames;
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    mel_db_kernel<<<(unsigned)blocks, 256, 0, st>>>(melp, clipmax, a->b, a->n_mels, frames, a->crop_frames, a->out, a->db_out);
    MSTK_CHECK_LAUNCH("mel_db");
    return MSTK_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA-graph capture, tight serving loops, harnesses that forbid it).  The front end's
// scratch (mel power, per-clip maximum, band ranges) lives in ONE file-scope device array instead of a caller buffer.
// Every GPU that loads the library holds MEL_STATIC_BYTES until exit.  In host code `g_mel_scratch` is only a host-side
// stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------------------------------------------
namespace mstk {
constexpr size_t MEL_STATIC_BYTES = (size_t)8 << 20;      // 8 MiB: e.g. 64 clips x 128 bands x 130 frames (3 s at 22.05 kHz) = 4.1 MiB
static __device__ __align__(256) unsigned char g_mel_scratch[MEL_STATIC_BYTES];

constexpr int MEL_MAX_DEVICES = 64;
static void* g_mel_scratch_ptr[MEL_MAX_DEVICES];            // device address of g_mel_scratch on device i (0 = not looked up yet)
static std::mutex g_mel_scratch_mu;

static int mel_static_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "mel_static: cudaGetDevice");
    MSTK_REQUIRE(dev >= 0 && dev < MEL_MAX_DEVICES, "mel_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_mel_scratch_mu);
    if (!g_mel_scratch_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_mel_scratch);       // the current device's copy of the array
        if (e != cudaSuccess) return cuda_fail(e, "mel_static: cudaGetSymbolAddress");
        MSTK_REQUIRE(p != nullptr, "mel_static: cudaGetSymbolAddress returned a null address");
        g_mel_scratch_ptr[dev] = p;
    }
    *out = g_mel_scratch_ptr[dev];
    return MSTK_OK;
}
}  // namespace mstk

extern "C" {

size_t mstk_mel_static_scratch_bytes(void) { return MEL_STATIC_BYTES; }

int mstk_mel_frontend_static(const mstk_mel_args* a, void* stream) {
    MSTK_REQUIRE(a != nullptr, "mel_static: null");
    const size_t need = mstk_mel_workspace_bytes(a->b, a->len, a->n_mels);
    MSTK_REQUIRE(need > 0, "mel_static: bad dims");
    MSTK_REQUIRE(need <= MEL_STATIC_BYTES,
                 "mel_static: b %d x n_mels %d x frames %d needs %zu bytes of scratch, the built-in array holds %zu; "
                 "use mstk_mel_frontend with a caller workspace", a->b, a->n_mels, 1 + a->len / HOP, need, MEL_STATIC_BYTES);
    void* scratch = nullptr;
    const int rc = mel_static_scratch(&scratch);
    if (rc != MSTK_OK) return rc;
    mstk_mel_args b = *a;              // a->work / a->work_bytes are ignored
    b.work = scratch;
    b.work_bytes = MEL_STATIC_BYTES;
    // same launches as the caller-workspace entry point: it clears the per-clip maxima (the only scratch that must start at
    // zero) on `stream` and overwrites everything else it reads, so the previous call's contents never leak into this one
    return mstk_mel_frontend(&b, stream);
}

}  // extern "C"
