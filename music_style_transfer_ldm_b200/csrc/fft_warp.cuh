// Warp-level FFT primitives shared by the mel front end (mel.cu) and the inverse audio path (griffinlim.cu):
// a 32-point radix-2 DFT in registers and the 1024-point complex transform of one warp built from two of them
// (n = 32a + b, k = k1 + 32 k2; see mel.cu for the derivation), with a shared-memory table e^{-2 pi i m / 2048}.
#pragma once
#include "common.cuh"

namespace mstk {

constexpr int NFFT = 2048, HOP = 512, NBINS = NFFT / 2 + 1, NC = NFFT / 2;
constexpr int FFT_TW = NFFT + NFFT / 32;          // twiddle table entries, one pad slot per 32
constexpr int FFT_BUF = 32 * 33;                  // per-warp transposition tile (float2)

__device__ __forceinline__ float2 cmul(const float2 a, const float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__host__ __device__ constexpr int brev5(int r) { return ((r & 1) << 4) | ((r & 2) << 2) | (r & 4) | ((r & 8) >> 2) | ((r & 16) >> 4); }

// e^{-2 pi i j / 32}, j = 0..15, folded to literals when j is a compile-time constant (every call site is unrolled)
__device__ __forceinline__ float2 w32(const int j) {
    switch (j) {
        case 0: return make_float2(1.f, 0.f);
        case 1: return make_float2(0.98078528040323043f, -0.19509032201612825f);
        case 2: return make_float2(0.92387953251128674f, -0.38268343236508978f);
        case 3: return make_float2(0.83146961230254524f, -0.55557023301960218f);
        case 4: return make_float2(0.70710678118654757f, -0.70710678118654757f);
        case 5: return make_float2(0.55557023301960229f, -0.83146961230254524f);
        case 6: return make_float2(0.38268343236508984f, -0.92387953251128674f);
        case 7: return make_float2(0.19509032201612833f, -0.98078528040323043f);
        case 8: return make_float2(0.f, -1.f);
        case 9: return make_float2(-0.19509032201612819f, -0.98078528040323043f);
        case 10: return make_float2(-0.38268343236508973f, -0.92387953251128674f);
        case 11: return make_float2(-0.55557023301960196f, -0.83146961230254535f);
        case 12: return make_float2(-0.70710678118654746f, -0.70710678118654757f);
        case 13: return make_float2(-0.83146961230254535f, -0.55557023301960218f);
        case 14: return make_float2(-0.92387953251128674f, -0.38268343236508989f);
        default: return make_float2(-0.98078528040323043f, -0.19509032201612861f);
    }
}

// 32-point DFT in registers, radix-2 decimation in frequency: on return v[r] holds X[brev5(r)]
__device__ __forceinline__ void fft32(float2 (&v)[32]) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
#pragma unroll
        for (int g = 0; g < 32; g += 2 * s) {
#pragma unroll
            for (int j = 0; j < s; ++j) {
                const float2 a = v[g + j], b = v[g + j + s];
                v[g + j] = make_float2(a.x + b.x, a.y + b.y);
                const float2 d = make_float2(a.x - b.x, a.y - b.y);
                const int tw = j * (16 / s);
                if (tw == 0) v[g + j + s] = d;
                else if (tw == 8) v[g + j + s] = make_float2(d.y, -d.x);          // * (-i)
                else v[g + j + s] = cmul(d, w32(tw));
            }
        }
    }
}

__device__ __forceinline__ int tw_slot(const int m) { return m + (m >> 5); }

// table e^{-2 pi i m / 2048}, m < 2048, filled by the whole CTA (caller synchronises)
__device__ __forceinline__ void fft_fill_twiddles(float2* tw, const int tid, const int nthreads) {
    for (int m = tid; m < NFFT; m += nthreads) {
        float sn, cs;
        sincospif(-(float)m * (2.0f / (float)NFFT), &sn, &cs);
        tw[tw_slot(m)] = make_float2(cs, sn);
    }
}

// Forward 1024-point complex DFT of one warp.  In: lane = b, v[a] = z[32a + b].  Out: v[r] = Z[lane + 32 brev5(r)].
// buf: this warp's FFT_BUF tile.  Ends with the tile free for reuse after a __syncwarp().
__device__ __forceinline__ void fft1024_warp(float2 (&v)[32], const float2* __restrict__ tw, float2* __restrict__ buf, const int lane) {
    fft32(v);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 32; ++r) {
        const int k1 = brev5(r);
        const float2 y = (k1 == 0) ? v[r] : cmul(v[r], tw[tw_slot(2 * lane * k1)]);       // W_1024^(b k1)
        buf[k1 * 33 + lane] = y;
    }
    __syncwarp();
#pragma unroll
    for (int b = 0; b < 32; ++b) v[b] = buf[lane * 33 + b];
    __syncwarp();
    fft32(v);
}

}  // namespace mstk
