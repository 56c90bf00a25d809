// HBM-streaming kernels of the VGGish feature loss (reference models/loss.py:64-101):
//
//   sample_stats_kernel   per-sample sum / sum of squares of a feature map [S, inner]
//   feat_mse3_kernel      ONE pass over up to three feature maps f0, f1, f2 [B, inner] that
//                         accumulates  A = sum (f0/(std0+1e-8) - f1/(std1+1e-8))^2   (perceptual pair)
//                                      C = sum (f1/(std1+1e-8) - f2/(std2+1e-8))^2   (style pair)
//                         the training step evaluates both terms on (x, recon) and (recon, style), so
//                         recon's features are read once instead of twice.
//   maxpool2x2_vec_kernel 2x2/2 max pool on NHWC, 8 channels per thread
//
// All of them move 16 bytes per thread and load (8 bf16 / 4 fp32), keep fp32 partial sums over short
// runs and fold them into doubles, use 32-bit index arithmetic inside a sample, and end in one double
// atomic per CTA and quantity.
#include "common.cuh"
#include "mma_sync.cuh"
#include <stdlib.h>

namespace mstk {

template <int DT> struct VecT;
template <> struct VecT<MSTK_BF16> { static constexpr int N = 8; };
template <> struct VecT<MSTK_F32> { static constexpr int N = 4; };

template <int DT>
__device__ __forceinline__ void load_vec(const void* p, size_t vec_idx, float (&o)[VecT<DT>::N]) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(p) + vec_idx);
    if (DT == MSTK_BF16) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 f = __bfloat1622float2(h[e]);
            o[2 * e] = f.x; o[2 * e + 1] = f.y;
        }
    } else {
        o[0] = __uint_as_float(q.x); o[1] = __uint_as_float(q.y); o[2] = __uint_as_float(q.z); o[3] = __uint_as_float(q.w);
    }
}

__device__ __forceinline__ double block_sum_d(double v, double* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double r = 0.0;
    if (warp == 0) {
        r = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.0;
        r = warp_sum(r);
    }
    return r;   // valid in thread 0
}

// stats[s*2 + {0,1}] += sum, sum of squares of f[s, :]   (grid: x = chunks of a sample, y = sample)
template <int DT>
__global__ void __launch_bounds__(256) sample_stats_kernel(const void* __restrict__ f, long long inner, double* stats) {
    __shared__ double red[32];
    constexpr int V = VecT<DT>::N;
    const int s = blockIdx.y;
    const size_t vbase = (size_t)s * (size_t)(inner / V);
    const unsigned nvec = (unsigned)(inner / V);
    double ds = 0.0, dq = 0.0;
    float ls = 0.f, lq = 0.f;
    int run = 0;
    for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < nvec; i += gridDim.x * 256u) {
        float v[V];
        load_vec<DT>(f, vbase + i, v);
#pragma unroll
        for (int j = 0; j < V; ++j) { ls += v[j]; lq = fmaf(v[j], v[j], lq); }
        if (++run == 16) { ds += ls; dq += lq; ls = lq = 0.f; run = 0; }
    }
    ds += ls; dq += lq;
    ds = block_sum_d(ds, red);
    dq = block_sum_d(dq, red);
    if (threadIdx.x == 0) {
        atomicAdd(stats + 2 * s, ds);
        atomicAdd(stats + 2 * s + 1, dq);
    }
}

__device__ __forceinline__ float inv_std(const double* st, int b, double n) {
    const double m = st[2 * b] / n;
    double var = (st[2 * b + 1] - n * m * m) / (n - 1.0);      // unbiased, as torch.std (loss.py:94)
    if (var < 0) var = 0;
    return 1.f / ((float)sqrt(var) + 1e-8f);
}

// scratch: [acc_a, acc_c, counter]; the last CTA finalises, adds the weighted means into out_a / out_c and
// re-zeroes scratch and the three statistics arrays.
template <int DT>
__global__ void __launch_bounds__(256) feat_mse3_kernel(const void* __restrict__ f0, const void* __restrict__ f1, const void* __restrict__ f2,
                                                       int nb, long long inner, double* st0, double* st1, double* st2, double* scratch,
                                                       float weight, float* out_a, float* out_c) {
    __shared__ double red[32];
    constexpr int V = VecT<DT>::N;
    const int b = blockIdx.y;
    const double n = (double)inner;
    const float r0 = inv_std(st0, b, n), r1 = inv_std(st1, b, n);
    const float r2 = f2 ? inv_std(st2, b, n) : 0.f;
    const size_t vbase = (size_t)b * (size_t)(inner / V);
    const unsigned nvec = (unsigned)(inner / V);
    double da = 0.0, dc = 0.0;
    float la = 0.f, lc = 0.f;
    int run = 0;
    for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < nvec; i += gridDim.x * 256u) {
        float a[V], m[V];
        load_vec<DT>(f0, vbase + i, a);
        load_vec<DT>(f1, vbase + i, m);
        if (f2) {
            float c[V];
            load_vec<DT>(f2, vbase + i, c);
#pragma unroll
            for (int j = 0; j < V; ++j) { const float d = m[j] * r1 - c[j] * r2; lc = fmaf(d, d, lc); }
        }
#pragma unroll
        for (int j = 0; j < V; ++j) { const float d = a[j] * r0 - m[j] * r1; la = fmaf(d, d, la); }
        if (++run == 16) { da += la; dc += lc; la = lc = 0.f; run = 0; }
    }
    da += la; dc += lc;
    da = block_sum_d(da, red);
    dc = block_sum_d(dc, red);
    if (threadIdx.x == 0) {
        atomicAdd(scratch, da);
        if (f2) atomicAdd(scratch + 1, dc);
        __threadfence();
        unsigned int* ctr = reinterpret_cast<unsigned int*>(scratch + 2);
        const unsigned int old = atomicAdd(ctr, 1u);
        if (old == gridDim.x * gridDim.y - 1) {
            __threadfence();
            const double sa = atomicAdd(scratch, 0.0), sc = atomicAdd(scratch + 1, 0.0);
            const double den = (double)nb * n;
            atomicAdd(out_a, weight * (float)(sa / den));
            if (f2) atomicAdd(out_c, weight * (float)(sc / den));
            scratch[0] = 0.0; scratch[1] = 0.0;
            *ctr = 0u;
            for (int i = 0; i < 2 * nb; ++i) {
                st0[i] = 0.0; st1[i] = 0.0;
                if (f2) st2[i] = 0.0;
            }
        }
    }
}

// x [n,h,w,c] -> y [n,h/2,w/2,c]; one thread = 8 (bf16) or 4 (fp32) channels of one output pixel
template <int DT>
__global__ void __launch_bounds__(256) maxpool2x2_vec_kernel(const void* __restrict__ x, void* y, int n, int h, int w, int c) {
    constexpr int V = VecT<DT>::N;
    const int ho = h / 2, wo = w / 2, cv = c / V;
    const unsigned per_img = (unsigned)ho * wo * cv;
    const int img = blockIdx.y;
    const size_t in_base = (size_t)img * h * w * cv;      // in vectors
    const size_t out_base = (size_t)img * per_img;
    for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < per_img; i += gridDim.x * 256u) {
        const unsigned ch = i % cv;
        const unsigned r = i / cv;
        const unsigned ox = r % wo, oy = r / wo;
        const size_t b00 = in_base + ((size_t)(2 * oy) * w + 2 * ox) * cv + ch;
        float a[V], b[V], d[V], e[V];
        load_vec<DT>(x, b00, a);
        load_vec<DT>(x, b00 + cv, b);
        load_vec<DT>(x, b00 + (size_t)w * cv, d);
        load_vec<DT>(x, b00 + (size_t)w * cv + cv, e);
        uint4 o;
        if (DT == MSTK_BF16) {
            __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                hh[j] = __floats2bfloat162_rn(fmaxf(fmaxf(a[2 * j], b[2 * j]), fmaxf(d[2 * j], e[2 * j])),
                                              fmaxf(fmaxf(a[2 * j + 1], b[2 * j + 1]), fmaxf(d[2 * j + 1], e[2 * j + 1])));
        } else {
            o.x = __float_as_uint(fmaxf(fmaxf(a[0], b[0]), fmaxf(d[0], e[0])));
            o.y = __float_as_uint(fmaxf(fmaxf(a[1], b[1]), fmaxf(d[1], e[1])));
            o.z = __float_as_uint(fmaxf(fmaxf(a[2 % V], b[2 % V]), fmaxf(d[2 % V], e[2 % V])));
            o.w = __float_as_uint(fmaxf(fmaxf(a[3 % V], b[3 % V]), fmaxf(d[3 % V], e[3 % V])));
        }
        reinterpret_cast<uint4*>(y)[out_base + i] = o;
    }
}

static int fl_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) sms = v;
        else sms = 148;
    }
    return sms;
}
// grid.x so that grid.x * samples ~ 8 CTAs per SM, at most one CTA per 256 vectors
static unsigned fl_grid_x(long long nvec, int samples) {
    long long gx = (nvec + 255) / 256;
    long long cap = ((long long)fl_sms() * 8 + samples - 1) / samples;
    if (cap < 1) cap = 1;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    return (unsigned)gx;
}

// ---------------------------------------------------------------------------
// The whole VGGish tap of the training step in ONE pass over the feature map of the three image batches
// x (0), recon (1), style (2), stacked as f[3*nb, h, w, c]:
//   per sample b: sums  Sx, Sxx, Sr, Srr, Ss, Sss  and the cross terms  Sxr, Srs   -> acc[b][8] (double)
//   optionally the 2x2/2 max-pooled map (the next conv's input) for all three batches.
// With r_i = 1/(std_i + 1e-8) the tap terms of loss.py:94-99 follow algebraically, with no second pass:
//   sum (x r_x - r r_r)^2 = r_x^2 Sxx - 2 r_x r_r Sxr + r_r^2 Srr      (likewise for the (recon, style) pair)
// The last CTA to finish evaluates them for every sample, adds weight * mean into out_a / out_c and re-zeroes
// acc and the counter.  Traffic: one read of f (+ a quarter-size write) instead of three reads (+ the write).
// ---------------------------------------------------------------------------
// Common end of the one-pass tap kernels: block-reduce the eight per-sample sums (l: fp32 run, d: double), add them to
// acc[b], and let the last CTA of the grid turn the sums into the two loss terms (see feat_fused3_kernel).
__device__ __forceinline__ void fused3_finish(float (&l)[8], double (&d)[8], double* red, const int b, const int nb, const int h, const int w,
                                              const int c, double* acc, double* scratch, const float weight, float* out_a, float* out_c) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        d[k] += l[k];
        d[k] = block_sum_d(d[k], red);
    }
    __shared__ unsigned int s_last;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(acc + (size_t)b * 8 + k, d[k]);
        __threadfence();
        unsigned int* ctr = reinterpret_cast<unsigned int*>(scratch);
        s_last = (atomicAdd(ctr, 1u) == gridDim.x * gridDim.y - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        const double n = (double)h * w * c;
        double ta = 0.0, tc = 0.0;
        for (int s = threadIdx.x; s < nb; s += 256) {
            double* a = acc + (size_t)s * 8;
            double v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) { v[k] = __ldcg(a + k); a[k] = 0.0; }
            double r3[3];
#pragma unroll
            for (int im = 0; im < 3; ++im) {
                const double mean = v[2 * im] / n;
                double var = (v[2 * im + 1] - n * mean * mean) / (n - 1.0);
                if (var < 0) var = 0;
                r3[im] = 1.0 / ((double)((float)sqrt(var)) + 1e-8);
            }
            ta += r3[0] * r3[0] * v[1] - 2.0 * r3[0] * r3[1] * v[6] + r3[1] * r3[1] * v[3];
            tc += r3[1] * r3[1] * v[3] - 2.0 * r3[1] * r3[2] * v[7] + r3[2] * r3[2] * v[5];
        }
        ta = block_sum_d(ta, red);
        tc = block_sum_d(tc, red);
        if (threadIdx.x == 0) {
            const double den = (double)nb * n;
            atomicAdd(out_a, weight * (float)(ta / den));
            atomicAdd(out_c, weight * (float)(tc / den));
            *reinterpret_cast<unsigned int*>(scratch) = 0u;
        }
    }
}

template <int DT, bool POOL>
__global__ void __launch_bounds__(256) feat_fused3_kernel(const void* __restrict__ f, void* pooled, int nb, int h, int w, int c, double* acc,
                                                         double* scratch, float weight, float* out_a, float* out_c) {
    __shared__ double red[32];
    constexpr int V = VecT<DT>::N;
    const int b = blockIdx.y;
    const int cv = c / V;
    const size_t img_vecs = (size_t)h * w * cv;                    // vectors per image
    float l[8];
    double d[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { l[k] = 0.f; d[k] = 0.0; }
    int run = 0;
    if (POOL) {
        const int ho = h / 2, wo = w / 2;
        const unsigned items = (unsigned)ho * wo * cv;
        for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < items; i += gridDim.x * 256u) {
            const unsigned ch = i % cv;
            const unsigned r = i / cv;
            const unsigned ox = r % wo, oy = r / wo;
            const size_t q00 = ((size_t)(2 * oy) * w + 2 * ox) * cv + ch;
            float m[3][V];
#pragma unroll
            for (int im = 0; im < 3; ++im)
#pragma unroll
                for (int j = 0; j < V; ++j) m[im][j] = -INFINITY;
#pragma unroll
            for (int p4 = 0; p4 < 4; ++p4) {
                const size_t off = q00 + (size_t)(p4 >> 1) * w * cv + (size_t)(p4 & 1) * cv;
                float x[V], rr[V], s[V];
                load_vec<DT>(f, (size_t)b * img_vecs + off, x);
                load_vec<DT>(f, (size_t)(nb + b) * img_vecs + off, rr);
                load_vec<DT>(f, (size_t)(2 * nb + b) * img_vecs + off, s);
#pragma unroll
                for (int j = 0; j < V; ++j) {
                    l[0] += x[j]; l[1] = fmaf(x[j], x[j], l[1]);
                    l[2] += rr[j]; l[3] = fmaf(rr[j], rr[j], l[3]);
                    l[4] += s[j]; l[5] = fmaf(s[j], s[j], l[5]);
                    l[6] = fmaf(x[j], rr[j], l[6]); l[7] = fmaf(rr[j], s[j], l[7]);
                    m[0][j] = fmaxf(m[0][j], x[j]); m[1][j] = fmaxf(m[1][j], rr[j]); m[2][j] = fmaxf(m[2][j], s[j]);
                }
            }
            const size_t out_img = (size_t)ho * wo * cv;
#pragma unroll
            for (int im = 0; im < 3; ++im) {
                uint4 o;
                if (DT == MSTK_BF16) {
                    __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
                    for (int j = 0; j < 4; ++j) hh[j] = __floats2bfloat162_rn(m[im][(2 * j) % V], m[im][(2 * j + 1) % V]);
                } else {
                    o.x = __float_as_uint(m[im][0]); o.y = __float_as_uint(m[im][1]); o.z = __float_as_uint(m[im][2 % V]); o.w = __float_as_uint(m[im][3 % V]);
                }
                reinterpret_cast<uint4*>(pooled)[(size_t)(im * nb + b) * out_img + i] = o;
            }
            if (++run == 4) {
#pragma unroll
                for (int k = 0; k < 8; ++k) { d[k] += l[k]; l[k] = 0.f; }
                run = 0;
            }
        }
    } else {
        const unsigned nvec = (unsigned)img_vecs;
        for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < nvec; i += gridDim.x * 256u) {
            float x[V], rr[V], s[V];
            load_vec<DT>(f, (size_t)b * img_vecs + i, x);
            load_vec<DT>(f, (size_t)(nb + b) * img_vecs + i, rr);
            load_vec<DT>(f, (size_t)(2 * nb + b) * img_vecs + i, s);
#pragma unroll
            for (int j = 0; j < V; ++j) {
                l[0] += x[j]; l[1] = fmaf(x[j], x[j], l[1]);
                l[2] += rr[j]; l[3] = fmaf(rr[j], rr[j], l[3]);
                l[4] += s[j]; l[5] = fmaf(s[j], s[j], l[5]);
                l[6] = fmaf(x[j], rr[j], l[6]); l[7] = fmaf(rr[j], s[j], l[7]);
            }
            if (++run == 16) {
#pragma unroll
                for (int k = 0; k < 8; ++k) { d[k] += l[k]; l[k] = 0.f; }
                run = 0;
            }
        }
    }
    fused3_finish(l, d, red, b, nb, h, w, c, acc, scratch, weight, out_a, out_c);
}

// ---------------------------------------------------------------------------
// First VGGish tap of the training step WITHOUT materialising it: conv3x3(1 -> 64) + bias + ReLU of the three stacked
// image batches (x, recon, style) is recomputed on the fly from the single-channel images (9 FMAs per value), the eight
// per-sample sums of the tap and its 2x2 max-pool come out of the same pass.  The stand-alone form wrote the
// [3b,128,128,64] map (1.6 GB bf16 at batch 256, 499 us at ~50 % of the HBM peak) and read it back once for the
// statistics; here only the pooled quarter-size map is written.  The statistics are taken on the fp32 values (the
// materialised form took them on the bf16-rounded map); the pooled map is identical (rounding is monotonic).
// Thread = 4 consecutive channels (weights + bias in 40 registers, two CTAs per SM) of one pooled pixel; its 4x4 input
// patch per image is loaded once (unpredicated in the interior) and shared by the 16 threads of a pixel through L1.
// ---------------------------------------------------------------------------
template <int DT>
__global__ void __launch_bounds__(256, 2) feat_first3_kernel(const float* __restrict__ img, const float* __restrict__ wgt,
                                                            const float* __restrict__ bias, void* pooled, int nb, int h, int w, double* acc,
                                                            double* scratch, float weight, float* out_a, float* out_c) {
    __shared__ double red[32];
    constexpr int C = 64, NC = 4, CV = C / NC;          // thread = 4 consecutive channels; 16 threads per pooled pixel
    const int b = blockIdx.y;
    const int ch = threadIdx.x & (CV - 1);              // constant per thread: the item stride is a multiple of 256
    float wr[9][NC], br[NC];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int j = 0; j < NC; ++j) wr[t][j] = __ldg(wgt + t * C + ch * NC + j);
#pragma unroll
    for (int j = 0; j < NC; ++j) br[j] = bias ? __ldg(bias + ch * NC + j) : 0.f;
    float l[8];
    double d[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { l[k] = 0.f; d[k] = 0.0; }
    const int ho = h / 2, wo = w / 2;
    const unsigned items = (unsigned)ho * wo * CV;
    const size_t plane = (size_t)h * w, out_img = (size_t)ho * wo * CV;
    int run = 0;
    for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < items; i += gridDim.x * 256u) {
        const unsigned r = i / CV;
        const int ox = (int)(r % wo), oy = (int)(r / wo);
        const int y0 = 2 * oy - 1, x0 = 2 * ox - 1;      // top-left of the 4x4 input patch of this 2x2 output block
        const bool interior = y0 >= 0 && x0 >= 0 && y0 + 3 < h && x0 + 3 < w;
        float v[3][4][NC];                               // [image][pixel of the 2x2 block][channel]
#pragma unroll
        for (int im = 0; im < 3; ++im) {
            const float* pl = img + (size_t)(im * nb + b) * plane;
            float pt[4][4];
            if (interior) {
#pragma unroll
                for (int yy = 0; yy < 4; ++yy) {
                    const float* rowp = pl + (size_t)(y0 + yy) * w + x0;
                    pt[yy][0] = __ldg(rowp); pt[yy][1] = __ldg(rowp + 1); pt[yy][2] = __ldg(rowp + 2); pt[yy][3] = __ldg(rowp + 3);
                }
            } else {
#pragma unroll
                for (int yy = 0; yy < 4; ++yy)
#pragma unroll
                    for (int xx = 0; xx < 4; ++xx) {
                        const int gy = y0 + yy, gx = x0 + xx;
                        pt[yy][xx] = (gy >= 0 && gy < h && gx >= 0 && gx < w) ? __ldg(pl + (size_t)gy * w + gx) : 0.f;
                    }
            }
#pragma unroll
            for (int p4 = 0; p4 < 4; ++p4) {
                const int py = p4 >> 1, px = p4 & 1;
#pragma unroll
                for (int j = 0; j < NC; ++j) v[im][p4][j] = br[j];
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    const float pv = pt[py + t / 3][px + t % 3];
#pragma unroll
                    for (int j = 0; j < NC; ++j) v[im][p4][j] = fmaf(pv, wr[t][j], v[im][p4][j]);
                }
#pragma unroll
                for (int j = 0; j < NC; ++j) v[im][p4][j] = fmaxf(v[im][p4][j], 0.f);
            }
        }
#pragma unroll
        for (int p4 = 0; p4 < 4; ++p4)
#pragma unroll
            for (int j = 0; j < NC; ++j) {
                const float xv = v[0][p4][j], rv = v[1][p4][j], sv = v[2][p4][j];
                l[0] += xv; l[1] = fmaf(xv, xv, l[1]);
                l[2] += rv; l[3] = fmaf(rv, rv, l[3]);
                l[4] += sv; l[5] = fmaf(sv, sv, l[5]);
                l[6] = fmaf(xv, rv, l[6]); l[7] = fmaf(rv, sv, l[7]);
            }
#pragma unroll
        for (int im = 0; im < 3; ++im) {
            float m[NC];
#pragma unroll
            for (int j = 0; j < NC; ++j) m[j] = fmaxf(fmaxf(v[im][0][j], v[im][1][j]), fmaxf(v[im][2][j], v[im][3][j]));
            if (DT == MSTK_BF16) {
                uint2 o;
                __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&o);
                hh[0] = __floats2bfloat162_rn(m[0], m[1]); hh[1] = __floats2bfloat162_rn(m[2], m[3]);
                reinterpret_cast<uint2*>(pooled)[(size_t)(im * nb + b) * out_img + i] = o;
            } else {
                reinterpret_cast<float4*>(pooled)[(size_t)(im * nb + b) * out_img + i] = make_float4(m[0], m[1], m[2], m[3]);
            }
        }
        if (++run == 4) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { d[k] += l[k]; l[k] = 0.f; }
            run = 0;
        }
    }
    fused3_finish(l, d, red, b, nb, h, w, C, acc, scratch, weight, out_a, out_c);
}

// ---------------------------------------------------------------------------
// The same first tap on the tensor cores (bf16 mode).  feat_first3_kernel is bound by the FMA pipe: 816 FMA-class
// instructions per thread item (432 of them the 9-tap convolution) at one per two cycles and scheduler = the measured
// 734 us.  Here a warp owns a 2 x 8 pixel patch and computes its 16 pixels x 64 channels as eight mma.sync.m16n8k16 tiles
// per image (K = 9 taps padded to 16): the im2col fragment is built from <= 6 pixel loads per lane; pixels AND weights are
// carried as bf16 hi + lo (three MMAs: hi*hi + lo*hi + hi*lo, ~16 mantissa bits), so the statistics keep the accuracy of the
// fp32 kernel (tests hold the two loss terms to 1e-4).  Accumulator layout: a thread holds pixel (y, x0 + g) in c0/c1 and (y + 1, x0 + g) in c2/c3
// (g = lane / 4), so the vertical half of the 2x2 max-pool is in-thread and the horizontal half one shuffle (lane ^ 4);
// the pooled 4 pixels x 64 channels of a patch are 512 contiguous bytes: staged per warp, written as 16 bytes per lane.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 2) feat_first3_mma_kernel(const float* __restrict__ img, const float* __restrict__ wgt,
                                                                const float* __restrict__ bias, __nv_bfloat16* __restrict__ pooled, int nb, int h,
                                                                int w, double* acc, double* scratch, float weight, float* out_a, float* out_c) {
    __shared__ double red[32];
    __shared__ __align__(16) uint8_t stage[8][3][512];       // per warp and image: pooled 4 pixels x 64 channels bf16
    constexpr int C = 64;
    const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    // B fragments (weights as bf16 hi + lo, so that the statistics keep fp32-level accuracy): tile j = channels 8j..8j+7;
    // b0 = taps (2t, 2t+1), b1 = taps (2t+8, 2t+9) of channel 8j + g.  Kept in shared memory (one word per lane and
    // fragment: conflict-free), not in 48 registers.
    __shared__ uint32_t sbw[4][8][32];                       // [hi b0, hi b1, lo b0, lo b1][tile][lane]
    __shared__ __align__(8) float sbias[64];
    if (warp == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int ch = 8 * j + g;
            uint32_t h0, l0, h1 = 0u, l1 = 0u;
            split_pack(__ldg(wgt + (2 * t) * C + ch), __ldg(wgt + (2 * t + 1) * C + ch), h0, l0);
            if (t == 0) split_pack(__ldg(wgt + 8 * C + ch), 0.f, h1, l1);
            sbw[0][j][lane] = h0; sbw[1][j][lane] = h1; sbw[2][j][lane] = l0; sbw[3][j][lane] = l1;
        }
    }
    if (threadIdx.x < 64) sbias[threadIdx.x] = bias ? __ldg(bias + threadIdx.x) : 0.f;
    __syncthreads();
    float l[8];
    double d[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { l[k] = 0.f; d[k] = 0.0; }
    const int ho = h / 2, wo = w / 2, ux = w / 8;
    const int units = ho * ux;
    const size_t plane = (size_t)h * w;
    // taps of this lane: k0 = 2t, k1 = 2t + 1 (and k = 8 for t == 0)
    const int ky0 = (2 * t) / 3, kx0 = (2 * t) % 3, ky1 = (2 * t + 1) / 3, kx1 = (2 * t + 1) % 3;
    int run = 0;
    for (int u = blockIdx.x * 8 + warp; u < units; u += gridDim.x * 8) {
        const int ry = u / ux, cx = u - ry * ux;
        const int y0 = 2 * ry, x = 8 * cx + g;
        uint32_t ah[3][4], al[3][4];
#pragma unroll
        for (int im = 0; im < 3; ++im) {
            const float* pl = img + (size_t)(im * nb + b) * plane;
            auto px = [&](int yy, int xx) -> float { return (yy >= 0 && yy < h && xx >= 0 && xx < w) ? __ldg(pl + (size_t)yy * w + xx) : 0.f; };
            // rows of the fragment: g -> pixel (y0, x), g + 8 -> pixel (y0 + 1, x); input of tap (ky, kx) = (y + ky - 1, x + kx - 1)
            const float p00 = px(y0 + ky0 - 1, x + kx0 - 1), p01 = px(y0 + ky1 - 1, x + kx1 - 1);
            const float p10 = px(y0 + ky0, x + kx0 - 1), p11 = px(y0 + ky1, x + kx1 - 1);
            split_pack(p00, p01, ah[im][0], al[im][0]);
            split_pack(p10, p11, ah[im][1], al[im][1]);
            if (t == 0) {                                    // tap 8 = (2, 2)
                split_pack(px(y0 + 1, x + 1), 0.f, ah[im][2], al[im][2]);
                split_pack(px(y0 + 2, x + 1), 0.f, ah[im][3], al[im][3]);
            } else {
                ah[im][2] = ah[im][3] = al[im][2] = al[im][3] = 0u;
            }
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            float v[3][4][4];                                // [image][tile of this half][c0..c3]
#pragma unroll
            for (int im = 0; im < 3; ++im)
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const int j = 4 * half + jj;
                    const float2 bj = *reinterpret_cast<const float2*>(&sbias[8 * j + 2 * t]);
                    float c[4] = {bj.x, bj.y, bj.x, bj.y};
                    const uint32_t wh0 = sbw[0][j][lane], wh1 = sbw[1][j][lane];
                    mma_bf16_16816(c, ah[im], wh0, wh1);
                    mma_bf16_16816(c, al[im], wh0, wh1);
                    mma_bf16_16816(c, ah[im], sbw[2][j][lane], sbw[3][j][lane]);
#pragma unroll
                    for (int e = 0; e < 4; ++e) v[im][jj][e] = fmaxf(c[e], 0.f);
                }
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float xv = v[0][jj][e], rv = v[1][jj][e], sv = v[2][jj][e];
                    l[0] += xv; l[1] = fmaf(xv, xv, l[1]);
                    l[2] += rv; l[3] = fmaf(rv, rv, l[3]);
                    l[4] += sv; l[5] = fmaf(sv, sv, l[5]);
                    l[6] = fmaf(xv, rv, l[6]); l[7] = fmaf(rv, sv, l[7]);
                }
            // 2x2 max-pool: rows g / g + 8 are the vertical pair, lane ^ 4 holds the horizontal neighbour
#pragma unroll
            for (int im = 0; im < 3; ++im) {
                __syncwarp();
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    float m0 = fmaxf(v[im][jj][0], v[im][jj][2]), m1 = fmaxf(v[im][jj][1], v[im][jj][3]);
                    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 4));
                    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 4));
                    if (!(g & 1)) {
                        const int ch = 8 * (4 * half + jj) + 2 * t;
                        *reinterpret_cast<uint32_t*>(&stage[warp][im][(g >> 1) * 128 + ch * 2]) = pack_bf16x2(m0, m1);
                    }
                }
                if (half == 1) {
                    // both halves of the 64 channels are staged: 4 pooled pixels x 128 bytes, contiguous in the pooled map
                    __syncwarp();
                    const uint4 o = *reinterpret_cast<const uint4*>(&stage[warp][im][lane * 16]);
                    uint8_t* dst = reinterpret_cast<uint8_t*>(pooled + (((size_t)(im * nb + b) * ho + ry) * wo + 4 * cx) * C);
                    *reinterpret_cast<uint4*>(dst + lane * 16) = o;
                }
            }
        }
        if (++run == 4) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { d[k] += l[k]; l[k] = 0.f; }
            run = 0;
        }
    }
    fused3_finish(l, d, red, b, nb, h, w, C, acc, scratch, weight, out_a, out_c);
}

int launch_feat_first3(const float* img, const float* wgt, const float* bias, void* pooled, int dtype, int b, int h, int w, double* acc,
                       double* scratch, float weight, float* out_a, float* out_c, cudaStream_t st) {
    const long long items = (long long)(h / 2) * (w / 2) * 16;
    dim3 grid(fl_grid_x(items, b), (unsigned)b);
    static int use_mma = -1;
    if (use_mma < 0) { const char* e = getenv("MSTK_FIRST3_MMA"); use_mma = e ? atoi(e) : 1; }
    if (dtype == MSTK_BF16 && use_mma && (w % 8) == 0 && (h % 2) == 0 && (reinterpret_cast<uintptr_t>(pooled) & 15) == 0) {
        const long long units = (long long)(h / 2) * (w / 8);
        dim3 g2(fl_grid_x(units * 32, b), (unsigned)b);
        feat_first3_mma_kernel<<<g2, 256, 0, st>>>(img, wgt, bias, reinterpret_cast<__nv_bfloat16*>(pooled), b, h, w, acc, scratch, weight, out_a, out_c);
        MSTK_CHECK_LAUNCH("feature_first3(mma)");
        return MSTK_OK;
    }
    if (dtype == MSTK_BF16) feat_first3_kernel<MSTK_BF16><<<grid, 256, 0, st>>>(img, wgt, bias, pooled, b, h, w, acc, scratch, weight, out_a, out_c);
    else feat_first3_kernel<MSTK_F32><<<grid, 256, 0, st>>>(img, wgt, bias, pooled, b, h, w, acc, scratch, weight, out_a, out_c);
    MSTK_CHECK_LAUNCH("feature_first3");
    return MSTK_OK;
}

int launch_feat_fused3(const void* f, void* pooled, int dtype, int b, int h, int w, int c, double* acc, double* scratch, float weight,
                       float* out_a, float* out_c, cudaStream_t st) {
    const int v = dtype == MSTK_BF16 ? 8 : 4;
    const long long items = pooled ? (long long)(h / 2) * (w / 2) * (c / v) : (long long)h * w * (c / v);
    dim3 grid(fl_grid_x(items, b), (unsigned)b);
    if (dtype == MSTK_BF16) {
        if (pooled) feat_fused3_kernel<MSTK_BF16, true><<<grid, 256, 0, st>>>(f, pooled, b, h, w, c, acc, scratch, weight, out_a, out_c);
        else feat_fused3_kernel<MSTK_BF16, false><<<grid, 256, 0, st>>>(f, pooled, b, h, w, c, acc, scratch, weight, out_a, out_c);
    } else {
        if (pooled) feat_fused3_kernel<MSTK_F32, true><<<grid, 256, 0, st>>>(f, pooled, b, h, w, c, acc, scratch, weight, out_a, out_c);
        else feat_fused3_kernel<MSTK_F32, false><<<grid, 256, 0, st>>>(f, pooled, b, h, w, c, acc, scratch, weight, out_a, out_c);
    }
    MSTK_CHECK_LAUNCH("feature_fused3");
    return MSTK_OK;
}

bool feat_vec_ok(const void* f, int dtype, long long inner) {
    const int v = dtype == MSTK_BF16 ? 8 : 4;
    return (inner % v) == 0 && (reinterpret_cast<uintptr_t>(f) & 15) == 0 && inner / v < (1ll << 31);
}

int launch_sample_stats(const void* f, int dtype, int samples, long long inner, double* stats, cudaStream_t st) {
    const int v = dtype == MSTK_BF16 ? 8 : 4;
    dim3 grid(fl_grid_x(inner / v, samples), (unsigned)samples);
    if (dtype == MSTK_BF16) sample_stats_kernel<MSTK_BF16><<<grid, 256, 0, st>>>(f, inner, stats);
    else sample_stats_kernel<MSTK_F32><<<grid, 256, 0, st>>>(f, inner, stats);
    MSTK_CHECK_LAUNCH("sample_stats");
    return MSTK_OK;
}

int launch_feat_mse3(const void* f0, const void* f1, const void* f2, int dtype, int b, long long inner, double* st0, double* st1,
                     double* st2, double* scratch, float weight, float* out_a, float* out_c, cudaStream_t st) {
    const int v = dtype == MSTK_BF16 ? 8 : 4;
    dim3 grid(fl_grid_x(inner / v, b), (unsigned)b);
    if (dtype == MSTK_BF16)
        feat_mse3_kernel<MSTK_BF16><<<grid, 256, 0, st>>>(f0, f1, f2, b, inner, st0, st1, st2, scratch, weight, out_a, out_c);
    else
        feat_mse3_kernel<MSTK_F32><<<grid, 256, 0, st>>>(f0, f1, f2, b, inner, st0, st1, st2, scratch, weight, out_a, out_c);
    MSTK_CHECK_LAUNCH("feature_mse3");
    return MSTK_OK;
}

int launch_maxpool_vec(const void* x, void* y, int n, int h, int w, int c, int dtype, cudaStream_t st) {
    const int v = dtype == MSTK_BF16 ? 8 : 4;
    const long long per_img = (long long)(h / 2) * (w / 2) * (c / v);
    dim3 grid(fl_grid_x(per_img, n), (unsigned)n);
    if (dtype == MSTK_BF16) maxpool2x2_vec_kernel<MSTK_BF16><<<grid, 256, 0, st>>>(x, y, n, h, w, c);
    else maxpool2x2_vec_kernel<MSTK_F32><<<grid, 256, 0, st>>>(x, y, n, h, w, c);
    MSTK_CHECK_LAUNCH("maxpool(vec)");
    return MSTK_OK;
}

}  // namespace mstk
