// HBM-streaming kernels for the two degenerate contractions of the hot path, which are
// bandwidth problems, not MMA problems (the round-1a launch list showed the generic 64x64
// CUDA-core GEMM tile spending 1.09 ms on a layer whose traffic is worth ~6 us):
//
//   conv_cin1_kernel      Conv2d k3 p1 (s1/s2) with Cin == 1: the first layer of the VAE
//                         encoder (model.py:17), the style encoder (model.py:61) and the VGGish
//                         stack (vggish.py:108-118).  9 input floats -> Cout outputs per pixel;
//                         one thread owns 8 consecutive output channels of one pixel, so a warp
//                         writes 512 contiguous bytes; BatchNorm statistics are reduced in
//                         registers -> shared memory -> one double atomic per channel and CTA.
//   convt4_cout1_kernel   ConvTranspose2d k4 s2 p1 with Cout == 1: the last decoder layer
//                         (model.py:45) + tanh (+ the callers' (y+1)/2).  One thread owns one
//                         input-grid position = a 2x2 block of output pixels and walks its 3x3
//                         input neighbourhood with 128-bit loads (neighbouring threads share
//                         them through L1); weights live in shared memory as fp32.
#include <cstdlib>
#include "common.cuh"

namespace mstk {

struct EdgeP {
    const void* x; const void* w; const float* bias; void* y; double* stats;
    int N, H, W, Cin, Cout, Ho, Wo, stride, act, affine, dtype, out_dtype;
    float out_scale, out_shift;
};

template <typename TO>
__global__ void __launch_bounds__(256) conv_cin1_kernel(const EdgeP a) {
    extern __shared__ float sm[];                 // w[9][Cout], bias[Cout], ssum[Cout], ssq[Cout]
    float* sw = sm;
    float* sb = sw + 9 * a.Cout;
    float* ssum = sb + a.Cout;
    float* ssq = ssum + a.Cout;
    const float* __restrict__ wsrc = reinterpret_cast<const float*>(a.w);      // packed [9][Cout][1]
    for (int i = threadIdx.x; i < 9 * a.Cout; i += blockDim.x) sw[i] = wsrc[i];
    for (int i = threadIdx.x; i < a.Cout; i += blockDim.x) {
        sb[i] = a.bias ? a.bias[i] : 0.f;
        ssum[i] = 0.f; ssq[i] = 0.f;
    }
    __syncthreads();
    const float* __restrict__ X = reinterpret_cast<const float*>(a.x);
    TO* __restrict__ Y = reinterpret_cast<TO*>(a.y);
    const int CG = a.Cout >> 3;
    const size_t items = (size_t)a.N * a.Ho * a.Wo * CG;
    float s1[8], s2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
    int cg_prev = -1;
    const unsigned per_img = (unsigned)(a.Ho * a.Wo);
    for (size_t it = blockIdx.x * (size_t)blockDim.x + threadIdx.x; it < items; it += (size_t)gridDim.x * blockDim.x) {
        // 32-bit index arithmetic (64-bit div/mod cost more than the 72 FMAs of the item); CG is a power of two
        // for every layer of the model, but stay generic
        int cg, n;
        unsigned r;
        if (items <= 0xFFFFFFFFull) {
            const unsigned i32 = (unsigned)it, pix = i32 / (unsigned)CG;
            cg = (int)(i32 - pix * (unsigned)CG);
            n = (int)(pix / per_img);
            r = pix - (unsigned)n * per_img;
        } else {
            const size_t pix64 = it / (unsigned)CG;
            cg = (int)(it - pix64 * (unsigned)CG);
            n = (int)(pix64 / per_img);
            r = (unsigned)(pix64 - (size_t)n * per_img);
        }
        const int oh = (int)(r / (unsigned)a.Wo);
        const int ow = (int)(r - (unsigned)oh * (unsigned)a.Wo);
        if (a.stats && cg != cg_prev && cg_prev >= 0) {
            // the channel group of this thread changed (only when gridDim*blockDim % CG != 0): flush
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(&ssum[cg_prev * 8 + j], s1[j]); atomicAdd(&ssq[cg_prev * 8 + j], s2[j]);
                s1[j] = 0.f; s2[j] = 0.f;
            }
        }
        cg_prev = cg;
        float xin[9];
        const float* xb = X + (size_t)n * a.H * a.W;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const int ih = oh * a.stride + ky - 1, iw = ow * a.stride + kx - 1;
                xin[ky * 3 + kx] = (ih >= 0 && ih < a.H && iw >= 0 && iw < a.W) ? __ldg(xb + (size_t)ih * a.W + iw) : 0.f;
            }
        float v[8];
        const int c0 = cg * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = sb[c0 + j];
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            const float4 w0 = *reinterpret_cast<const float4*>(sw + t * a.Cout + c0);
            const float4 w1 = *reinterpret_cast<const float4*>(sw + t * a.Cout + c0 + 4);
            v[0] = fmaf(xin[t], w0.x, v[0]); v[1] = fmaf(xin[t], w0.y, v[1]); v[2] = fmaf(xin[t], w0.z, v[2]); v[3] = fmaf(xin[t], w0.w, v[3]);
            v[4] = fmaf(xin[t], w1.x, v[4]); v[5] = fmaf(xin[t], w1.y, v[5]); v[6] = fmaf(xin[t], w1.z, v[6]); v[7] = fmaf(xin[t], w1.w, v[7]);
        }
        if (a.stats) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { s1[j] += v[j]; s2[j] = fmaf(v[j], v[j], s2[j]); }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float x = apply_act(v[j], a.act);
            if (a.affine) x = x * a.out_scale + a.out_shift;
            v[j] = x;
        }
        const size_t off = ((((size_t)n * a.Ho + oh) * a.Wo + ow) * a.Cout) + c0;
        if (sizeof(TO) == 4) {
            float4* o = reinterpret_cast<float4*>(Y + off);
            o[0] = make_float4(v[0], v[1], v[2], v[3]);
            o[1] = make_float4(v[4], v[5], v[6], v[7]);
        } else {
            uint4 pk;
            __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            *reinterpret_cast<uint4*>(Y + off) = pk;
        }
    }
    if (a.stats) {
        if (cg_prev >= 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { atomicAdd(&ssum[cg_prev * 8 + j], s1[j]); atomicAdd(&ssq[cg_prev * 8 + j], s2[j]); }
        }
        __syncthreads();
        for (int c = threadIdx.x; c < a.Cout; c += blockDim.x) {
            atomicAdd(a.stats + c, (double)ssum[c]);
            atomicAdd(a.stats + a.Cout + c, (double)ssq[c]);
        }
    }
}

// Row-walking variant (192 threads; used whenever Cout/8 divides 192, i.e. every layer of the model): a thread keeps
// the 9x8 weights and the bias of ITS channel group in registers and walks SEG consecutive output pixels
// of one row with a sliding 3x3 input window, so an output pixel costs 3 (stride 1) or 6 (stride 2) L1
// loads + 72 FMA + one 16-byte store instead of 9 loads + 18 shared-memory loads + index arithmetic.
template <typename TO, int STRIDE>
__global__ void __launch_bounds__(192, 2) conv_cin1_row_kernel(const EdgeP a) {
    extern __shared__ float sm[];                 // w[9][Cout], bias[Cout], ssum[Cout], ssq[Cout]
    float* sw = sm;
    float* sb = sw + 9 * a.Cout;
    float* ssum = sb + a.Cout;
    float* ssq = ssum + a.Cout;
    const float* __restrict__ wsrc = reinterpret_cast<const float*>(a.w);      // packed [9][Cout][1]
    for (int i = threadIdx.x; i < 9 * a.Cout; i += blockDim.x) sw[i] = wsrc[i];
    for (int i = threadIdx.x; i < a.Cout; i += blockDim.x) {
        sb[i] = a.bias ? a.bias[i] : 0.f;
        ssum[i] = 0.f; ssq[i] = 0.f;
    }
    __syncthreads();
    constexpr int SEG = 16;
    const int CG = a.Cout >> 3;
    const int cg = threadIdx.x % CG;
    const int c0 = cg * 8;
    float w[9][8], bias[8];
#pragma unroll
    for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) w[t][j] = sw[t * a.Cout + c0 + j];
#pragma unroll
    for (int j = 0; j < 8; ++j) bias[j] = sb[c0 + j];
    const float* __restrict__ X = reinterpret_cast<const float*>(a.x);
    TO* __restrict__ Y = reinterpret_cast<TO*>(a.y);
    const unsigned segs = (unsigned)((a.Wo + SEG - 1) / SEG);
    const unsigned runs_per_img = (unsigned)a.Ho * segs;
    const unsigned long long runs = (unsigned long long)a.N * runs_per_img;
    const unsigned rpb = blockDim.x / (unsigned)CG;            // runs per block iteration
    float s1[8], s2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s1[j] = 0.f; s2[j] = 0.f; }
    for (unsigned long long run = (unsigned long long)blockIdx.x * rpb + threadIdx.x / (unsigned)CG; run < runs;
         run += (unsigned long long)gridDim.x * rpb) {
        const int n = (int)(run / runs_per_img);
        const unsigned rr = (unsigned)(run - (unsigned long long)n * runs_per_img);
        const int oh = (int)(rr / segs);
        const int ow0 = (int)(rr - (unsigned)oh * segs) * SEG;
        const float* xb = X + (size_t)n * a.H * a.W;
        const float* rowp[3];
        bool rok[3];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const int ih = oh * STRIDE + ky - 1;
            rok[ky] = ih >= 0 && ih < a.H;
            rowp[ky] = xb + (size_t)(rok[ky] ? ih : 0) * a.W;
        }
        float win[3][3];
        {
            const int iw = ow0 * STRIDE - 1;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                win[ky][0] = (rok[ky] && iw >= 0) ? __ldg(rowp[ky] + iw) : 0.f;
                if (STRIDE == 1) win[ky][1] = rok[ky] ? __ldg(rowp[ky] + iw + 1) : 0.f;
            }
        }
        TO* yrow = Y + (((size_t)n * a.Ho + oh) * a.Wo + ow0) * a.Cout + c0;
        const int npx = a.Wo - ow0 < SEG ? a.Wo - ow0 : SEG;
        // one output pixel from the 3x3 window: 72 FMA, statistics, activation, one 16-byte store, window shift
        auto emit = [&](int i) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = bias[j];
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx)
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] = fmaf(win[ky][kx], w[ky * 3 + kx][j], v[j]);
            if (a.stats) {
#pragma unroll
                for (int j = 0; j < 8; ++j) { s1[j] += v[j]; s2[j] = fmaf(v[j], v[j], s2[j]); }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float x = apply_act(v[j], a.act);
                if (a.affine) x = x * a.out_scale + a.out_shift;
                v[j] = x;
            }
            TO* o = yrow + (unsigned)i * (unsigned)a.Cout;
            if (sizeof(TO) == 4) {
                reinterpret_cast<float4*>(o)[0] = make_float4(v[0], v[1], v[2], v[3]);
                reinterpret_cast<float4*>(o)[1] = make_float4(v[4], v[5], v[6], v[7]);
            } else {
                uint4 pk;
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
                *reinterpret_cast<uint4*>(o) = pk;
            }
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                if (STRIDE == 1) { win[ky][0] = win[ky][1]; win[ky][1] = win[ky][2]; }
                else win[ky][0] = win[ky][2];
            }
        };
        // interior runs (all three input rows exist, no load of the run crosses the right edge): loads through
        // three walking pointers with compile-time offsets, no predicates; stride 2 fetches its two new columns
        // as ONE 8-byte load per row (2*ow is even and W is even, so the pair is 8-byte aligned)
        // (stride 1: only the LAST pixel of an image row reads past the right edge, so its load alone is predicated --
        // the last segment of a row shares its warp with interior segments and must not diverge from them)
        const bool interior = rok[0] && rok[1] && rok[2] && npx == SEG && (STRIDE == 2 ? (a.W & 1) == 0 : true);
        if (interior) {
            if (STRIDE == 2) {
                const float2* q0 = reinterpret_cast<const float2*>(rowp[0] + 2 * ow0);
                const float2* q1 = reinterpret_cast<const float2*>(rowp[1] + 2 * ow0);
                const float2* q2 = reinterpret_cast<const float2*>(rowp[2] + 2 * ow0);
#pragma unroll 4
                for (int i = 0; i < SEG; ++i) {
                    const float2 t0 = __ldg(q0 + i), t1 = __ldg(q1 + i), t2 = __ldg(q2 + i);
                    win[0][1] = t0.x; win[0][2] = t0.y; win[1][1] = t1.x; win[1][2] = t1.y; win[2][1] = t2.x; win[2][2] = t2.y;
                    emit(i);
                }
            } else {
                const float* q0 = rowp[0] + ow0 + 1;
                const float* q1 = rowp[1] + ow0 + 1;
                const float* q2 = rowp[2] + ow0 + 1;
#pragma unroll 5
                for (int i = 0; i < SEG - 1; ++i) {
                    win[0][2] = __ldg(q0 + i); win[1][2] = __ldg(q1 + i); win[2][2] = __ldg(q2 + i);
                    emit(i);
                }
                const bool more = ow0 + SEG < a.W;
                win[0][2] = more ? __ldg(q0 + SEG - 1) : 0.f;
                win[1][2] = more ? __ldg(q1 + SEG - 1) : 0.f;
                win[2][2] = more ? __ldg(q2 + SEG - 1) : 0.f;
                emit(SEG - 1);
            }
            continue;
        }
#pragma unroll 2
        for (int i = 0; i < npx; ++i) {
            const int iw = (ow0 + i) * STRIDE - 1;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                if (STRIDE == 1) {
                    win[ky][2] = (rok[ky] && iw + 2 < a.W) ? __ldg(rowp[ky] + iw + 2) : 0.f;
                } else {
                    win[ky][1] = rok[ky] ? __ldg(rowp[ky] + iw + 1) : 0.f;            // iw + 1 = 2*ow < W always
                    win[ky][2] = (rok[ky] && iw + 2 < a.W) ? __ldg(rowp[ky] + iw + 2) : 0.f;
                }
            }
            emit(i);
        }
    }
    if (a.stats) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { atomicAdd(&ssum[c0 + j], s1[j]); atomicAdd(&ssq[c0 + j], s2[j]); }
        __syncthreads();
        for (int c = threadIdx.x; c < a.Cout; c += blockDim.x) {
            atomicAdd(a.stats + c, (double)ssum[c]);
            atomicAdd(a.stats + a.Cout + c, (double)ssq[c]);
        }
    }
}

template <typename T> __device__ __forceinline__ void load8(const T* p, float (&o)[8]);
template <> __device__ __forceinline__ void load8<float>(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&o)[8]) {
    const uint4 q = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(h[e]); o[2 * e] = f.x; o[2 * e + 1] = f.y; }
}

// x [N,H,W,Cin] (T), w packed [16][1][Cin] (T), y [N,2H,2W,1] (TO)
// One thread = 8 channels of one input-grid position (Cin/8 consecutive lanes share a position, so a warp reads
// whole 128-byte pixel rows); it loads its 3x3 neighbourhood as nine independent 16-byte loads, accumulates the
// four output phases over its 8 channels and the lanes of a position are summed with shuffles.
template <typename T, typename TO>
__global__ void __launch_bounds__(192, 2) convt4_cout1_kernel(const EdgeP a) {
    extern __shared__ float sw[];                 // [16][Cin] fp32
    const T* __restrict__ wsrc = reinterpret_cast<const T*>(a.w);
    for (int i = threadIdx.x; i < 16 * a.Cin; i += blockDim.x) sw[i] = to_f(wsrc[i]);
    __syncthreads();
    const T* __restrict__ X = reinterpret_cast<const T*>(a.x);
    TO* __restrict__ Y = reinterpret_cast<TO*>(a.y);
    const float b0 = a.bias ? a.bias[0] : 0.f;
    const int tpp = a.Cin >> 3;                   // lanes per position: 1, 2, 4, 8, 16 or 32 (host guarantees a power of two)
    const int cg = threadIdx.x & (tpp - 1);
    const int c0 = cg << 3;
    const unsigned ppb = blockDim.x / tpp;
    const unsigned slot = threadIdx.x / tpp;
    const unsigned hw = (unsigned)(a.W * a.H);
    const unsigned long long items = (unsigned long long)a.N * hw;
    // all lanes of a warp run the same number of iterations (positions are padded per warp), so the shuffles are safe
    const unsigned long long iters = (items + (unsigned long long)gridDim.x * ppb - 1) / ((unsigned long long)gridDim.x * ppb);
    for (unsigned long long k = 0; k < iters; ++k) {
        const unsigned long long it = (k * gridDim.x + blockIdx.x) * ppb + slot;
        const bool live = it < items;
        const unsigned long long itc = live ? it : 0;
        const int n = (int)(itc / hw);
        const unsigned r = (unsigned)(itc - (unsigned long long)n * hw);
        const int h = (int)(r / (unsigned)a.W);
        const int w = (int)(r - (unsigned)h * (unsigned)a.W);
        float xv[9][8];
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                const int ih = h + dy, iw = w + dx;
                const int ihc = min(max(ih, 0), a.H - 1), iwc = min(max(iw, 0), a.W - 1);
                load8<T>(X + (((size_t)n * a.H + ihc) * a.W + iwc) * a.Cin + c0, xv[(dy + 1) * 3 + dx + 1]);
                if (ih != ihc || iw != iwc) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) xv[(dy + 1) * 3 + dx + 1][j] = 0.f;
                }
            }
        float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
            for (int dx = -1; dx <= 1; ++dx) {
                const float* xr = xv[(dy + 1) * 3 + dx + 1];
                // phase py receives this row through kernel row ky = py + 1 - 2*dy (valid iff 0 <= ky < 4)
#pragma unroll
                for (int py = 0; py < 2; ++py) {
                    const int ky = py + 1 - 2 * dy;
                    if (ky < 0 || ky > 3) continue;
#pragma unroll
                    for (int px = 0; px < 2; ++px) {
                        const int kx = px + 1 - 2 * dx;
                        if (kx < 0 || kx > 3) continue;
                        const float* wr = sw + (ky * 4 + kx) * a.Cin + c0;
                        const float4 w0 = *reinterpret_cast<const float4*>(wr), w1 = *reinterpret_cast<const float4*>(wr + 4);
                        float s = acc[py][px];
                        s = fmaf(xr[0], w0.x, s); s = fmaf(xr[1], w0.y, s); s = fmaf(xr[2], w0.z, s); s = fmaf(xr[3], w0.w, s);
                        s = fmaf(xr[4], w1.x, s); s = fmaf(xr[5], w1.y, s); s = fmaf(xr[6], w1.z, s); s = fmaf(xr[7], w1.w, s);
                        acc[py][px] = s;
                    }
                }
            }
        // sum over the tpp lanes of this position (tpp is a power of two <= 32 and positions are tpp-aligned in the warp)
        for (int o = tpp >> 1; o > 0; o >>= 1) {
#pragma unroll
            for (int py = 0; py < 2; ++py)
#pragma unroll
                for (int px = 0; px < 2; ++px) acc[py][px] += __shfl_xor_sync(0xffffffffu, acc[py][px], o);
        }
        if (live && cg == 0) {
#pragma unroll
            for (int py = 0; py < 2; ++py) {
                float o[2];
#pragma unroll
                for (int px = 0; px < 2; ++px) {
                    float x = apply_act(acc[py][px] + b0, a.act);
                    if (a.affine) x = x * a.out_scale + a.out_shift;
                    o[px] = x;
                }
                const size_t off = ((size_t)n * a.Ho + (2 * h + py)) * a.Wo + 2 * w;
                if (sizeof(TO) == 4) *reinterpret_cast<float2*>(Y + off) = make_float2(o[0], o[1]);
                else *reinterpret_cast<__nv_bfloat162*>(Y + off) = __floats2bfloat162_rn(o[0], o[1]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// convt4_cout1_mma_kernel: the same layer for bf16 activations with Cin == 64 (the decoder's last layer in bf16 mode)
// as a skinny GEMM + col2im, so that x is read ONCE from HBM (the walking kernel above re-reads every pixel row
// through nine 16-byte loads per position and 32 LDS.128 of weights: 143 us for 16 x 64 x 512 x 64, 0.5 TB/s).
//   P[pos][tap] = sum_c x[pos][c] * w[tap][c]        M = positions of an 8 x 32 tile + 1-pixel halo (340), N = 16 taps, K = 64
//   y[2h+py][2w+px] = sum over the 2 x 2 (dy, dx) neighbours whose tap (ky, kx) = (py+1-2dy, px+1-2dx) exists of P[h+dy][w+dx][ky*4+kx]
// The GEMM runs on mma.sync m16n8k16 (bf16 x bf16 -> fp32) straight from global memory: the contraction does not care
// about the order of the channels, so lane q of a quad owns the 16 CONSECUTIVE channels [16q, 16q+16) of its two rows
// (two 16-byte loads per row) and the weight fragments use the same permutation (16 registers, loaded once).
// P lives in shared memory as [tap][pos] (pos stride 340 = 4 mod 16: the fragment stores and the col2im reads are
// conflict-free).  HBM bound: 128 B read per position + 16 B (fp32) written.
constexpr int CT_TH = 8, CT_TW = 32, CT_PW = CT_TW + 2, CT_NP = (CT_TH + 2) * CT_PW, CT_MT = (CT_NP + 15) / 16;

template <typename TO>
__global__ void __launch_bounds__(256) convt4_cout1_mma_kernel(const EdgeP a, int tiles_w, int tiles_h) {
    __shared__ float P[16 * CT_NP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, q = lane & 3;
    unsigned bid = blockIdx.x;
    const int twi = (int)(bid % (unsigned)tiles_w); bid /= (unsigned)tiles_w;
    const int thi = (int)(bid % (unsigned)tiles_h);
    const int n = (int)(bid / (unsigned)tiles_h);
    const int h0 = thi * CT_TH, w0 = twi * CT_TW;
    const __nv_bfloat16* __restrict__ X = reinterpret_cast<const __nv_bfloat16*>(a.x) + (size_t)n * a.H * a.W * 64;
    // weight fragments: tap = nt*8 + g, channels [16q, 16q+16)
    uint4 wv[2][2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const uint4* wp = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(a.w) + (nt * 8 + g) * 64 + 16 * q);
        wv[nt][0] = __ldg(wp); wv[nt][1] = __ldg(wp + 1);
    }
    for (int mt = warp; mt < CT_MT; mt += 8) {
        uint4 r[2][2];
        int tpos[2];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int t = mt * 16 + g + 8 * half;
            tpos[half] = t;
            const int ty = t / CT_PW, tx = t - ty * CT_PW;
            const int gh = h0 - 1 + ty, gw = w0 - 1 + tx;
            if (t < CT_NP && gh >= 0 && gh < a.H && gw >= 0 && gw < a.W) {
                const uint4* xp = reinterpret_cast<const uint4*>(X + ((size_t)gh * a.W + gw) * 64 + 16 * q);
                r[half][0] = __ldg(xp); r[half][1] = __ldg(xp + 1);
            } else {
                r[half][0] = make_uint4(0u, 0u, 0u, 0u); r[half][1] = r[half][0];
            }
        }
        float c[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        const unsigned* r0 = reinterpret_cast<const unsigned*>(&r[0][0]);
        const unsigned* r1 = reinterpret_cast<const unsigned*>(&r[1][0]);
#pragma unroll
        for (int s4 = 0; s4 < 4; ++s4) {
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const unsigned* wr = reinterpret_cast<const unsigned*>(&wv[nt][0]);
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[nt][0]), "+f"(c[nt][1]), "+f"(c[nt][2]), "+f"(c[nt][3])
                             : "r"(r0[2 * s4]), "r"(r1[2 * s4]), "r"(r0[2 * s4 + 1]), "r"(r1[2 * s4 + 1]), "r"(wr[2 * s4]), "r"(wr[2 * s4 + 1]));
            }
        }
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int half = 0; half < 2; ++half)
                if (tpos[half] < CT_NP) {
                    P[(nt * 8 + 2 * q) * CT_NP + tpos[half]] = c[nt][2 * half];
                    P[(nt * 8 + 2 * q + 1) * CT_NP + tpos[half]] = c[nt][2 * half + 1];
                }
    }
    __syncthreads();
    const int iy = threadIdx.x >> 5, ix = threadIdx.x & 31;
    const int h = h0 + iy, w = w0 + ix;
    if (h >= a.H || w >= a.W) return;
    const float b0 = a.bias ? a.bias[0] : 0.f;
    const int tc = (iy + 1) * CT_PW + ix + 1;
    TO* __restrict__ Y = reinterpret_cast<TO*>(a.y);
#pragma unroll
    for (int py = 0; py < 2; ++py) {
        // (dy, ky) pairs of this output row: (0, py + 1) and (py ? +1 : -1, py ? 0 : 3)
        const int dyb = py ? 1 : -1, kyb = py ? 0 : 3, kya = py + 1;
        float o[2];
#pragma unroll
        for (int px = 0; px < 2; ++px) {
            const int dxb = px ? 1 : -1, kxb = px ? 0 : 3, kxa = px + 1;
            float v = P[(kya * 4 + kxa) * CT_NP + tc] + P[(kya * 4 + kxb) * CT_NP + tc + dxb] +
                      P[(kyb * 4 + kxa) * CT_NP + tc + dyb * CT_PW] + P[(kyb * 4 + kxb) * CT_NP + tc + dyb * CT_PW + dxb];
            v = apply_act(v + b0, a.act);
            if (a.affine) v = v * a.out_scale + a.out_shift;
            o[px] = v;
        }
        const size_t off = ((size_t)n * a.Ho + (2 * h + py)) * a.Wo + 2 * w;
        if (sizeof(TO) == 4) *reinterpret_cast<float2*>(Y + off) = make_float2(o[0], o[1]);
        else *reinterpret_cast<__nv_bfloat162*>(Y + off) = __floats2bfloat162_rn(o[0], o[1]);
    }
}

// ---------------------------------------------------------------------------------------------------------
// conv_cin1_k4s2_mma_kernel: Conv2d(1, 64, k4, s2, p1) with bf16 output -- the data gradient of the decoder's last
// layer (ConvTranspose2d(64, 1, 4, 2, 1) read backwards: dh[pix][c] = sum_tap dy[2oh+ky-1][2ow+kx-1] * w[tap][c]).  It ran on
// the generic CUDA-core tile (114 us at batch 256 for 134 MB of output).  One k16 step of mma.sync per 16 output pixels:
// A = the 16 gathered taps of each pixel, B = the [16 taps][64 channels] weight (n-tile j, column g <-> channel 8g+j, so a
// lane ends up with 2 x 8 consecutive output channels = 16-byte stores, a quad with a whole 128-byte pixel).
// fp32 operands are split into bf16 hi + lo and three MMAs (hi*hi + lo*hi + hi*lo) keep ~16 mantissa bits.
// HBM bound: 128 B written per output pixel, the scalar input stays in L1/L2.
template <typename TO>
__global__ void __launch_bounds__(256) conv_cin1_k4s2_mma_kernel(const EdgeP a) {
    const int lane = threadIdx.x & 31, g = lane >> 2, qd = lane & 3;
    const float* __restrict__ Wt = reinterpret_cast<const float*>(a.w);       // packed [16][64][1]
    unsigned bh[8][2], bl[8][2];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int k = 2 * qd + 8 * r;
            const float w0 = Wt[k * 64 + 8 * g + j], w1 = Wt[(k + 1) * 64 + 8 * g + j];
            const __nv_bfloat162 hi = __floats2bfloat162_rn(w0, w1);
            const float2 hf = __bfloat1622float2(hi);
            const __nv_bfloat162 lo = __floats2bfloat162_rn(w0 - hf.x, w1 - hf.y);
            bh[j][r] = *reinterpret_cast<const unsigned*>(&hi);
            bl[j][r] = *reinterpret_cast<const unsigned*>(&lo);
        }
    float bias0[8], bias1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        bias0[j] = a.bias ? a.bias[16 * qd + j] : 0.f;
        bias1[j] = a.bias ? a.bias[16 * qd + 8 + j] : 0.f;
    }
    const float* __restrict__ X = reinterpret_cast<const float*>(a.x);
    TO* __restrict__ Y = reinterpret_cast<TO*>(a.y);
    const int segs = a.Wo >> 4;
    const long long steps = (long long)a.N * a.Ho * segs;
    const int kyb = qd >> 1, kxb = 2 * (qd & 1);            // taps 2q, 2q+1: (kyb, kxb), (kyb, kxb+1); taps 2q+8, 2q+9: ky + 2
    for (long long st = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); st < steps; st += (long long)gridDim.x * 8) {
        const int seg = (int)(st % segs);
        const long long row = st / segs;
        const int oh = (int)(row % a.Ho), n = (int)(row / a.Ho);
        const int ow0 = seg << 4;
        float sv[4][2];                                      // [a-register r: (pixel g | g+8, ky | ky+2)][column pair]
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int ow = ow0 + g + 8 * (r & 1);
            const int ih = 2 * oh + kyb + 2 * (r >> 1) - 1;
            const bool rok = ih >= 0 && ih < a.H;
            const float* xr = X + ((size_t)n * a.H + (rok ? ih : 0)) * a.W;
            const int iw = 2 * ow + kxb - 1;
            sv[r][0] = (rok && iw >= 0) ? __ldg(xr + (iw >= 0 ? iw : 0)) : 0.f;
            sv[r][1] = (rok && iw + 1 < a.W) ? __ldg(xr + (iw + 1 < a.W ? iw + 1 : 0)) : 0.f;
        }
        unsigned ah[4], al[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const __nv_bfloat162 hi = __floats2bfloat162_rn(sv[r][0], sv[r][1]);
            const float2 hf = __bfloat1622float2(hi);
            const __nv_bfloat162 lo = __floats2bfloat162_rn(sv[r][0] - hf.x, sv[r][1] - hf.y);
            ah[r] = *reinterpret_cast<const unsigned*>(&hi);
            al[r] = *reinterpret_cast<const unsigned*>(&lo);
        }
        float c[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            c[j][0] = bias0[j]; c[j][1] = bias1[j]; c[j][2] = bias0[j]; c[j][3] = bias1[j];
#define MSTK_MMA(A, B)                                                                                                      \
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" \
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])                                        \
                         : "r"(A[0]), "r"(A[1]), "r"(A[2]), "r"(A[3]), "r"(B[j][0]), "r"(B[j][1]))
            MSTK_MMA(al, bh); MSTK_MMA(ah, bl); MSTK_MMA(ah, bh);
#undef MSTK_MMA
        }
        // c[j][0]: pixel g, channel 16q+j; c[j][1]: pixel g, channel 16q+8+j; c[j][2], c[j][3]: pixel g+8
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            TO* o = Y + (((size_t)row * a.Wo + ow0 + g + 8 * half) * 64 + 16 * qd);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                float v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float x = apply_act(c[j][2 * half + e], a.act);
                    if (a.affine) x = x * a.out_scale + a.out_shift;
                    v[j] = x;
                }
                if (sizeof(TO) == 4) {
                    reinterpret_cast<float4*>(o + 8 * e)[0] = make_float4(v[0], v[1], v[2], v[3]);
                    reinterpret_cast<float4*>(o + 8 * e)[1] = make_float4(v[4], v[5], v[6], v[7]);
                } else {
                    uint4 pk;
                    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                    for (int t = 0; t < 4; ++t) h[t] = __floats2bfloat162_rn(v[2 * t], v[2 * t + 1]);
                    *reinterpret_cast<uint4*>(o + 8 * e) = pk;
                }
            }
        }
    }
}

static bool cin1_k4s2_plan(const ConvP& p) {
    if (!(p.nphases == 1 && p.Cin == 1 && p.Cout == 64 && p.ph[0].ntaps == 16 && p.s_in == 2 && p.dtype == MSTK_F32 && !p.stats &&
          p.out_dtype == MSTK_BF16 && (p.Wo & 15) == 0 && p.Ho * 2 == p.H && p.Wo * 2 == p.W)) return false;
    for (int t = 0; t < 16; ++t)
        if (p.ph[0].taps[t].slice != t || p.ph[0].taps[t].dy != t / 4 - 1 || p.ph[0].taps[t].dx != t % 4 - 1) return false;
    return getenv("MSTK_CIN1_K4_SIMT") == nullptr;
}

static EdgeP make_edge(const ConvP& p) {
    EdgeP a;
    a.x = p.x; a.w = p.w; a.bias = p.bias; a.y = p.y; a.stats = p.stats;
    a.N = p.N; a.H = p.H; a.W = p.W; a.Cin = p.Cin; a.Cout = p.Cout; a.Ho = p.Ho; a.Wo = p.Wo;
    a.stride = p.s_in; a.act = p.act; a.affine = p.affine; a.dtype = p.dtype; a.out_dtype = p.out_dtype;
    a.out_scale = p.out_scale; a.out_shift = p.out_shift;
    return a;
}

bool conv_edge_supported(const ConvP& p) {
    if (p.add || p.add_nc || p.has_ddim || !p.y) return false;
    if ((reinterpret_cast<uintptr_t>(p.x) & 15) || (reinterpret_cast<uintptr_t>(p.y) & 15) || (reinterpret_cast<uintptr_t>(p.w) & 15)) return false;
    if (p.nphases == 1 && p.Cin == 1 && p.ph[0].ntaps == 9 && p.dtype == MSTK_F32 && (p.Cout % 8) == 0 && p.Cout <= 1024) return true;
    if (cin1_k4s2_plan(p)) return true;
    if (p.nphases == 4 && p.ph[3].ntaps == 4 && p.ph[0].ntaps == 4 && p.Cout == 1 && (p.Cin % 8) == 0 && p.Cin <= 256 &&
        ((p.Cin / 8) & (p.Cin / 8 - 1)) == 0 && !p.stats) return true;      // Cin/8 lanes per position: a power of two <= 32
    return false;
}

int launch_conv_edge(const ConvP& p, cudaStream_t st) {
    const EdgeP a = make_edge(p);
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (p.nphases == 1 && p.ph[0].ntaps == 16) {
        const long long steps = (long long)p.N * p.Ho * (p.Wo / 16);
        long long blocks = (steps + 7) / 8;
        if (blocks > 4LL * sms) blocks = 4LL * sms;
        conv_cin1_k4s2_mma_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(a);
        MSTK_CHECK_LAUNCH("conv_cin1_k4s2_mma launch");
        return MSTK_OK;
    }
    if (p.nphases == 1) {
        const size_t items = (size_t)p.N * p.Ho * p.Wo * (p.Cout / 8);
        size_t blocks = (items + 255) / 256;
        // 256 threads: a thread keeps its channel group across the grid-stride loop whenever Cout/8 divides 256
        const size_t cap = (size_t)sms * 8;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        const size_t smem = (size_t)(9 + 3) * p.Cout * sizeof(float);
        const int cgn = p.Cout / 8;
        if (192 % cgn == 0) {
            // row-walking kernel: one thread = 16 consecutive output pixels of its channel group
            const size_t runs = (size_t)p.N * p.Ho * ((p.Wo + 15) / 16);
            const size_t rpb = 192 / cgn;
            size_t rb = (runs + rpb - 1) / rpb;
            const size_t rcap = (size_t)sms * 6;
            if (rb > rcap) rb = rcap;
            if (rb < 1) rb = 1;
            const bool of = p.out_dtype == MSTK_F32;
            if (a.stride == 1) {
                if (of) conv_cin1_row_kernel<float, 1><<<(unsigned)rb, 192, smem, st>>>(a);
                else conv_cin1_row_kernel<__nv_bfloat16, 1><<<(unsigned)rb, 192, smem, st>>>(a);
            } else {
                if (of) conv_cin1_row_kernel<float, 2><<<(unsigned)rb, 192, smem, st>>>(a);
                else conv_cin1_row_kernel<__nv_bfloat16, 2><<<(unsigned)rb, 192, smem, st>>>(a);
            }
            MSTK_CHECK_LAUNCH("conv_cin1_row launch");
            return MSTK_OK;
        }
        if (p.out_dtype == MSTK_F32) conv_cin1_kernel<float><<<(unsigned)blocks, 256, smem, st>>>(a);
        else conv_cin1_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, smem, st>>>(a);
        MSTK_CHECK_LAUNCH("conv_cin1 launch");
    } else {
        const size_t items = (size_t)p.N * p.H * p.W;
        const size_t ppb = 192 / (p.Cin / 8);
        size_t blocks = (items + ppb - 1) / ppb;
        const size_t cap = (size_t)sms * 6;
        if (blocks > cap) blocks = cap;
        const size_t smem = (size_t)16 * p.Cin * sizeof(float);
        const bool xf = p.dtype == MSTK_F32, of = p.out_dtype == MSTK_F32;
        const size_t tw = (p.W + CT_TW - 1) / CT_TW, th = (p.H + CT_TH - 1) / CT_TH;
        static const bool no_mma = getenv("MSTK_CONVT_WALK") != nullptr;      // A/B switch: the walking kernel for every shape
        if (!xf && p.Cin == 64 && !no_mma && (size_t)p.N * tw * th < 0x7FFFFFFFull) {
            const unsigned tiles = (unsigned)((size_t)p.N * tw * th);
            if (of) convt4_cout1_mma_kernel<float><<<tiles, 256, 0, st>>>(a, (int)tw, (int)th);
            else convt4_cout1_mma_kernel<__nv_bfloat16><<<tiles, 256, 0, st>>>(a, (int)tw, (int)th);
            MSTK_CHECK_LAUNCH("convt4_cout1_mma launch");
            return MSTK_OK;
        }
        if (xf && of) convt4_cout1_kernel<float, float><<<(unsigned)blocks, 192, smem, st>>>(a);
        else if (xf) convt4_cout1_kernel<float, __nv_bfloat16><<<(unsigned)blocks, 192, smem, st>>>(a);
        else if (of) convt4_cout1_kernel<__nv_bfloat16, float><<<(unsigned)blocks, 192, smem, st>>>(a);
        else convt4_cout1_kernel<__nv_bfloat16, __nv_bfloat16><<<(unsigned)blocks, 192, smem, st>>>(a);
        MSTK_CHECK_LAUNCH("convt4_cout1 launch");
    }
    return MSTK_OK;
}

}  // namespace mstk
