// Backward pass and optimizer kernels of the training step (reference: autograd over
// models/model.py + torch.optim.Adam at models/train.py:156-201).
//
//   wgrad_simt_kernel     weight gradient of the whole convolution family as a tap-gather
//                         GEMM with K = pixels, CUDA cores, fp32 accumulate, split over pixel
//                         ranges, atomically accumulated in the parameter's own layout
//   act_bwd_kernel        ReLU / tanh backward (+ skip-gradient sum) + bias gradient
//   bn_bwd_*              BatchNorm2d training-mode backward (reduce, apply)
//   attention_bwd_kernel  softmax-attention backward per (sample, head)
//   small streaming ops   sum over pixels, GELU', broadcast add, loss gradients, Adam
#include <cstdlib>
#include "common.cuh"

namespace mstk {

static int bw_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) sms = v;
        else sms = 148;
    }
    return sms;
}
static unsigned bw_grid(size_t items, int threads, int per_sm = 8) {
    size_t blocks = (items + threads - 1) / threads;
    const size_t cap = (size_t)bw_sms() * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}

__device__ __forceinline__ void load4_rt(const void* p, int dtype, size_t idx, float (&o)[4]) {
    if (dtype == MSTK_F32) {
        const float4 q = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + idx);
        o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w;
    } else {
        const uint2 q = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p) + idx);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
        const float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
        o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
    }
}

// ---------------------------------------------------------------------------
// weight gradient
// ---------------------------------------------------------------------------
struct WgradP {
    ConvP p;              // forward plan; p.x = forward input, p.y = dy (Ho x Wo x Cout)
    float* dw;
    int x_dtype, dy_dtype;
    long long s_co, s_ci; // strides of (co, ci) in the parameter layout; the tap index is innermost
    int ksplit;
    int npt;              // number of (phase, tap) pairs
};

constexpr int WG_BM = 64, WG_BN = 64, WG_BK = 16, WG_LD = 68;

__global__ void __launch_bounds__(256) wgrad_simt_kernel(const WgradP q) {
    __shared__ __align__(16) float As[WG_BK][WG_LD];   // dy rows  [pixel][co]
    __shared__ __align__(16) float Bs[WG_BK][WG_LD];   // x rows   [pixel][ci]
    const ConvP& p = q.p;
    int pt = blockIdx.z / q.ksplit;
    const int ks = blockIdx.z - pt * q.ksplit;
    int phase_idx = 0;
    while (pt >= p.ph[phase_idx].ntaps) { pt -= p.ph[phase_idx].ntaps; ++phase_idx; }
    const Phase& ph = p.ph[phase_idx];
    const Tap tap = ph.taps[pt];
    const long long K = (long long)p.N * p.Hg * p.Wg;
    const long long k_begin = K * ks / q.ksplit, k_end = K * (ks + 1) / q.ksplit;
    const int co0 = blockIdx.x * WG_BM, ci0 = blockIdx.y * WG_BN;
    const int tid = threadIdx.x;
    const int lrow = tid >> 4;            // pixel row of the k-chunk this thread loads
    const int lc = (tid & 15) * 4;        // first of its 4 channels inside the 64-wide tile
    const bool vec_a = (p.Cout & 3) == 0, vec_b = (p.Cin & 3) == 0;
    const int ty = tid >> 4, tx = tid & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (long long k0 = k_begin; k0 < k_end; k0 += WG_BK) {
        float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
        const long long kp = k0 + lrow;
        if (kp < k_end) {
            const int n = (int)(kp / ((long long)p.Hg * p.Wg));
            const int r = (int)(kp - (long long)n * p.Hg * p.Wg);
            const int hg = r / p.Wg, wg = r - hg * p.Wg;
            const int oh = hg * p.s_out + ph.oy, ow = wg * p.s_out + ph.ox;
            const int ih = hg * p.s_in + tap.dy, iw = wg * p.s_in + tap.dx;
            const size_t yo = (((size_t)n * p.Ho + oh) * p.Wo + ow) * (size_t)p.Cout;
            const int ca = co0 + lc;
            if (vec_a) {
                if (ca < p.Cout) load4_rt(p.y, q.dy_dtype, yo + ca, av);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (ca + j < p.Cout) av[j] = load_as_f(p.y, q.dy_dtype, yo + ca + j);
            }
            if (ih >= 0 && ih < p.H && iw >= 0 && iw < p.W) {
                const size_t xo = (((size_t)n * p.H + ih) * p.W + iw) * (size_t)p.Cin;
                const int cb = ci0 + lc;
                if (vec_b) {
                    if (cb < p.Cin) load4_rt(p.x, q.x_dtype, xo + cb, bv);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (cb + j < p.Cin) bv[j] = load_as_f(p.x, q.x_dtype, xo + cb + j);
                }
            }
        }
        *reinterpret_cast<float4*>(&As[lrow][lc]) = make_float4(av[0], av[1], av[2], av[3]);
        *reinterpret_cast<float4*>(&Bs[lrow][lc]) = make_float4(bv[0], bv[1], bv[2], bv[3]);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < WG_BK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float ar[4] = {a.x, a.y, a.z, a.w};
            const float br[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int co = co0 + ty * 4 + i;
        if (co >= p.Cout) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int ci = ci0 + tx * 4 + j;
            if (ci < p.Cin) atomicAdd(q.dw + (long long)co * q.s_co + (long long)ci * q.s_ci + tap.slice, acc[i][j]);
        }
    }
}

// ---------------------------------------------------------------------------
// vector helpers: 8 consecutive elements of a runtime-typed tensor
// ---------------------------------------------------------------------------
__device__ __forceinline__ void load8_rt(const void* p, int dtype, size_t i8, float (&v)[8]) {
    if (dtype == MSTK_F32) {
        const float4 a = reinterpret_cast<const float4*>(p)[2 * i8], b = reinterpret_cast<const float4*>(p)[2 * i8 + 1];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
        const uint4 q = reinterpret_cast<const uint4*>(p)[i8];
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float2 f = __bfloat1622float2(h[k]); v[2 * k] = f.x; v[2 * k + 1] = f.y; }
    }
}
__device__ __forceinline__ void store8_rt(void* p, int dtype, size_t i8, const float (&v)[8]) {
    if (dtype == MSTK_F32) {
        reinterpret_cast<float4*>(p)[2 * i8] = make_float4(v[0], v[1], v[2], v[3]);
        reinterpret_cast<float4*>(p)[2 * i8 + 1] = make_float4(v[4], v[5], v[6], v[7]);
    } else {
        uint4 q;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
        for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
        reinterpret_cast<uint4*>(p)[i8] = q;
    }
}

// ---------------------------------------------------------------------------
// activation backward + bias gradient
// ---------------------------------------------------------------------------
struct ActBwdP {
    const void* dy; const void* dy2; const void* y; void* g; float* dbias;
    long long rows; int c, act; float y_scale, y_shift; int dy_dtype, y_dtype, g_dtype;
    int vec;            // 8 / 4: the 128-bit paths (c % 8 == 0 / single channel) are usable -- sizes AND base addresses allow them; 0: scalar
};
__device__ __forceinline__ float act_deriv(float y, int act, float y_scale, float y_shift) {
    if (act == MSTK_ACT_RELU) return y > 0.f ? y_scale : 0.f;
    if (act == MSTK_ACT_TANH) { const float t = (y - y_shift) / y_scale; return (1.f - t * t) * y_scale; }
    return y_scale;
}
// vector path (a.vec == 8): c % 8 == 0 and 16-byte aligned tensors; per-thread column sums are flushed through shared memory
__global__ void __launch_bounds__(256) act_bwd_kernel(const ActBwdP a) {
    extern __shared__ float scol[];     // [c] when dbias
    if (a.dbias) {
        for (int i = threadIdx.x; i < a.c; i += blockDim.x) scol[i] = 0.f;
        __syncthreads();
    }
    const size_t total = (size_t)a.rows * a.c;
    if (a.vec == 8) {
        const size_t nvec = total >> 3;
        float s[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) s[k] = 0.f;
        int ch_prev = -1;
        ChanWalk cw((blockIdx.x * (size_t)blockDim.x + threadIdx.x) << 3, ((size_t)gridDim.x * blockDim.x) << 3, a.c);
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x, cw.next()) {
            const int ch = cw.ch;
            float d[8], d2[8], y[8];
            load8_rt(a.dy, a.dy_dtype, i, d);
            if (a.dy2) {
                load8_rt(a.dy2, a.dy_dtype, i, d2);
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] += d2[k];
            }
            if (a.act != MSTK_ACT_NONE) {
                load8_rt(a.y, a.y_dtype, i, y);
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] *= act_deriv(y[k], a.act, a.y_scale, a.y_shift);
            } else if (a.y_scale != 1.f) {
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] *= a.y_scale;
            }
            if (a.g) store8_rt(a.g, a.g_dtype, i, d);
            if (a.dbias) {
                if (ch != ch_prev && ch_prev >= 0) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) { atomicAdd(&scol[ch_prev + k], s[k]); s[k] = 0.f; }
                }
                ch_prev = ch;
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] += d[k];
            }
        }
        if (a.dbias && ch_prev >= 0) {
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(&scol[ch_prev + k], s[k]);
        }
    } else if (a.vec == 4) {
        // single channel (the decoder's output layer): 4 elements per thread, the bias gradient is ONE sum --
        // registers -> warp shuffle -> one shared-memory atomic per warp (the generic path below would issue one
        // shared-memory atomic per ELEMENT on the same address)
        const size_t nvec = total >> 2;
        float s = 0.f;
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
            float d[4], d2[4], y[4];
            load4_rt(a.dy, a.dy_dtype, i << 2, d);
            if (a.dy2) {
                load4_rt(a.dy2, a.dy_dtype, i << 2, d2);
#pragma unroll
                for (int k = 0; k < 4; ++k) d[k] += d2[k];
            }
            if (a.act != MSTK_ACT_NONE) {
                load4_rt(a.y, a.y_dtype, i << 2, y);
#pragma unroll
                for (int k = 0; k < 4; ++k) d[k] *= act_deriv(y[k], a.act, a.y_scale, a.y_shift);
            } else if (a.y_scale != 1.f) {
#pragma unroll
                for (int k = 0; k < 4; ++k) d[k] *= a.y_scale;
            }
            if (a.g) {
                if (a.g_dtype == MSTK_F32) {
                    *reinterpret_cast<float4*>(reinterpret_cast<float*>(a.g) + (i << 2)) = make_float4(d[0], d[1], d[2], d[3]);
                } else {
                    uint2 pk;
                    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
                    h[0] = __floats2bfloat162_rn(d[0], d[1]); h[1] = __floats2bfloat162_rn(d[2], d[3]);
                    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(a.g) + (i << 2)) = pk;
                }
            }
            s += (d[0] + d[1]) + (d[2] + d[3]);
        }
        if (a.dbias) {
            s = warp_sum(s);
            if ((threadIdx.x & 31) == 0) atomicAdd(&scol[0], s);
        }
    } else {
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
            float d = load_as_f(a.dy, a.dy_dtype, i);
            if (a.dy2) d += load_as_f(a.dy2, a.dy_dtype, i);
            const float yv = a.act != MSTK_ACT_NONE ? load_as_f(a.y, a.y_dtype, i) : 0.f;
            d *= act_deriv(yv, a.act, a.y_scale, a.y_shift);
            if (a.g) store_from_f(a.g, a.g_dtype, i, d);
            if (a.dbias) atomicAdd(&scol[(int)(i % (size_t)a.c)], d);
        }
    }
    if (a.dbias) {
        __syncthreads();
        for (int i = threadIdx.x; i < a.c; i += blockDim.x) atomicAdd(a.dbias + i, scol[i]);
    }
}

// ---------------------------------------------------------------------------
// BatchNorm backward
// ---------------------------------------------------------------------------
struct BnBwdP {
    const void* dy; const void* v; void* dv; const float* mean; const float* invstd; const float* gamma; const float* beta;
    double* sums; float* dgamma; float* dbeta; long long rows, count; int c, act, dy_dtype, v_dtype, dv_dtype;
};
// g = dy * relu-mask(gamma*xhat+beta); requires c % 8 == 0
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const BnBwdP a) {
    extern __shared__ float sm[];        // s0[c], s1[c]
    float* s0 = sm; float* s1 = sm + a.c;
    for (int i = threadIdx.x; i < 2 * a.c; i += blockDim.x) sm[i] = 0.f;
    __syncthreads();
    const size_t nvec = ((size_t)a.rows * a.c) >> 3;
    float p0[8], p1[8], pm[8], pis[8], pg[8], pb[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { p0[k] = 0.f; p1[k] = 0.f; pm[k] = 0.f; pis[k] = 0.f; pg[k] = 0.f; pb[k] = 0.f; }
    int ch_prev = -1;
    ChanWalk cw((blockIdx.x * (size_t)blockDim.x + threadIdx.x) << 3, ((size_t)gridDim.x * blockDim.x) << 3, a.c);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x, cw.next()) {
        const int ch = cw.ch;
        if (ch != ch_prev && ch_prev >= 0) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { atomicAdd(&s0[ch_prev + k], p0[k]); atomicAdd(&s1[ch_prev + k], p1[k]); p0[k] = 0.f; p1[k] = 0.f; }
        }
        if (ch != ch_prev) {
            // per-channel parameters in registers: the grid stride is a multiple of c whenever c divides 2048 * gridDim, so a
            // thread normally keeps its 8 channels for the whole pass (the loop did 4 global loads per ELEMENT)
#pragma unroll
            for (int k = 0; k < 8; ++k) { pm[k] = a.mean[ch + k]; pis[k] = a.invstd[ch + k]; pg[k] = a.gamma[ch + k]; pb[k] = a.beta[ch + k]; }
        }
        ch_prev = ch;
        float d[8], v[8];
        load8_rt(a.dy, a.dy_dtype, i, d);
        load8_rt(a.v, a.v_dtype, i, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float xh = (v[k] - pm[k]) * pis[k];
            float g = d[k];
            if (a.act == MSTK_ACT_RELU && !(pg[k] * xh + pb[k] > 0.f)) g = 0.f;
            p0[k] += g; p1[k] = fmaf(g, xh, p1[k]);
        }
    }
    if (ch_prev >= 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { atomicAdd(&s0[ch_prev + k], p0[k]); atomicAdd(&s1[ch_prev + k], p1[k]); }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.c; i += blockDim.x) {
        atomicAdd(a.sums + i, (double)s0[i]);
        atomicAdd(a.sums + a.c + i, (double)s1[i]);
    }
}
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const BnBwdP a) {
    extern __shared__ float sm[];        // m0[c] = sum g / count, m1[c] = sum g*xhat / count
    float* m0 = sm; float* m1 = sm + a.c;
    for (int i = threadIdx.x; i < a.c; i += blockDim.x) {
        m0[i] = (float)(a.sums[i] / (double)a.count);
        m1[i] = (float)(a.sums[a.c + i] / (double)a.count);
        if (blockIdx.x == 0) {
            if (a.dbeta) a.dbeta[i] += (float)a.sums[i];
            if (a.dgamma) a.dgamma[i] += (float)a.sums[a.c + i];
        }
    }
    __syncthreads();
    const size_t nvec = ((size_t)a.rows * a.c) >> 3;
    float pm[8], pis[8], pg[8], pb[8], q0[8], q1[8];
    int chc = -1;
    ChanWalk cw((blockIdx.x * (size_t)blockDim.x + threadIdx.x) << 3, ((size_t)gridDim.x * blockDim.x) << 3, a.c);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x, cw.next()) {
        const int ch = cw.ch;
        if (ch != chc) {                     // this thread's per-channel parameters in registers (normally loaded once)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                pm[k] = a.mean[ch + k]; pis[k] = a.invstd[ch + k]; pg[k] = a.gamma[ch + k]; pb[k] = a.beta[ch + k];
                q0[k] = m0[ch + k]; q1[k] = m1[ch + k];
            }
            chc = ch;
        }
        float d[8], v[8];
        load8_rt(a.dy, a.dy_dtype, i, d);
        load8_rt(a.v, a.v_dtype, i, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float xh = (v[k] - pm[k]) * pis[k];
            float g = d[k];
            if (a.act == MSTK_ACT_RELU && !(pg[k] * xh + pb[k] > 0.f)) g = 0.f;
            d[k] = pg[k] * pis[k] * (g - q0[k] - xh * q1[k]);
        }
        store8_rt(a.dv, a.dv_dtype, i, d);
    }
}

// ---------------------------------------------------------------------------
// attention backward: one CTA per (sample, head); K, V staged as fp32, dK / dV accumulated
// in shared memory, one warp per query row (probabilities recomputed).
// ---------------------------------------------------------------------------
struct AttnBwdP {
    const void* q; const void* k; const void* v; const void* d_out; void* dq; void* dk; void* dv;
    int b, l, s, e, heads, qs, ks, vs, dos, dqs, dks, dvs, dtype;
};
constexpr int AB_WARPS = 8;
__global__ void __launch_bounds__(AB_WARPS * 32) attention_bwd_kernel(const AttnBwdP a) {
    extern __shared__ float sm[];
    const int d = a.e / a.heads;
    float* Ks = sm;                                  // [s][d+1]
    float* Vs = Ks + (size_t)a.s * (d + 1);          // [s][d+1]
    float* dKs = Vs + (size_t)a.s * (d + 1);         // [s][d]
    float* dVs = dKs + (size_t)a.s * d;              // [s][d]
    float* Ps = dVs + (size_t)a.s * d;               // [warps][s]  probabilities, then dS
    float* Qs = Ps + AB_WARPS * a.s;                 // [warps][d]  scaled q
    float* Os = Qs + AB_WARPS * d;                   // [warps][d]  d_out row
    const int b = blockIdx.x / a.heads, hd = blockIdx.x - b * a.heads;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < a.s * d; i += blockDim.x) {
        const int si = i / d, di = i - si * d;
        Ks[si * (d + 1) + di] = load_as_f(a.k, a.dtype, ((size_t)b * a.s + si) * a.ks + hd * d + di);
        Vs[si * (d + 1) + di] = load_as_f(a.v, a.dtype, ((size_t)b * a.s + si) * a.vs + hd * d + di);
        dKs[i] = 0.f; dVs[i] = 0.f;
    }
    __syncthreads();
    const float scaling = 1.0f / sqrtf((float)d);
    float* pw = Ps + warp * a.s;
    float* qw = Qs + warp * d;
    float* ow = Os + warp * d;
    for (int qi = warp; qi < a.l; qi += AB_WARPS) {
        for (int di = lane; di < d; di += 32) {
            qw[di] = load_as_f(a.q, a.dtype, ((size_t)b * a.l + qi) * a.qs + hd * d + di) * scaling;
            ow[di] = load_as_f(a.d_out, a.dtype, ((size_t)b * a.l + qi) * a.dos + hd * d + di);
        }
        __syncwarp();
        // probabilities
        float mx = -INFINITY;
        for (int si = lane; si < a.s; si += 32) {
            float acc = 0.f;
            const float* kr = Ks + si * (d + 1);
            for (int di = 0; di < d; ++di) acc = fmaf(qw[di], kr[di], acc);
            pw[si] = acc;
            mx = fmaxf(mx, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int si = lane; si < a.s; si += 32) { const float ev = expf(pw[si] - mx); pw[si] = ev; sum += ev; }
        sum = warp_sum(sum);
        const float inv = 1.f / sum;
        // dP = d_out . V^T ; rowdot = sum p*dP ; dS = p*(dP - rowdot)
        float rowdot = 0.f;
        float dp_local[4];                           // s <= 128 -> at most 4 keys per lane
        int cnt = 0;
        for (int si = lane; si < a.s; si += 32, ++cnt) {
            const float pr = pw[si] * inv;
            float acc = 0.f;
            const float* vr = Vs + si * (d + 1);
            for (int di = 0; di < d; ++di) acc = fmaf(ow[di], vr[di], acc);
            dp_local[cnt] = acc;
            rowdot = fmaf(pr, acc, rowdot);
            // dV[s] += p * d_out
            for (int di = 0; di < d; ++di) atomicAdd(&dVs[si * d + di], pr * ow[di]);
            pw[si] = pr;
        }
        rowdot = warp_sum(rowdot);
        cnt = 0;
        for (int si = lane; si < a.s; si += 32, ++cnt) {
            const float ds = pw[si] * (dp_local[cnt] - rowdot);
            pw[si] = ds;
            // dK[s] += dS * (scaled q)   (scores = (q*scaling) . k)
            for (int di = 0; di < d; ++di) atomicAdd(&dKs[si * d + di], ds * qw[di]);
        }
        __syncwarp();
        // dq = scaling * sum_s dS[s] * k[s]
        for (int di = lane; di < d; di += 32) {
            float acc = 0.f;
            for (int si = 0; si < a.s; ++si) acc = fmaf(pw[si], Ks[si * (d + 1) + di], acc);
            store_from_f(a.dq, a.dtype, ((size_t)b * a.l + qi) * a.dqs + hd * d + di, acc * scaling);
        }
        __syncwarp();
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.s * d; i += blockDim.x) {
        const int si = i / d, di = i - si * d;
        store_from_f(a.dk, a.dtype, ((size_t)b * a.s + si) * a.dks + hd * d + di, dKs[i]);
        store_from_f(a.dv, a.dtype, ((size_t)b * a.s + si) * a.dvs + hd * d + di, dVs[i]);
    }
}

// Block form for the UNet's tiny cross-attention (L, S <= 16 tokens, d = 64 / 128): one CTA per (sample, head) holds Q,
// K, V and dO in shared memory (16-byte staging loads), one thread per (query, key) pair for the scores and dP, one warp
// per softmax row, one thread per element of dV = P^T dO, dK = dS^T Qs and dQ = scaling * dS K: no atomics, nothing
// serialised over the head dimension.  The warp-per-query kernel above stays for shapes that do not fit.
constexpr int AB2_THREADS = 256;
template <typename T>
__device__ __forceinline__ void ab_stage8(float* dst, const T* src, float scale);
template <>
__device__ __forceinline__ void ab_stage8<__nv_bfloat16>(float* dst, const __nv_bfloat16* src, float scale) {
    const uint4 r = __ldg(reinterpret_cast<const uint4*>(src));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        dst[2 * e] = f.x * scale; dst[2 * e + 1] = f.y * scale;
    }
}
template <>
__device__ __forceinline__ void ab_stage8<float>(float* dst, const float* src, float scale) {
    const float4 r0 = __ldg(reinterpret_cast<const float4*>(src)), r1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
    dst[0] = r0.x * scale; dst[1] = r0.y * scale; dst[2] = r0.z * scale; dst[3] = r0.w * scale;
    dst[4] = r1.x * scale; dst[5] = r1.y * scale; dst[6] = r1.z * scale; dst[7] = r1.w * scale;
}
template <typename T>
__device__ __forceinline__ void ab_store(T* p, float v);
template <> __device__ __forceinline__ void ab_store<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void ab_store<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

__host__ __device__ inline size_t ab2_smem_floats(int l, int s, int d) {
    return (size_t)2 * s * (d + 4) + (size_t)2 * l * (d + 4) + (size_t)2 * l * (s + 1);
}
template <typename T>
__device__ __forceinline__ void ab_store4(T* p, float4 v);
template <> __device__ __forceinline__ void ab_store4<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void ab_store4<__nv_bfloat16>(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = pk;
}
template <typename T>
__global__ void __launch_bounds__(AB2_THREADS) attention_bwd_block_kernel(const AttnBwdP a) {
    extern __shared__ __align__(16) float smv[];
    float* sm = smv;
    const int d = a.e / a.heads, d8 = d >> 3, l = a.l, s = a.s;
    const int dp = d + 4;                            // 16-byte aligned rows, 8 consecutive rows on distinct banks
    float* Ks = sm;                                  // [s][dp]
    float* Vs = Ks + (size_t)s * dp;                 // [s][dp]
    float* Qs = Vs + (size_t)s * dp;                 // [l][dp]  scaled q
    float* Os = Qs + (size_t)l * dp;                 // [l][dp]  d_out
    float* Ps = Os + (size_t)l * dp;                 // [l][s+1]  scores -> probabilities
    float* Ds = Ps + (size_t)l * (s + 1);            // [l][s+1]  dP -> dS
    const int b = blockIdx.x / a.heads, hd = blockIdx.x - b * a.heads;
    const float scaling = 1.0f / sqrtf((float)d);
    const T* kp = reinterpret_cast<const T*>(a.k) + (size_t)b * s * a.ks + hd * d;
    const T* vp = reinterpret_cast<const T*>(a.v) + (size_t)b * s * a.vs + hd * d;
    const T* qp = reinterpret_cast<const T*>(a.q) + (size_t)b * l * a.qs + hd * d;
    const T* op = reinterpret_cast<const T*>(a.d_out) + (size_t)b * l * a.dos + hd * d;
    const int nk = s * d8, nq = l * d8;
    for (int i = threadIdx.x; i < 2 * nk + 2 * nq; i += AB2_THREADS) {
        float f[8];
        float* dst;
        if (i < nk) {
            const int r = i / d8, c = (i - r * d8) * 8;
            ab_stage8<T>(f, kp + (size_t)r * a.ks + c, 1.f); dst = Ks + r * dp + c;
        } else if (i < 2 * nk) {
            const int ii = i - nk, r = ii / d8, c = (ii - r * d8) * 8;
            ab_stage8<T>(f, vp + (size_t)r * a.vs + c, 1.f); dst = Vs + r * dp + c;
        } else if (i < 2 * nk + nq) {
            const int ii = i - 2 * nk, r = ii / d8, c = (ii - r * d8) * 8;
            ab_stage8<T>(f, qp + (size_t)r * a.qs + c, scaling); dst = Qs + r * dp + c;
        } else {
            const int ii = i - 2 * nk - nq, r = ii / d8, c = (ii - r * d8) * 8;
            ab_stage8<T>(f, op + (size_t)r * a.dos + c, 1.f); dst = Os + r * dp + c;
        }
        *reinterpret_cast<float4*>(dst) = make_float4(f[0], f[1], f[2], f[3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(f[4], f[5], f[6], f[7]);
    }
    __syncthreads();
    // scores and dP = dO V^T (128-bit shared-memory reads: the kernel is bound by LDS instructions)
    for (int pr = threadIdx.x; pr < l * s; pr += AB2_THREADS) {
        const int qi = pr / s, si = pr - qi * s;
        const float4* qr = reinterpret_cast<const float4*>(Qs + qi * dp);
        const float4* orow = reinterpret_cast<const float4*>(Os + qi * dp);
        const float4* kr = reinterpret_cast<const float4*>(Ks + si * dp);
        const float4* vr = reinterpret_cast<const float4*>(Vs + si * dp);
        float a1 = 0.f, a2 = 0.f;
#pragma unroll 4
        for (int c = 0; c < (d >> 2); ++c) {
            const float4 qv = qr[c], kv = kr[c], ov = orow[c], vv = vr[c];
            a1 = fmaf(qv.x, kv.x, a1); a1 = fmaf(qv.y, kv.y, a1); a1 = fmaf(qv.z, kv.z, a1); a1 = fmaf(qv.w, kv.w, a1);
            a2 = fmaf(ov.x, vv.x, a2); a2 = fmaf(ov.y, vv.y, a2); a2 = fmaf(ov.z, vv.z, a2); a2 = fmaf(ov.w, vv.w, a2);
        }
        Ps[qi * (s + 1) + si] = a1;
        Ds[qi * (s + 1) + si] = a2;
    }
    __syncthreads();
    // softmax rows, rowdot = sum p dP, dS = p (dP - rowdot)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int qi = warp; qi < l; qi += AB2_THREADS / 32) {
        float* pw = Ps + qi * (s + 1);
        float* dw = Ds + qi * (s + 1);
        float mx = -INFINITY;
        for (int si = lane; si < s; si += 32) mx = fmaxf(mx, pw[si]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int si = lane; si < s; si += 32) { const float ev = expf(pw[si] - mx); pw[si] = ev; sum += ev; }
        sum = warp_sum(sum);
        const float inv = 1.f / sum;
        float rowdot = 0.f;
        for (int si = lane; si < s; si += 32) { const float pr = pw[si] * inv; pw[si] = pr; rowdot = fmaf(pr, dw[si], rowdot); }
        rowdot = warp_sum(rowdot);
        for (int si = lane; si < s; si += 32) dw[si] = pw[si] * (dw[si] - rowdot);
    }
    __syncthreads();
    T* dkp = reinterpret_cast<T*>(a.dk) + (size_t)b * s * a.dks + hd * d;
    T* dvp = reinterpret_cast<T*>(a.dv) + (size_t)b * s * a.dvs + hd * d;
    T* dqp = reinterpret_cast<T*>(a.dq) + (size_t)b * l * a.dqs + hd * d;
    const int d4 = d >> 2, nkv = s * d4, nqd = l * d4;
    for (int o = threadIdx.x; o < 2 * nkv + nqd; o += AB2_THREADS) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (o < nkv) {                       // dV[si][di..] = sum_q P[q][si] dO[q][di..]
            const int si = o / d4, di = (o - si * d4) * 4;
            for (int qi = 0; qi < l; ++qi) {
                const float w = Ps[qi * (s + 1) + si];
                const float4 x = *reinterpret_cast<const float4*>(Os + qi * dp + di);
                acc.x = fmaf(w, x.x, acc.x); acc.y = fmaf(w, x.y, acc.y); acc.z = fmaf(w, x.z, acc.z); acc.w = fmaf(w, x.w, acc.w);
            }
            ab_store4<T>(dvp + (size_t)si * a.dvs + di, acc);
        } else if (o < 2 * nkv) {            // dK[si][di..] = sum_q dS[q][si] Qs[q][di..]   (scores = (q*scaling) . k)
            const int oo = o - nkv, si = oo / d4, di = (oo - si * d4) * 4;
            for (int qi = 0; qi < l; ++qi) {
                const float w = Ds[qi * (s + 1) + si];
                const float4 x = *reinterpret_cast<const float4*>(Qs + qi * dp + di);
                acc.x = fmaf(w, x.x, acc.x); acc.y = fmaf(w, x.y, acc.y); acc.z = fmaf(w, x.z, acc.z); acc.w = fmaf(w, x.w, acc.w);
            }
            ab_store4<T>(dkp + (size_t)si * a.dks + di, acc);
        } else {                             // dQ[qi][di..] = scaling * sum_s dS[qi][s] K[s][di..]
            const int oo = o - 2 * nkv, qi = oo / d4, di = (oo - qi * d4) * 4;
            for (int si = 0; si < s; ++si) {
                const float w = Ds[qi * (s + 1) + si];
                const float4 x = *reinterpret_cast<const float4*>(Ks + si * dp + di);
                acc.x = fmaf(w, x.x, acc.x); acc.y = fmaf(w, x.y, acc.y); acc.z = fmaf(w, x.z, acc.z); acc.w = fmaf(w, x.w, acc.w);
            }
            ab_store4<T>(dqp + (size_t)qi * a.dqs + di, make_float4(acc.x * scaling, acc.y * scaling, acc.z * scaling, acc.w * scaling));
        }
    }
}

// ---------------------------------------------------------------------------
// small streaming ops
// ---------------------------------------------------------------------------
__global__ void sum_hw_kernel(const void* __restrict__ a, const void* __restrict__ a2, int dtype, long long hw, int c, float* __restrict__ out) {
    // grid (ceil(c/blockDim.x), b); thread = channel, loop over pixels (coalesced across channels)
    const int ch = blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= c) return;
    const size_t base = (size_t)blockIdx.y * hw * c + ch;
    float acc = 0.f;
    for (long long p = 0; p < hw; ++p) {
        acc += load_as_f(a, dtype, base + (size_t)p * c);
        if (a2) acc += load_as_f(a2, dtype, base + (size_t)p * c);
    }
    out[(size_t)blockIdx.y * c + ch] = acc;
}
__global__ void gelu_bwd_kernel(const float* __restrict__ dh, const float* __restrict__ pre, float* __restrict__ dpre, size_t count) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        const float x = pre[i];
        const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752440f));
        const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
        dpre[i] = dh[i] * (cdf + x * pdf);
    }
}
__global__ void add_bcast_kernel(const void* __restrict__ y, const void* __restrict__ add, const float* __restrict__ add_nc, void* out,
                                 long long hw, int c, int dtype, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        float v = load_as_f(y, dtype, i);
        if (add) v += load_as_f(add, dtype, i);
        if (add_nc) {
            if (total < (1ull << 32)) {
                const unsigned i32 = (unsigned)i, per = (unsigned)(hw * c);
                v += add_nc[(i32 / per) * (unsigned)c + (i32 % (unsigned)c)];
            } else {
                v += add_nc[(i / ((size_t)hw * c)) * c + (i % c)];
            }
        }
        store_from_f(out, dtype, i, v);
    }
}
// x *= *factor (the upstream scalar of a loss's autograd node stays on the device: no host sync in loss.backward())
__global__ void scale_dev_kernel(float* __restrict__ x, const float* __restrict__ factor, size_t count) {
    const float f = __ldg(factor);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) x[i] *= f;
}

__global__ void mse_grad_kernel(const float* __restrict__ a, const float* __restrict__ b, float scale, void* g, int g_dtype, size_t count) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x)
        store_from_f(g, g_dtype, i, scale * (a[i] - b[i]));
}
// d/dz of w * mean(0.5 (z^2 - 1 - log(z^2 + 1e-8)))  (loss.py:31-32) = scale * (z - z / (z^2 + 1e-8)), scale = w / count;
// added to the gradient arriving from the decoder (optional)
__global__ void kl_grad_kernel(const float* __restrict__ z, const void* __restrict__ dz, int dz_dtype, float scale, float* __restrict__ g,
                               size_t count) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        const float zz = z[i];
        const float k = scale * (zz - zz / (zz * zz + 1e-8f));
        g[i] = dz ? load_as_f(dz, dz_dtype, i) + k : k;
    }
}
__global__ void eps_grad_kernel(const void* __restrict__ dx0, int dx0_dtype, const float* __restrict__ eh, const float* __restrict__ eps,
                                const float* __restrict__ ab, const long long* __restrict__ t, float scale_d, void* g, int g_dtype,
                                long long inner, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const float abv = ab[t[i / (size_t)inner]];
        const float coef = -sqrtf(1.f - abv) / sqrtf(abv);
        store_from_f(g, g_dtype, i, scale_d * (eh[i] - eps[i]) + coef * load_as_f(dx0, dx0_dtype, i));
    }
}
// torch.optim.Adam single-tensor semantics: exp_avg.lerp_(g, 1-b1); exp_avg_sq = b2*v + (1-b2) g^2;
// denom = sqrt(v)/sqrt(bc2) + eps; p -= (lr/bc1) * m/denom
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                                                   size_t count, float lr, float b1, float b2, float eps, float bc1, float bc2_sqrt, float gs,
                                                   const long long* __restrict__ step_dev, float decay) {
    if (step_dev) {
        // step counter in device memory (CUDA-graph replays): bias corrections computed here, in double as on the host
        const double s = (double)step_dev[0];
        bc1 = (float)(1.0 - pow((double)b1, s));
        bc2_sqrt = (float)sqrt(1.0 - pow((double)b2, s));
    }
    const size_t nvec = count >> 2;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
        const float4 gg = reinterpret_cast<const float4*>(g)[i];
        float* pe = reinterpret_cast<float*>(&pp); float* me = reinterpret_cast<float*>(&mm); float* ve = reinterpret_cast<float*>(&vv);
        const float* ge = reinterpret_cast<const float*>(&gg);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float gr = ge[k] * gs;
            me[k] = me[k] + (gr - me[k]) * (1.f - b1);
            ve[k] = ve[k] * b2 + (1.f - b2) * gr * gr;
            const float denom = sqrtf(ve[k]) / bc2_sqrt + eps;
            pe[k] = pe[k] * decay - (lr / bc1) * (me[k] / denom);
        }
        reinterpret_cast<float4*>(p)[i] = pp; reinterpret_cast<float4*>(m)[i] = mm; reinterpret_cast<float4*>(v)[i] = vv;
    }
    for (size_t i = (nvec << 2) + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        const float gr = g[i] * gs;
        const float mn = m[i] + (gr - m[i]) * (1.f - b1);
        const float vn = v[i] * b2 + (1.f - b2) * gr * gr;
        m[i] = mn; v[i] = vn;
        p[i] = p[i] * decay - (lr / bc1) * (mn / (sqrtf(vn) / bc2_sqrt + eps));
    }
}
// dst [k*k][cin][cout] = w[cout][cin][k-1-ky][k-1-kx]
__global__ void pack_dgrad_kernel(const float* __restrict__ src, void* dst, int cout, int cin, int k, int dtype) {
    const int kk = k * k;
    const size_t total = (size_t)kk * cout * cin;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int co = (int)(i % cout);
        const int ci = (int)((i / cout) % cin);
        const int s = (int)(i / ((size_t)cin * cout));
        store_from_f(dst, dtype, i, src[((size_t)co * cin + ci) * kk + (kk - 1 - s)]);
    }
}

// Every packed copy of the trainable weights of a training step in ONE launch (the step re-packed each weight with its
// own 3-5 us kernel right in front of the convolution that reads it: 52 launches, and a weight prefetch ahead of
// griddepcontrol.wait that relied on the pack kernel's exit).  Table in device memory; block b serves the entry whose
// [first_block, first_block + blocks) range holds it, PACK_PER_BLOCK elements each.
constexpr int PACK_PER_BLOCK = 2048;
__global__ void __launch_bounds__(256) pack_multi_kernel(const mstk_pack_entry* __restrict__ tab, int n) {
    __shared__ int s_e;
    if (threadIdx.x < 32) {
        // entry of this block = the last one whose first_block <= blockIdx.x: one coalesced probe of the table per 32 entries
        // (a serial scan by one thread was ~25 dependent global loads per block)
        int e = 0;
        for (int base = 0; base < n; base += 32) {
            const int i = base + (int)threadIdx.x;
            const bool le = i < n && tab[i].first_block <= (int)blockIdx.x;
            e += __popc(__ballot_sync(0xffffffffu, le));
        }
        if (threadIdx.x == 0) s_e = e - 1;
    }
    __syncthreads();
    const mstk_pack_entry t = tab[s_e];
    const uint32_t total = (uint32_t)t.kk * (uint32_t)t.cout * (uint32_t)t.cin;
    const uint32_t i0 = (uint32_t)((int)blockIdx.x - t.first_block) * PACK_PER_BLOCK;
    const uint32_t i1 = i0 + PACK_PER_BLOCK < total ? i0 + PACK_PER_BLOCK : total;
    const float* __restrict__ src = t.src;
    const uint32_t kk = (uint32_t)t.kk, cin = (uint32_t)t.cin, cout = (uint32_t)t.cout;
#pragma unroll 4
    for (uint32_t i = i0 + threadIdx.x; i < i1; i += 256) {
        uint32_t si;
        if (t.mode == 2) {              // dgrad: dst [kk][cin][cout], taps flipped
            const uint32_t co = i % cout, r = i / cout;
            const uint32_t ci = r % cin, sl = r / cin;
            si = (co * cin + ci) * kk + (kk - 1 - sl);
        } else {                        // dst [kk][cout][cin]; src [cout][cin][kk] (0) or [cin][cout][kk] (1)
            const uint32_t ci = i % cin, r = i / cin;
            const uint32_t co = r % cout, sl = r / cout;
            si = t.mode == 1 ? (ci * cout + co) * kk + sl : (co * cin + ci) * kk + sl;
        }
        store_from_f(t.dst, t.dtype, i, __ldg(src + si));
    }
}

// ---------------------------------------------------------------------------
// weight gradient of the single-channel edge layers (HBM-streaming, CUDA cores):
//   Conv2d(1, C, k3, stride s, p1):          V = dy [N,Hv,Wv,C], S = x  [N,Hs,Ws]
//   ConvTranspose2d(C, 1, k4, s2, p1):       V = x  [N,Hv,Wv,C], S = dy [N,Hs,Ws]
// Both are  dw[c][ky][kx] += sum_{n,vh,vw} V[n,vh,vw,c] * S[n, vh*ss + ky - 1, vw*ss + kx - 1]  (zero outside S).
// The vector operand is read exactly once (CH channels = one 8/16-byte load per thread and pixel), the
// scalar window comes from L1; K*K*CH accumulators per thread, shared-memory then global RED at the end.
// ---------------------------------------------------------------------------
struct EdgeWgradP {
    const __nv_bfloat16* v;
    const void* s;
    float* dw;
    int s_dtype;
    int N, Hv, Wv, C, Hs, Ws, ss;
};

template <int K, int CH, bool SF32>
__global__ void __launch_bounds__(256) wgrad_edge_kernel(const EdgeWgradP q) {
    extern __shared__ float acc_s[];                  // [C][K*K]
    const int tpp = q.C / CH;                         // threads per pixel
    const int ppb = 256 / tpp;                        // pixels per block iteration
    const int cg = threadIdx.x % tpp;                 // channel group
    const int slot = threadIdx.x / tpp;
    for (int i = threadIdx.x; i < q.C * K * K; i += 256) acc_s[i] = 0.f;
    __syncthreads();
    float acc[K * K][CH];
#pragma unroll
    for (int t = 0; t < K * K; ++t)
#pragma unroll
        for (int c = 0; c < CH; ++c) acc[t][c] = 0.f;
    const long long npix = (long long)q.N * q.Hv * q.Wv;
    if (slot < ppb) {
        for (long long pix = (long long)blockIdx.x * ppb + slot; pix < npix; pix += (long long)gridDim.x * ppb) {
            int n, r;
            if (npix < (1ll << 31)) {          // 32-bit index arithmetic whenever it fits (a 64-bit division is ~100 instructions)
                const unsigned p32 = (unsigned)pix, hw = (unsigned)(q.Hv * q.Wv);
                n = (int)(p32 / hw);
                r = (int)(p32 - (unsigned)n * hw);
            } else {
                n = (int)(pix / ((long long)q.Hv * q.Wv));
                r = (int)(pix - (long long)n * q.Hv * q.Wv);
            }
            const int vh = (int)((unsigned)r / (unsigned)q.Wv), vw = r - vh * q.Wv;
            float vv[CH];
            const __nv_bfloat16* vp = q.v + pix * q.C + cg * CH;
            if (CH == 8) {
                const uint4 u = *reinterpret_cast<const uint4*>(vp);
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(h[e]); vv[2 * e] = f.x; vv[2 * e + 1] = f.y; }
            } else {
                const uint2 u = *reinterpret_cast<const uint2*>(vp);
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
                for (int e = 0; e < 2; ++e) { const float2 f = __bfloat1622float2(h[e]); vv[2 * e] = f.x; vv[2 * e + 1] = f.y; }
            }
            const int sh0 = vh * q.ss - 1, sw0 = vw * q.ss - 1;
            const size_t sbase = (size_t)n * q.Hs * q.Ws;
            // the K*K scalar window FIRST, as independent branch-free loads (clamped address, zero by select): with a
            // branch per tap the compiler kept load -> FMA -> load order and every tap paid a serial L2 round trip
            float sv[K * K];
#pragma unroll
            for (int ky = 0; ky < K; ++ky) {
                const int sh = sh0 + ky;
                const int shc = min(max(sh, 0), q.Hs - 1);
#pragma unroll
                for (int kx = 0; kx < K; ++kx) {
                    const int sw = sw0 + kx;
                    const int swc = min(max(sw, 0), q.Ws - 1);
                    const size_t si = sbase + (size_t)shc * q.Ws + swc;
                    const float raw = SF32 ? __ldg(reinterpret_cast<const float*>(q.s) + si)
                                           : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(q.s)[si]);
                    sv[ky * K + kx] = (sh == shc && sw == swc) ? raw : 0.f;
                }
            }
#pragma unroll
            for (int t = 0; t < K * K; ++t)
#pragma unroll
                for (int c = 0; c < CH; ++c) acc[t][c] = fmaf(vv[c], sv[t], acc[t][c]);
        }
#pragma unroll
        for (int t = 0; t < K * K; ++t)
#pragma unroll
            for (int c = 0; c < CH; ++c) atomicAdd(&acc_s[(cg * CH + c) * K * K + t], acc[t][c]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < q.C * K * K; i += 256) atomicAdd(q.dw + i, acc_s[i]);
}

// ---------------------------------------------------------------------------
// wgrad_edge_mma_kernel: the ConvTranspose2d(64, 1, k4, s2, p1) weight gradient (decoder last layer) as a skinny GEMM on mma.sync m16n8k16:
//   dw^T[tap][c] = sum_pix S_tap[pix] * V[pix][c]          M = 16 taps, N = 64 channels, K = pixels
// The CUDA-core kernel above spends ~200 instructions per (pixel, 4 channels) on window addressing that all 16 threads
// of a pixel repeat (337 us for 256 x 64 x 64 x 64, 0.4 TB/s).  Here a warp takes 16 consecutive pixels of a row per step:
//   B (V, bf16): lane (g, q) loads the 8 CONSECUTIVE channels [8g, 8g+8) of pixels 2q, 2q+1, 2q+8, 2q+9 (four 16-byte loads,
//      a quad-row of lanes covers whole 128-byte pixels) and interleaves pixel pairs with PRMT; n-tile j, column g <-> channel 8g+j
//   A (S = dy, fp32 or bf16): 8 gathered scalars per lane, split into bf16 hi + lo parts (two MMAs) so fp32 dy keeps ~16 bits
// 32 fp32 accumulators per lane; warps -> shared memory -> one global RED per weight and CTA.  V is read exactly once.
// ---------------------------------------------------------------------------
// The same kernel serves Conv2d(1, 64, k3, stride s, p1) (first layer of the style encoder / VAE encoder): K = 3, nine taps in
// the 16-row A tile (rows 9..15 are zero), V = dy, S = x.
template <int K, bool SF32>
__global__ void __launch_bounds__(256) wgrad_edge_mma_kernel(const EdgeWgradP q) {
    constexpr int KK = K * K;
    __shared__ float acc_s[64 * KK];                  // [c][tap] = the parameter's own layout
    for (int i = threadIdx.x; i < 64 * KK; i += 256) acc_s[i] = 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 2, qd = lane & 3;
    const int segs = q.Wv >> 4;
    const long long steps = (long long)q.N * q.Hv * segs;
    const long long wstride = (long long)gridDim.x * 8;
    float c[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j][0] = 0.f; c[j][1] = 0.f; c[j][2] = 0.f; c[j][3] = 0.f; }
    // taps g and g + 8 of the A tile
    const int tky[2] = {g / K, (g + 8) / K}, tkx[2] = {g % K, (g + 8) % K};
    const bool tok[2] = {g < KK, g + 8 < KK};
    for (long long st = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); st < steps; st += wstride) {
        const int seg = (int)(st % segs);
        const long long row = st / segs;
        const int vh = (int)(row % q.Hv), n = (int)(row / q.Hv);
        const int vw0 = seg << 4;
        // ---- V: four pixels x 8 channels ----
        const __nv_bfloat16* vp = q.v + ((row * q.Wv + vw0) * 64 + 8 * g);
        const uint4 p0 = *reinterpret_cast<const uint4*>(vp + (2 * qd) * 64);
        const uint4 p1 = *reinterpret_cast<const uint4*>(vp + (2 * qd + 1) * 64);
        const uint4 p2 = *reinterpret_cast<const uint4*>(vp + (2 * qd + 8) * 64);
        const uint4 p3 = *reinterpret_cast<const uint4*>(vp + (2 * qd + 9) * 64);
        // ---- S: taps (g, g+8) at pixels 2q, 2q+1, 2q+8, 2q+9 ----
        float sv[2][4];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int sh = q.ss * vh + tky[half] - 1;
            const bool rok = tok[half] && sh >= 0 && sh < q.Hs;
            const size_t rbase = ((size_t)n * q.Hs + (rok ? sh : 0)) * q.Ws;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int k = 2 * qd + (e & 1) + 8 * (e >> 1);
                const int sw = q.ss * (vw0 + k) + tkx[half] - 1;
                const bool ok = rok && sw >= 0 && sw < q.Ws;
                const size_t si = rbase + (ok ? sw : 0);
                const float raw = SF32 ? __ldg(reinterpret_cast<const float*>(q.s) + si)
                                       : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(q.s)[si]);
                sv[half][e] = ok ? raw : 0.f;
            }
        }
        unsigned ah[4], al[4];                          // a0: (tap g, k 2q..), a1: (tap g+8, k 2q..), a2: (tap g, k 2q+8..), a3: (tap g+8, k 2q+8..)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int half = r & 1, e0 = (r >> 1) * 2;
            const __nv_bfloat162 hi = __floats2bfloat162_rn(sv[half][e0], sv[half][e0 + 1]);
            ah[r] = *reinterpret_cast<const unsigned*>(&hi);
            if (SF32) {
                const float2 hf = __bfloat1622float2(hi);
                const __nv_bfloat162 lo = __floats2bfloat162_rn(sv[half][e0] - hf.x, sv[half][e0 + 1] - hf.y);
                al[r] = *reinterpret_cast<const unsigned*>(&lo);
            }
        }
        const unsigned* w0 = reinterpret_cast<const unsigned*>(&p0);
        const unsigned* w1 = reinterpret_cast<const unsigned*>(&p1);
        const unsigned* w2 = reinterpret_cast<const unsigned*>(&p2);
        const unsigned* w3 = reinterpret_cast<const unsigned*>(&p3);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            // channel 8g + j of the pixel pairs: low halves for even j, high halves for odd j
            const unsigned sel = (j & 1) ? 0x7632u : 0x5410u;
            const unsigned b0 = __byte_perm(w0[j >> 1], w1[j >> 1], sel);
            const unsigned b1 = __byte_perm(w2[j >> 1], w3[j >> 1], sel);
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])
                         : "r"(ah[0]), "r"(ah[1]), "r"(ah[2]), "r"(ah[3]), "r"(b0), "r"(b1));
            if (SF32)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])
                             : "r"(al[0]), "r"(al[1]), "r"(al[2]), "r"(al[3]), "r"(b0), "r"(b1));
        }
    }
    // c[j][0..1]: tap g, columns 2q, 2q+1 of n-tile j -> channels 8*(2q) + j, 8*(2q+1) + j; c[j][2..3]: tap g + 8
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (tok[0]) {
            atomicAdd(&acc_s[(16 * qd + j) * KK + g], c[j][0]);
            atomicAdd(&acc_s[(16 * qd + 8 + j) * KK + g], c[j][1]);
        }
        if (tok[1]) {
            atomicAdd(&acc_s[(16 * qd + j) * KK + g + 8], c[j][2]);
            atomicAdd(&acc_s[(16 * qd + 8 + j) * KK + g + 8], c[j][3]);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * KK; i += 256) atomicAdd(q.dw + i, acc_s[i]);
}

// returns MSTK_EUNSUPPORTED when the layer is not one of the two edge shapes
static int launch_wgrad_edge(const mstk_wgrad_args* a, cudaStream_t st) {
    EdgeWgradP q;
    memset(&q, 0, sizeof(q));
    q.dw = a->dw; q.N = a->n;
    if (!a->transposed && a->cin == 1 && a->ksize == 3 && a->dy_dtype == MSTK_BF16 && (a->cout % 8) == 0 && a->cout <= 256 &&
        (256 % (a->cout / 8)) == 0) {
        q.v = reinterpret_cast<const __nv_bfloat16*>(a->dy); q.s = a->x; q.s_dtype = a->x_dtype;
        q.Hv = (a->h + 2 - 3) / a->stride + 1; q.Wv = (a->w + 2 - 3) / a->stride + 1; q.C = a->cout;
        q.Hs = a->h; q.Ws = a->w; q.ss = a->stride;
        const long long npix = (long long)q.N * q.Hv * q.Wv;
        static const bool no_mma3 = getenv("MSTK_WGRAD_EDGE_WALK") != nullptr;
        if (q.C == 64 && (q.Wv & 15) == 0 && !no_mma3 && (reinterpret_cast<uintptr_t>(q.v) & 15) == 0) {
            long long mb = (npix / 16 + 7) / 8;
            if (mb > 4LL * bw_sms()) mb = 4LL * bw_sms();
            if (q.s_dtype == MSTK_F32) wgrad_edge_mma_kernel<3, true><<<(unsigned)mb, 256, 0, st>>>(q);
            else wgrad_edge_mma_kernel<3, false><<<(unsigned)mb, 256, 0, st>>>(q);
            MSTK_CHECK_LAUNCH("wgrad_edge_mma(cin=1)");
            return MSTK_OK;
        }
        const int ppb = 256 / (q.C / 8);
        long long blocks = (npix + ppb - 1) / ppb;
        if (blocks > 4LL * bw_sms()) blocks = 4LL * bw_sms();
        if (q.s_dtype == MSTK_F32) wgrad_edge_kernel<3, 8, true><<<(unsigned)blocks, 256, (size_t)q.C * 9 * sizeof(float), st>>>(q);
        else wgrad_edge_kernel<3, 8, false><<<(unsigned)blocks, 256, (size_t)q.C * 9 * sizeof(float), st>>>(q);
        MSTK_CHECK_LAUNCH("wgrad_edge(cin=1)");
        return MSTK_OK;
    }
    if (a->transposed && a->cout == 1 && a->ksize == 4 && a->stride == 2 && a->x_dtype == MSTK_BF16 && (a->cin % 4) == 0 &&
        a->cin <= 512 && (256 % (a->cin / 4)) == 0) {
        q.v = reinterpret_cast<const __nv_bfloat16*>(a->x); q.s = a->dy; q.s_dtype = a->dy_dtype;
        q.Hv = a->h; q.Wv = a->w; q.C = a->cin; q.Hs = 2 * a->h; q.Ws = 2 * a->w; q.ss = 2;
        const long long npix = (long long)q.N * q.Hv * q.Wv;
        static const bool no_mma = getenv("MSTK_WGRAD_EDGE_WALK") != nullptr;      // A/B switch: the CUDA-core kernel for every shape
        if (q.C == 64 && (q.Wv & 15) == 0 && !no_mma && (reinterpret_cast<uintptr_t>(q.v) & 15) == 0) {
            long long mb = (npix / 16 + 7) / 8;
            if (mb > 4LL * bw_sms()) mb = 4LL * bw_sms();
            if (q.s_dtype == MSTK_F32) wgrad_edge_mma_kernel<4, true><<<(unsigned)mb, 256, 0, st>>>(q);
            else wgrad_edge_mma_kernel<4, false><<<(unsigned)mb, 256, 0, st>>>(q);
            MSTK_CHECK_LAUNCH("wgrad_edge_mma(cout=1)");
            return MSTK_OK;
        }
        const int ppb = 256 / (q.C / 4);
        long long blocks = (npix + ppb - 1) / ppb;
        if (blocks > 4LL * bw_sms()) blocks = 4LL * bw_sms();
        if (q.s_dtype == MSTK_F32) wgrad_edge_kernel<4, 4, true><<<(unsigned)blocks, 256, (size_t)q.C * 16 * sizeof(float), st>>>(q);
        else wgrad_edge_kernel<4, 4, false><<<(unsigned)blocks, 256, (size_t)q.C * 16 * sizeof(float), st>>>(q);
        MSTK_CHECK_LAUNCH("wgrad_edge(cout=1)");
        return MSTK_OK;
    }
    return MSTK_EUNSUPPORTED;
}

}  // namespace mstk

using namespace mstk;

extern "C" {

int mstk_pack_conv_weight_dgrad(const float* w, void* dst, int cout, int cin, int k, int dtype, void* stream) {
    MSTK_REQUIRE(w && dst && cout > 0 && cin > 0 && (k == 1 || k == 3), "pack_conv_weight_dgrad: bad args");
    const size_t total = (size_t)k * k * cout * cin;
    pack_dgrad_kernel<<<bw_grid(total, 256), 256, 0, (cudaStream_t)stream>>>(w, dst, cout, cin, k, dtype);
    MSTK_CHECK_LAUNCH("pack_conv_weight_dgrad");
    return MSTK_OK;
}

int mstk_pack_blocks(int cout, int cin, int kk) {
    if (cout <= 0 || cin <= 0 || kk <= 0 || (long long)kk * cout * cin >= (1ll << 31)) return 0;
    return (int)(((size_t)kk * cout * cin + PACK_PER_BLOCK - 1) / PACK_PER_BLOCK);
}

int mstk_pack_multi(const mstk_pack_entry* entries_dev, int n_entries, int total_blocks, void* stream) {
    MSTK_REQUIRE(entries_dev && n_entries > 0 && total_blocks > 0, "pack_multi: bad args");
    pack_multi_kernel<<<(unsigned)total_blocks, 256, 0, (cudaStream_t)stream>>>(entries_dev, n_entries);
    MSTK_CHECK_LAUNCH("pack_multi");
    return MSTK_OK;
}

int mstk_conv2d_wgrad(const mstk_wgrad_args* a, void* stream) {
    MSTK_REQUIRE(a && a->x && a->dy && a->dw, "wgrad: null");
    MSTK_REQUIRE(aligned16({a->x, a->dy, a->dw}), "wgrad: x / dy / dw must be 16-byte aligned (TMA boxes, 128-bit loads)");
    // build the FORWARD plan (taps / phases); x stays the input, "y" is dy
    mstk_conv_args f;
    memset(&f, 0, sizeof(f));
    f.x = a->x; f.wpacked = a->x /* unused */; f.y = const_cast<void*>(a->dy);
    f.n = a->n; f.h = a->h; f.w = a->w; f.cin = a->cin; f.cout = a->cout; f.ksize = a->ksize; f.stride = a->stride;
    f.transposed = a->transposed; f.act = MSTK_ACT_NONE; f.out_scale = 1.f; f.out_shift = 0.f;
    f.dtype = a->x_dtype; f.out_dtype = a->dy_dtype;
    {
        const int rce = launch_wgrad_edge(a, (cudaStream_t)stream);
        if (rce != MSTK_EUNSUPPORTED) return rce;
    }
    WgradP q;
    int rc = build_conv_plan(&f, &q.p);
    if (rc != MSTK_OK) return rc;
    MSTK_REQUIRE((reinterpret_cast<uintptr_t>(a->x) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->dy) & 15) == 0, "wgrad: operands must be 16-byte aligned");
    q.dw = a->dw; q.x_dtype = a->x_dtype; q.dy_dtype = a->dy_dtype;
    const int kk = a->ksize * a->ksize;
    if (a->transposed) { q.s_ci = (long long)a->cout * kk; q.s_co = kk; }
    else { q.s_co = (long long)a->cin * kk; q.s_ci = kk; }
    if (wgrad_tc_supported(q.p, a->x_dtype, a->dy_dtype)) return launch_wgrad_tc(q.p, a->dw, q.s_co, q.s_ci, (cudaStream_t)stream);
    q.npt = 0;
    for (int i = 0; i < q.p.nphases; ++i) q.npt += q.p.ph[i].ntaps;
    const long long K = (long long)q.p.N * q.p.Hg * q.p.Wg;
    const long long tiles = (long long)((a->cout + WG_BM - 1) / WG_BM) * ((a->cin + WG_BN - 1) / WG_BN) * q.npt;
    long long ks = (4LL * bw_sms() + tiles - 1) / tiles;          // aim at ~4 CTAs per SM
    const long long ks_max = (K + 4 * WG_BK - 1) / (4 * WG_BK);    // at least 4 k-chunks per split
    if (ks > ks_max) ks = ks_max;
    if (ks < 1) ks = 1;
    MSTK_REQUIRE((long long)q.npt * ks <= 65535, "wgrad: grid too large");
    q.ksplit = (int)ks;
    dim3 grid((a->cout + WG_BM - 1) / WG_BM, (a->cin + WG_BN - 1) / WG_BN, (unsigned)(q.npt * ks));
    wgrad_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(q);
    MSTK_CHECK_LAUNCH("wgrad");
    return MSTK_OK;
}

int mstk_act_bwd(const mstk_act_bwd_args* a, void* stream) {
    MSTK_REQUIRE(a && a->dy && (a->g || a->dbias), "act_bwd: null");
    MSTK_REQUIRE(a->act == MSTK_ACT_NONE || a->y, "act_bwd: y needed for the activation derivative");
    MSTK_REQUIRE(a->rows > 0 && a->c > 0 && a->c <= 8192, "act_bwd: bad dims");
    MSTK_REQUIRE(a->y_scale != 0.f, "act_bwd: y_scale must be non-zero");
    ActBwdP p;
    p.dy = a->dy; p.dy2 = a->dy2; p.y = a->y; p.g = a->g; p.dbias = a->dbias; p.rows = a->rows; p.c = a->c; p.act = a->act;
    p.y_scale = a->y_scale; p.y_shift = a->y_shift; p.dy_dtype = a->dy_dtype; p.y_dtype = a->y_dtype; p.g_dtype = a->g_dtype;
    const size_t total = (size_t)a->rows * a->c;
    const bool al = aligned16({a->dy, a->dy2, a->y, a->g});
    p.vec = !al ? 0 : (a->c & 7) == 0 ? 8 : (a->c == 1 && (total & 3) == 0) ? 4 : 0;
    const size_t items = p.vec ? total / p.vec : total;
    act_bwd_kernel<<<bw_grid(items, 256, 4), 256, a->dbias ? a->c * sizeof(float) : 0, (cudaStream_t)stream>>>(p);
    MSTK_CHECK_LAUNCH("act_bwd");
    return MSTK_OK;
}

static int bn_bwd_common(const mstk_bn_bwd_args* a, BnBwdP* p) {
    MSTK_REQUIRE(a && a->dy && a->v && a->mean && a->invstd && a->gamma && a->beta && a->sums, "bn_bwd: null");
    MSTK_REQUIRE(a->c > 0 && (a->c & 7) == 0 && a->c <= 4096 && a->rows > 0 && a->count >= a->rows, "bn_bwd: bad dims (c must be a multiple of 8)");
    MSTK_REQUIRE(aligned16({a->dy, a->v, a->dv}), "bn_bwd: dy / v / dv must be 16-byte aligned (128-bit loads and stores)");
    p->dy = a->dy; p->v = a->v; p->dv = a->dv; p->mean = a->mean; p->invstd = a->invstd; p->gamma = a->gamma; p->beta = a->beta;
    p->sums = a->sums; p->dgamma = a->dgamma; p->dbeta = a->dbeta; p->rows = a->rows; p->count = a->count; p->c = a->c; p->act = a->act;
    p->dy_dtype = a->dy_dtype; p->v_dtype = a->v_dtype; p->dv_dtype = a->dv_dtype;
    return MSTK_OK;
}
int mstk_bn_bwd_reduce(const mstk_bn_bwd_args* a, void* stream) {
    BnBwdP p;
    int rc = bn_bwd_common(a, &p);
    if (rc != MSTK_OK) return rc;
    bn_bwd_reduce_kernel<<<bw_grid((size_t)a->rows * a->c / 8, 256, 4), 256, 2 * a->c * sizeof(float), (cudaStream_t)stream>>>(p);
    MSTK_CHECK_LAUNCH("bn_bwd_reduce");
    return MSTK_OK;
}
int mstk_bn_bwd_apply(const mstk_bn_bwd_args* a, void* stream) {
    BnBwdP p;
    int rc = bn_bwd_common(a, &p);
    if (rc != MSTK_OK) return rc;
    MSTK_REQUIRE(a->dv, "bn_bwd_apply: dv is null");
    bn_bwd_apply_kernel<<<bw_grid((size_t)a->rows * a->c / 8, 256, 8), 256, 2 * a->c * sizeof(float), (cudaStream_t)stream>>>(p);
    MSTK_CHECK_LAUNCH("bn_bwd_apply");
    return MSTK_OK;
}

int mstk_attention_bwd(const mstk_attn_bwd_args* a, void* stream) {
    MSTK_REQUIRE(a && a->q && a->k && a->v && a->d_out && a->dq && a->dk && a->dv, "attention_bwd: null");
    MSTK_REQUIRE(a->b > 0 && a->l > 0 && a->s > 0 && a->s <= 128 && a->heads > 0 && a->e % a->heads == 0, "attention_bwd: bad dims (s <= 128)");
    AttnBwdP p;
    p.q = a->q; p.k = a->k; p.v = a->v; p.d_out = a->d_out; p.dq = a->dq; p.dk = a->dk; p.dv = a->dv;
    p.b = a->b; p.l = a->l; p.s = a->s; p.e = a->e; p.heads = a->heads; p.qs = a->q_stride; p.ks = a->k_stride; p.vs = a->v_stride;
    p.dos = a->do_stride; p.dqs = a->dq_stride; p.dks = a->dk_stride; p.dvs = a->dv_stride; p.dtype = a->dtype;
    const int d = a->e / a->heads;
    {
        const int per16 = a->dtype == MSTK_F32 ? 4 : 8;
        const size_t smem2 = ab2_smem_floats(a->l, a->s, d) * sizeof(float);
        const bool aligned = (d % 8) == 0 && (a->q_stride % per16) == 0 && (a->k_stride % per16) == 0 && (a->v_stride % per16) == 0 &&
                             (a->do_stride % per16) == 0 && (a->dq_stride % per16) == 0 && (a->dk_stride % per16) == 0 &&
                             (a->dv_stride % per16) == 0 &&
                             ((reinterpret_cast<uintptr_t>(a->dq) | reinterpret_cast<uintptr_t>(a->dk) | reinterpret_cast<uintptr_t>(a->dv)) & 15) == 0 &&
                             ((reinterpret_cast<uintptr_t>(a->q) | reinterpret_cast<uintptr_t>(a->k) | reinterpret_cast<uintptr_t>(a->v) |
                               reinterpret_cast<uintptr_t>(a->d_out)) & 15) == 0;
        if (aligned && smem2 <= 48 * 1024 && a->l * a->s <= 4096) {
            if (a->dtype == MSTK_F32) attention_bwd_block_kernel<float><<<a->b * a->heads, AB2_THREADS, smem2, (cudaStream_t)stream>>>(p);
            else attention_bwd_block_kernel<__nv_bfloat16><<<a->b * a->heads, AB2_THREADS, smem2, (cudaStream_t)stream>>>(p);
            MSTK_CHECK_LAUNCH("attention_bwd");
            return MSTK_OK;
        }
    }
    const size_t smem = ((size_t)2 * a->s * (d + 1) + (size_t)2 * a->s * d + AB_WARPS * a->s + 2 * AB_WARPS * d) * sizeof(float);
    MSTK_REQUIRE(smem <= 200 * 1024, "attention_bwd: s*d too large (s=%d d=%d)", a->s, d);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "attention_bwd smem attr");
        smem_set = 200 * 1024;
    }
    attention_bwd_kernel<<<a->b * a->heads, AB_WARPS * 32, smem, (cudaStream_t)stream>>>(p);
    MSTK_CHECK_LAUNCH("attention_bwd");
    return MSTK_OK;
}

int mstk_gelu_bwd(const float* dh, const float* pre, float* dpre, size_t count, void* stream) {
    MSTK_REQUIRE(dh && pre && dpre && count > 0, "gelu_bwd: bad args");
    gelu_bwd_kernel<<<bw_grid(count, 256), 256, 0, (cudaStream_t)stream>>>(dh, pre, dpre, count);
    MSTK_CHECK_LAUNCH("gelu_bwd");
    return MSTK_OK;
}
int mstk_sum_hw(const void* a, const void* a2, int dtype, int b, int64_t hw, int c, float* out, void* stream) {
    MSTK_REQUIRE(a && out && b > 0 && b <= 65535 && hw > 0 && c > 0, "sum_hw: bad args");
    dim3 grid((c + 127) / 128, b);
    sum_hw_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(a, a2, dtype, hw, c, out);
    MSTK_CHECK_LAUNCH("sum_hw");
    return MSTK_OK;
}
int mstk_add_bcast(const void* y, const void* add, const float* add_nc, void* out, int n, int64_t hw, int c, int dtype, void* stream) {
    MSTK_REQUIRE(y && out && n > 0 && hw > 0 && c > 0, "add_bcast: bad args");
    const size_t total = (size_t)n * hw * c;
    add_bcast_kernel<<<bw_grid(total, 256), 256, 0, (cudaStream_t)stream>>>(y, add, add_nc, out, hw, c, dtype, total);
    MSTK_CHECK_LAUNCH("add_bcast");
    return MSTK_OK;
}
int mstk_scale_dev(float* x, const float* factor, size_t count, void* stream) {
    MSTK_REQUIRE(x && factor && count > 0, "scale_dev: bad args");
    scale_dev_kernel<<<bw_grid(count, 256), 256, 0, (cudaStream_t)stream>>>(x, factor, count);
    MSTK_CHECK_LAUNCH("scale_dev");
    return MSTK_OK;
}
int mstk_mse_grad(const float* a, const float* b, float scale, void* g, int g_dtype, size_t count, void* stream) {
    MSTK_REQUIRE(a && b && g && count > 0, "mse_grad: bad args");
    mse_grad_kernel<<<bw_grid(count, 256), 256, 0, (cudaStream_t)stream>>>(a, b, scale, g, g_dtype, count);
    MSTK_CHECK_LAUNCH("mse_grad");
    return MSTK_OK;
}
int mstk_eps_grad(const void* dx0, int dx0_dtype, const float* eps_hat, const float* eps, const float* alpha_bar, const int64_t* t,
                  float scale_d, void* g, int g_dtype, int b, int64_t inner, void* stream) {
    MSTK_REQUIRE(dx0 && eps_hat && eps && alpha_bar && t && g && b > 0 && inner > 0, "eps_grad: bad args");
    const size_t total = (size_t)b * inner;
    eps_grad_kernel<<<bw_grid(total, 256), 256, 0, (cudaStream_t)stream>>>(dx0, dx0_dtype, eps_hat, eps, alpha_bar, (const long long*)t, scale_d,
                                                                         g, g_dtype, inner, total);
    MSTK_CHECK_LAUNCH("eps_grad");
    return MSTK_OK;
}
int mstk_adam_step(float* p, const float* g, float* m, float* v, size_t count, float lr, float beta1, float beta2, float eps, int64_t step,
                   float grad_scale, void* stream) {
    MSTK_REQUIRE(p && g && m && v && count > 0 && step >= 1, "adam: bad args");
    MSTK_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
                 "adam: arenas must be 16-byte aligned");
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    adam_kernel<<<bw_grid(count / 4 + 1, 256, 8), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, count, lr, beta1, beta2, eps, (float)bc1,
                                                                               (float)sqrt(bc2), grad_scale, nullptr, 1.f);
    MSTK_CHECK_LAUNCH("adam");
    return MSTK_OK;
}

int mstk_adam_step_dev(float* p, const float* g, float* m, float* v, size_t count, float lr, float beta1, float beta2, float eps,
                       const int64_t* step_dev, float grad_scale, void* stream) {
    MSTK_REQUIRE(p && g && m && v && count > 0 && step_dev, "adam: bad args");
    MSTK_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
                 "adam: arenas must be 16-byte aligned");
    adam_kernel<<<bw_grid(count / 4 + 1, 256, 8), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, count, lr, beta1, beta2, eps, 1.f, 1.f, grad_scale,
                                                                               reinterpret_cast<const long long*>(step_dev), 1.f);
    MSTK_CHECK_LAUNCH("adam");
    return MSTK_OK;
}

int mstk_adamw_step_dev(float* p, const float* g, float* m, float* v, size_t count, float lr, float beta1, float beta2, float eps,
                        float weight_decay, const int64_t* step_dev, float grad_scale, void* stream) {
    MSTK_REQUIRE(p && g && m && v && count > 0 && step_dev && weight_decay >= 0.f, "adamw: bad args");
    MSTK_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
                 "adamw: arenas must be 16-byte aligned");
    adam_kernel<<<bw_grid(count / 4 + 1, 256, 8), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, count, lr, beta1, beta2, eps, 1.f, 1.f, grad_scale,
                                                                               reinterpret_cast<const long long*>(step_dev),
                                                                               1.f - lr * weight_decay);
    MSTK_CHECK_LAUNCH("adamw");
    return MSTK_OK;
}

int mstk_kl_grad(const float* z, const void* dz, int dz_dtype, float scale, float* g, size_t count, void* stream) {
    MSTK_REQUIRE(z && g && count > 0, "kl_grad: bad args");
    kl_grad_kernel<<<bw_grid(count, 256), 256, 0, (cudaStream_t)stream>>>(z, dz, dz_dtype, scale, g, count);
    MSTK_CHECK_LAUNCH("kl_grad");
    return MSTK_OK;
}

}  // extern "C"
