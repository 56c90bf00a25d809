// HBM-streaming kernels: weight packing, layout changes, BatchNorm apply, scheduler
// arithmetic, loss reductions, pooling.  128-bit vectorised where alignment allows,
// warp-shuffle reductions, grids sized in multiples of the SM count.
#include "common.cuh"
#include "attention_body.cuh"

namespace mstk {

static int g_sms = 0;
static int num_sms() {
    if (g_sms == 0) {
        int dev = 0, sms = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0)
            g_sms = sms;
        else
            g_sms = 148;
    }
    return g_sms;
}
static unsigned grid_for(size_t work_items, int threads, int per_sm = 8) {
    size_t blocks = (work_items + threads - 1) / threads;
    const size_t cap = (size_t)num_sms() * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}

// ---------------------------------------------------------------------------
// weight packing
// ---------------------------------------------------------------------------
// transposed == 0: src [cout][cin][k][k] ; transposed == 1: src [cin][cout][k][k]
// dst [k*k][cout][cin]
__global__ void pack_weight_kernel(const float* __restrict__ src, const float* __restrict__ scale, void* dst, int cout, int cin,
                                   int kk, int transposed, int dtype) {
    const size_t total = (size_t)kk * cout * cin;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int ci = (int)(i % cin);
        const int co = (int)((i / cin) % cout);
        const int s = (int)(i / ((size_t)cin * cout));
        const size_t si = transposed ? ((size_t)ci * cout + co) * kk + s : ((size_t)co * cin + ci) * kk + s;
        float v = src[si];
        if (scale) v *= scale[co];
        store_from_f(dst, dtype, i, v);
    }
}

// ---------------------------------------------------------------------------
// layout changes
// ---------------------------------------------------------------------------
// [n][c][hw] fp32 -> [n][hw][c] dtype, 32x32 shared-memory transpose per (n)
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, void* dst, int c, int hw, int dtype) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const float* s = src + (size_t)n * c * hw;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int cc = c0 + j, pp = p0 + threadIdx.x;
        tile[j][threadIdx.x] = (cc < c && pp < hw) ? s[(size_t)cc * hw + pp] : 0.f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int pp = p0 + j, cc = c0 + threadIdx.x;
        if (pp < hw && cc < c) store_from_f(dst, dtype, ((size_t)n * hw + pp) * c + cc, tile[threadIdx.x][j]);
    }
}
__global__ void nhwc_to_nchw_kernel(const void* __restrict__ src, float* dst, int c, int hw, int dtype) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int pp = p0 + j, cc = c0 + threadIdx.x;
        tile[j][threadIdx.x] = (cc < c && pp < hw) ? load_as_f(src, dtype, ((size_t)n * hw + pp) * c + cc) : 0.f;
    }
    __syncthreads();
    float* d = dst + (size_t)n * c * hw;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int cc = c0 + j, pp = p0 + threadIdx.x;
        if (cc < c && pp < hw) d[(size_t)cc * hw + pp] = tile[threadIdx.x][j];
    }
}

// Packed uint8 spectrogram cache -> training batch: out[i] = crop(pool[index[i]], crop_h x crop_w, top-left) / 255
// (models/dataset.py:51-53: img.crop((0, 0, 128, 128)), Grayscale, ToTensor).  One thread per output pixel; a warp reads
// 32 consecutive bytes of one image row and writes 128 consecutive bytes.
__global__ void u8_gather_crop_kernel(const uint8_t* __restrict__ pool, const long long* __restrict__ index, float* __restrict__ out,
                                      const int h, const int w, const int crop_h, const int crop_w, const size_t total) {
    const size_t per = (size_t)crop_h * crop_w;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t s = i / per, r = i - s * per;
        const int y = (int)(r / crop_w), x = (int)(r - (size_t)y * crop_w);
        out[i] = (float)pool[((size_t)index[s] * h + y) * w + x] / 255.0f;      // IEEE division, as ToTensor's .div(255)
    }
}

__global__ void cast_kernel(const void* __restrict__ src, int sd, void* dst, int dd, size_t count) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x)
        store_from_f(dst, dd, i, load_as_f(src, sd, i));
}

// ---------------------------------------------------------------------------
// BatchNorm apply
// ---------------------------------------------------------------------------
struct BnP {
    const void* x; void* y; const double* stats; const float* gamma; const float* beta;
    float* running_mean; float* running_var; float* save_mean; float* save_invstd;
    long long count, rows; int c; float eps, momentum; int act, dtype, out_dtype;
    int vec;            // 1: c % 8 == 0 and x / y 16-byte aligned -> 8 channels per thread with 128-bit accesses
};

// one thread handles 8 consecutive channels of one row (requires c % 8 == 0) or falls
// back to scalar indexing.  scale/shift are staged in shared memory once per CTA.
__global__ void __launch_bounds__(256) bn_apply_kernel(const BnP a) {
    extern __shared__ float sm[];   // scale[c], shift[c]
    float* scale = sm;
    float* shift = sm + a.c;
    const double inv_n = 1.0 / (double)a.count;
    for (int ch = threadIdx.x; ch < a.c; ch += blockDim.x) {
        const double mean = a.stats[ch] * inv_n;
        double var = a.stats[a.c + ch] * inv_n - mean * mean;
        if (var < 0.0) var = 0.0;
        const float invstd = (float)(1.0 / sqrt(var + (double)a.eps));
        const float g = a.gamma ? a.gamma[ch] : 1.f;
        const float b = a.beta ? a.beta[ch] : 0.f;
        scale[ch] = invstd * g;
        shift[ch] = b - (float)mean * invstd * g;
        if (blockIdx.x == 0) {
            if (a.save_mean) a.save_mean[ch] = (float)mean;
            if (a.save_invstd) a.save_invstd[ch] = invstd;
            if (a.running_mean) a.running_mean[ch] = (1.f - a.momentum) * a.running_mean[ch] + a.momentum * (float)mean;
            if (a.running_var) {
                const double unb = a.count > 1 ? var * (double)a.count / (double)(a.count - 1) : var;
                a.running_var[ch] = (1.f - a.momentum) * a.running_var[ch] + a.momentum * (float)unb;
            }
        }
    }
    __syncthreads();
    const size_t total = (size_t)a.rows * a.c;
    if (a.vec) {
        const size_t nvec = total >> 3;
        ChanWalk cw((blockIdx.x * (size_t)blockDim.x + threadIdx.x) << 3, ((size_t)gridDim.x * blockDim.x) << 3, a.c);
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x, cw.next()) {
            const int ch = cw.ch;
            float v[8];
            if (a.dtype == MSTK_F32) {
                const float4 q0 = reinterpret_cast<const float4*>(a.x)[2 * i];
                const float4 q1 = reinterpret_cast<const float4*>(a.x)[2 * i + 1];
                v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w; v[4] = q1.x; v[5] = q1.y; v[6] = q1.z; v[7] = q1.w;
            } else {
                const uint4 q = reinterpret_cast<const uint4*>(a.x)[i];
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(h[k]);
                    v[2 * k] = f.x; v[2 * k + 1] = f.y;
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float r = v[k] * scale[ch + k] + shift[ch + k];
                if (a.act == MSTK_ACT_RELU) r = fmaxf(r, 0.f);
                v[k] = r;
            }
            if (a.out_dtype == MSTK_F32) {
                reinterpret_cast<float4*>(a.y)[2 * i] = make_float4(v[0], v[1], v[2], v[3]);
                reinterpret_cast<float4*>(a.y)[2 * i + 1] = make_float4(v[4], v[5], v[6], v[7]);
            } else {
                uint4 q;
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
                for (int k = 0; k < 4; ++k) h[k] = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
                reinterpret_cast<uint4*>(a.y)[i] = q;
            }
        }
    } else {
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
            const int ch = (int)(i % a.c);
            float r = load_as_f(a.x, a.dtype, i) * scale[ch] + shift[ch];
            if (a.act == MSTK_ACT_RELU) r = fmaxf(r, 0.f);
            store_from_f(a.y, a.out_dtype, i, r);
        }
    }
}

// ---------------------------------------------------------------------------
// time embedding MLP: one CTA per sample, `dim` threads
// ---------------------------------------------------------------------------
// TE_SPB samples per CTA; each Linear's weight matrix is staged once per CTA in shared memory
// with coalesced reads ([dim][dim+1] padding: thread i then walks row i conflict-free).
constexpr int TE_SPB = 16;
__global__ void time_embed_kernel(const float* __restrict__ t, int nb, const float* __restrict__ w1, const float* __restrict__ b1,
                                  const float* __restrict__ w2, const float* __restrict__ b2, float* __restrict__ out,
                                  float* __restrict__ e_out, float* __restrict__ pre_out, float* __restrict__ h_out, int dim) {
    extern __shared__ float sm[];   // ws[dim][dim+1], e[TE_SPB][dim], h[TE_SPB][dim]
    float* ws = sm;
    float* e = ws + (size_t)dim * (dim + 1);
    float* h = e + TE_SPB * dim;
    const int i = threadIdx.x;
    const int s0 = blockIdx.x * TE_SPB;
    const int ns = min(TE_SPB, nb - s0);
    const int half = dim / 2;
    for (int idx = i; idx < dim * dim; idx += dim) ws[(idx / dim) * (dim + 1) + (idx % dim)] = w1[idx];
    {
        const int k = i < half ? i : i - half;
        // torch: exp(arange(half) * -(log(10000)/(half-1))) in fp32
        const float coef = -(float)(log(10000.0) / (double)(half - 1));
        const float fr = (float)exp((double)((float)k * coef));
        for (int s = 0; s < ns; ++s) {
            const float arg = t[s0 + s] * fr;
            e[s * dim + i] = i < half ? sinf(arg) : cosf(arg);
            if (e_out) e_out[(size_t)(s0 + s) * dim + i] = e[s * dim + i];
        }
    }
    __syncthreads();
    {
        float acc[TE_SPB];
        const float bb = b1[i];
#pragma unroll
        for (int s = 0; s < TE_SPB; ++s) acc[s] = bb;
        const float* wr = ws + (size_t)i * (dim + 1);
        for (int k = 0; k < dim; ++k) {
            const float w = wr[k];
#pragma unroll
            for (int s = 0; s < TE_SPB; ++s) acc[s] = fmaf(e[s * dim + k], w, acc[s]);
        }
        // exact GELU (nn.GELU default): 0.5 x (1 + erf(x / sqrt(2)))
#pragma unroll
        for (int s = 0; s < TE_SPB; ++s) h[s * dim + i] = 0.5f * acc[s] * (1.f + erff(acc[s] * 0.70710678118654752440f));
        if (pre_out)
            for (int s = 0; s < ns; ++s) pre_out[(size_t)(s0 + s) * dim + i] = acc[s];
        if (h_out)
            for (int s = 0; s < ns; ++s) h_out[(size_t)(s0 + s) * dim + i] = h[s * dim + i];
    }
    __syncthreads();
    for (int idx = i; idx < dim * dim; idx += dim) ws[(idx / dim) * (dim + 1) + (idx % dim)] = w2[idx];
    __syncthreads();
    {
        float acc[TE_SPB];
        const float bb = b2[i];
#pragma unroll
        for (int s = 0; s < TE_SPB; ++s) acc[s] = bb;
        const float* wr = ws + (size_t)i * (dim + 1);
        for (int k = 0; k < dim; ++k) {
            const float w = wr[k];
#pragma unroll
            for (int s = 0; s < TE_SPB; ++s) acc[s] = fmaf(h[s * dim + k], w, acc[s]);
        }
        for (int s = 0; s < ns; ++s) out[(size_t)(s0 + s) * dim + i] = acc[s];
    }
}

// ---------------------------------------------------------------------------
// attention core (body in attention_body.cuh): grid (b*heads, ceil(l / QPB))
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(ATTN_WARPS * 32) attention_kernel(const AttnP a) {
    extern __shared__ float sm[];
    // programmatic dependent launch: wait for the producer of q (the projection conv), then let the next kernel (the
    // output projection, which prefetches its weights before ITS wait) start its prologue under this one
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    attention_unit(a, (int)blockIdx.x, (int)blockIdx.y, sm, ATTN_WARPS);
}

// The UNet's cross-attention is tiny (L = 16 or 4 latent tokens against S = 16 or 4 style tokens, d = 64 / 128) and
// purely latency bound, so: one CTA per (batch, head, block of ATTN2_QPB queries) stages Q, K and V ONCE with 16-byte
// loads that are all in flight together, one thread per (query, key) pair for the scores, one warp per softmax row,
// one thread per output element for P V.  Same arithmetic order per dot product as attention_unit (q pre-scaled,
// ascending d, ascending s), so both paths agree to rounding of the softmax sum.
constexpr int ATTN2_THREADS = 256;
constexpr int ATTN2_QPB = 16;
template <typename T>
__device__ __forceinline__ void stage8(float* dst, const T* src);
template <>
__device__ __forceinline__ void stage8<__nv_bfloat16>(float* dst, const __nv_bfloat16* src) {
    const uint4 r = __ldg(reinterpret_cast<const uint4*>(src));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        dst[2 * e] = f.x; dst[2 * e + 1] = f.y;
    }
}
template <>
__device__ __forceinline__ void stage8<float>(float* dst, const float* src) {
    const float4 r0 = __ldg(reinterpret_cast<const float4*>(src)), r1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
    dst[0] = r0.x; dst[1] = r0.y; dst[2] = r0.z; dst[3] = r0.w;
    dst[4] = r1.x; dst[5] = r1.y; dst[6] = r1.z; dst[7] = r1.w;
}
template <typename T>
__global__ void __launch_bounds__(ATTN2_THREADS) attention_block_kernel(const AttnP a) {
    extern __shared__ __align__(16) float sm[];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int d = a.e / a.heads, d8 = d >> 3;
    const int dp = d + 4;                            // row pitch: 16-byte aligned rows, 8 consecutive rows on distinct banks
    const int bh = (int)blockIdx.x, b = bh / a.heads, hd = bh - b * a.heads;
    const int q0 = (int)blockIdx.y * ATTN2_QPB;
    const int lq = min(ATTN2_QPB, a.l - q0);
    float* Ks = sm;                                  // [s][dp]
    float* Vs = Ks + (size_t)a.s * dp;               // [s][dp]
    float* Qs = Vs + (size_t)a.s * dp;               // [lq][dp] (pre-scaled)
    float* Ps = Qs + (size_t)ATTN2_QPB * dp;         // [lq][s+1]
    const float scaling = 1.0f / sqrtf((float)d);
    const T* kp = reinterpret_cast<const T*>(a.k) + (size_t)b * a.s * a.ks + hd * d;
    const T* vp = reinterpret_cast<const T*>(a.v) + (size_t)b * a.s * a.vs + hd * d;
    const T* qp = reinterpret_cast<const T*>(a.q) + ((size_t)b * a.l + q0) * a.qs + hd * d;
    const int nk = a.s * d8, nq = lq * d8;
    for (int i = threadIdx.x; i < 2 * nk + nq; i += ATTN2_THREADS) {
        float f[8];
        float* dst;
        float sc = 1.f;
        if (i < nk) {
            const int r = i / d8, c = (i - r * d8) * 8;
            stage8<T>(f, kp + (size_t)r * a.ks + c); dst = Ks + r * dp + c;
        } else if (i < 2 * nk) {
            const int ii = i - nk, r = ii / d8, c = (ii - r * d8) * 8;
            stage8<T>(f, vp + (size_t)r * a.vs + c); dst = Vs + r * dp + c;
        } else {
            const int ii = i - 2 * nk, r = ii / d8, c = (ii - r * d8) * 8;
            stage8<T>(f, qp + (size_t)r * a.qs + c); dst = Qs + r * dp + c; sc = scaling;
        }
        *reinterpret_cast<float4*>(dst) = make_float4(f[0] * sc, f[1] * sc, f[2] * sc, f[3] * sc);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(f[4] * sc, f[5] * sc, f[6] * sc, f[7] * sc);
    }
    __syncthreads();
    // scores: one thread per (query, key); the kernel is bound by shared-memory instructions, so 128-bit reads
    // (same ascending-d fma order as the scalar form)
    for (int pr = threadIdx.x; pr < lq * a.s; pr += ATTN2_THREADS) {
        const int qi = pr / a.s, si = pr - qi * a.s;
        const float4* qr = reinterpret_cast<const float4*>(Qs + qi * dp);
        const float4* kr = reinterpret_cast<const float4*>(Ks + si * dp);
        float acc = 0.f;
#pragma unroll 4
        for (int c = 0; c < (d >> 2); ++c) {
            const float4 qv = qr[c], kv = kr[c];
            acc = fmaf(qv.x, kv.x, acc); acc = fmaf(qv.y, kv.y, acc); acc = fmaf(qv.z, kv.z, acc); acc = fmaf(qv.w, kv.w, acc);
        }
        Ps[qi * (a.s + 1) + si] = acc;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int qi = warp; qi < lq; qi += ATTN2_THREADS / 32) {
        float* pw = Ps + qi * (a.s + 1);
        float mx = -INFINITY;
        for (int si = lane; si < a.s; si += 32) mx = fmaxf(mx, pw[si]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int si = lane; si < a.s; si += 32) {
            const float ev = expf(pw[si] - mx);
            pw[si] = ev;
            sum += ev;
        }
        sum = warp_sum(sum);
        if (lane == 0) pw[a.s] = 1.f / sum;
    }
    __syncthreads();
    // P V: one thread per (query, 4 consecutive channels)
    T* op = reinterpret_cast<T*>(a.out) + ((size_t)b * a.l + q0) * a.os + hd * d;
    const int d4 = d >> 2;
    for (int o = threadIdx.x; o < lq * d4; o += ATTN2_THREADS) {
        const int qi = o / d4, di = (o - qi * d4) * 4;
        const float* pw = Ps + qi * (a.s + 1);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int si = 0; si < a.s; ++si) {
            const float pv = pw[si];
            const float4 vv = *reinterpret_cast<const float4*>(Vs + si * dp + di);
            acc.x = fmaf(pv, vv.x, acc.x); acc.y = fmaf(pv, vv.y, acc.y); acc.z = fmaf(pv, vv.z, acc.z); acc.w = fmaf(pv, vv.w, acc.w);
        }
        const float inv = pw[a.s];
        T* dst = op + (size_t)qi * a.os + di;
        if (sizeof(T) == 4) {
            *reinterpret_cast<float4*>(dst) = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
        } else {
            __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x * inv, acc.y * inv), hi = __floats2bfloat162_rn(acc.z * inv, acc.w * inv);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(dst) = pk;
        }
    }
}
__host__ __device__ inline size_t attn2_smem_floats(int s, int d) {
    return (size_t)2 * s * (d + 4) + (size_t)ATTN2_QPB * (d + 4) + (size_t)ATTN2_QPB * (s + 1);
}

// ---------------------------------------------------------------------------
// scheduler arithmetic
// ---------------------------------------------------------------------------
__global__ void q_sample_kernel(const float* __restrict__ x0, const float* __restrict__ eps, const float* __restrict__ ab,
                                const long long* __restrict__ t, float* __restrict__ z, __nv_bfloat16* __restrict__ zlp,
                                long long inner, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int b = (int)(i / (size_t)inner);
        const float abv = ab[t[b]];
        const float r = sqrtf(abv) * x0[i] + sqrtf(1.f - abv) * eps[i];
        z[i] = r;
        if (zlp) zlp[i] = __float2bfloat16_rn(r);
    }
}
__global__ void predict_x0_kernel(const float* __restrict__ z, const float* __restrict__ eh, const float* __restrict__ ab,
                                  const long long* __restrict__ t, float* __restrict__ x0, __nv_bfloat16* __restrict__ x0lp,
                                  long long inner, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int b = (int)(i / (size_t)inner);
        const float abv = ab[t[b]];
        const float r = (z[i] - sqrtf(1.f - abv) * eh[i]) / sqrtf(abv);
        x0[i] = r;
        if (x0lp) x0lp[i] = __float2bfloat16_rn(r);
    }
}
__global__ void ddim_step_kernel(const DdimP d, const float* __restrict__ eps, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const float e = eps[i];
        const float xc = d.x_cur[i];
        const float dt = d.s1m_t * e;
        const float x0 = (xc - dt) / d.sab_t;
        const float dn = d.s1m_n * e;
        const float xn = d.sab_n * x0 + dn + d.eta * (dn - dt);
        d.x_next[i] = xn;
        if (d.x_next_lp) d.x_next_lp[i] = __float2bfloat16_rn(xn);
        if (d.x0_pred) d.x0_pred[i] = x0;
    }
}

// ---------------------------------------------------------------------------
// reductions.  scratch: double acc[4] + uint32 counter (at byte offset 32).
// Last CTA to arrive finalises and re-zeroes the scratch.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double block_sum(double v) {
    __shared__ double red[32];
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

// mode 0: mse(a,b) ; mode 1: kl(a)
__global__ void __launch_bounds__(256) reduce_loss_kernel(const void* __restrict__ a, const void* __restrict__ b, int dtype,
                                                         size_t count, int mode, double* scratch, float* out) {
    float local = 0.f;
    double acc = 0.0;
    int chunk = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        const float x = load_as_f(a, dtype, i);
        if (mode == 0) {
            const float df = x - load_as_f(b, dtype, i);
            local = fmaf(df, df, local);
        } else {
            const float sq = x * x;
            local += 0.5f * (sq - 1.f - logf(sq + 1e-8f));
        }
        if (++chunk == 64) { acc += (double)local; local = 0.f; chunk = 0; }
    }
    acc += (double)local;
    const double tot = block_sum(acc);
    if (threadIdx.x == 0) {
        atomicAdd(scratch, tot);
        __threadfence();
        unsigned int* ctr = reinterpret_cast<unsigned int*>(scratch + 4);
        const unsigned int old = atomicAdd(ctr, 1u);
        if (old == gridDim.x - 1) {
            __threadfence();
            const double s = atomicAdd(scratch, 0.0);
            out[0] = (float)(s / (double)count);
            scratch[0] = 0.0;
            *ctr = 0u;
        }
    }
}

// per-sample sum / sumsq of f and g: tmp[b*4 + {0,1,2,3}]
__global__ void __launch_bounds__(256) feat_stats_kernel(const void* __restrict__ f, const void* __restrict__ g, int dtype,
                                                        long long inner, double* tmp) {
    const int b = blockIdx.y;
    const size_t base = (size_t)b * inner;
    double sf = 0, qf = 0, sg = 0, qg = 0;
    float lsf = 0, lqf = 0, lsg = 0, lqg = 0;
    int chunk = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < inner; i += (long long)gridDim.x * blockDim.x) {
        const float x = load_as_f(f, dtype, base + i), y = load_as_f(g, dtype, base + i);
        lsf += x; lqf = fmaf(x, x, lqf); lsg += y; lqg = fmaf(y, y, lqg);
        if (++chunk == 32) { sf += lsf; qf += lqf; sg += lsg; qg += lqg; lsf = lqf = lsg = lqg = 0.f; chunk = 0; }
    }
    sf += lsf; qf += lqf; sg += lsg; qg += lqg;
    sf = block_sum(sf); qf = block_sum(qf); sg = block_sum(sg); qg = block_sum(qg);
    if (threadIdx.x == 0) {
        atomicAdd(tmp + b * 4 + 0, sf); atomicAdd(tmp + b * 4 + 1, qf);
        atomicAdd(tmp + b * 4 + 2, sg); atomicAdd(tmp + b * 4 + 3, qg);
    }
}
// sum over everything of (f/(std_f+1e-8) - g/(std_g+1e-8))^2 ; tmp2 = [acc, -, -, -, counter]
__global__ void __launch_bounds__(256) feat_mse_kernel(const void* __restrict__ f, const void* __restrict__ g, int dtype, int nb,
                                                      long long inner, double* tmp, double* tmp2, float weight, float* out) {
    const int b = blockIdx.y;
    const size_t base = (size_t)b * inner;
    const double n = (double)inner;
    const double mf = tmp[b * 4 + 0] / n, mg = tmp[b * 4 + 2] / n;
    double vf = (tmp[b * 4 + 1] - n * mf * mf) / (n - 1.0), vg = (tmp[b * 4 + 3] - n * mg * mg) / (n - 1.0);
    if (vf < 0) vf = 0;
    if (vg < 0) vg = 0;
    const float rf = 1.f / ((float)sqrt(vf) + 1e-8f), rg = 1.f / ((float)sqrt(vg) + 1e-8f);
    double acc = 0.0;
    float local = 0.f;
    int chunk = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < inner; i += (long long)gridDim.x * blockDim.x) {
        const float df = load_as_f(f, dtype, base + i) * rf - load_as_f(g, dtype, base + i) * rg;
        local = fmaf(df, df, local);
        if (++chunk == 64) { acc += local; local = 0.f; chunk = 0; }
    }
    acc += local;
    acc = block_sum(acc);
    if (threadIdx.x == 0) {
        atomicAdd(tmp2, acc);
        __threadfence();
        unsigned int* ctr = reinterpret_cast<unsigned int*>(tmp2 + 1);
        const unsigned int old = atomicAdd(ctr, 1u);
        if (old == gridDim.x * gridDim.y - 1) {
            __threadfence();
            const double s = atomicAdd(tmp2, 0.0);
            atomicAdd(out, weight * (float)(s / ((double)nb * n)));
            tmp2[0] = 0.0;
            *ctr = 0u;
            for (int i = 0; i < nb * 4; ++i) tmp[i] = 0.0;
        }
    }
}

__global__ void maxpool2x2_kernel(const void* __restrict__ x, void* y, int n, int h, int w, int c, int dtype) {
    const int ho = h / 2, wo = w / 2;
    const size_t total = (size_t)n * ho * wo * c;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int ch = (int)(i % c);
        size_t r = i / c;
        const int ox = (int)(r % wo); r /= wo;
        const int oy = (int)(r % ho);
        const int ni = (int)(r / ho);
        const size_t b00 = (((size_t)ni * h + 2 * oy) * w + 2 * ox) * c + ch;
        const float v = fmaxf(fmaxf(load_as_f(x, dtype, b00), load_as_f(x, dtype, b00 + c)),
                              fmaxf(load_as_f(x, dtype, b00 + (size_t)w * c), load_as_f(x, dtype, b00 + (size_t)w * c + c)));
        store_from_f(y, dtype, i, v);
    }
}

}  // namespace mstk

using namespace mstk;

extern "C" {

int mstk_pack_conv_weight(const float* w, const float* scale, void* dst, int cout, int cin, int k, int dtype, void* stream) {
    MSTK_REQUIRE(w && dst && cout > 0 && cin > 0 && k > 0, "pack_conv_weight: bad args");
    const size_t total = (size_t)k * k * cout * cin;
    pack_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, scale, dst, cout, cin, k * k, 0, dtype);
    MSTK_CHECK_LAUNCH("pack_conv_weight");
    return MSTK_OK;
}
int mstk_pack_convt_weight(const float* w, const float* scale, void* dst, int cin, int cout, int k, int dtype, void* stream) {
    MSTK_REQUIRE(w && dst && cout > 0 && cin > 0 && k > 0, "pack_convt_weight: bad args");
    const size_t total = (size_t)k * k * cout * cin;
    pack_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(w, scale, dst, cout, cin, k * k, 1, dtype);
    MSTK_CHECK_LAUNCH("pack_convt_weight");
    return MSTK_OK;
}

int mstk_nchw_to_nhwc(const float* src, void* dst, int n, int c, int h, int w, int dst_dtype, void* stream) {
    MSTK_REQUIRE(src && dst && n > 0 && c > 0 && h > 0 && w > 0, "nchw_to_nhwc: bad args");
    MSTK_REQUIRE(n <= 65535, "nchw_to_nhwc: batch too large");
    const int hw = h * w;
    dim3 grid((hw + 31) / 32, (c + 31) / 32, n);
    nchw_to_nhwc_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(src, dst, c, hw, dst_dtype);
    MSTK_CHECK_LAUNCH("nchw_to_nhwc");
    return MSTK_OK;
}
int mstk_nhwc_to_nchw(const void* src, float* dst, int n, int c, int h, int w, int src_dtype, void* stream) {
    MSTK_REQUIRE(src && dst && n > 0 && c > 0 && h > 0 && w > 0, "nhwc_to_nchw: bad args");
    MSTK_REQUIRE(n <= 65535, "nhwc_to_nchw: batch too large");
    const int hw = h * w;
    dim3 grid((hw + 31) / 32, (c + 31) / 32, n);
    nhwc_to_nchw_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(src, dst, c, hw, src_dtype);
    MSTK_CHECK_LAUNCH("nhwc_to_nchw");
    return MSTK_OK;
}
int mstk_u8_gather_crop(const uint8_t* pool, const int64_t* index, float* out, int b, int h, int w, int crop_h, int crop_w, void* stream) {
    MSTK_REQUIRE(pool && index && out && b > 0 && h > 0 && w > 0, "u8_gather_crop: bad args");
    MSTK_REQUIRE(crop_h > 0 && crop_w > 0 && crop_h <= h && crop_w <= w, "u8_gather_crop: crop %dx%d outside the stored %dx%d image", crop_h,
                 crop_w, h, w);
    const size_t total = (size_t)b * crop_h * crop_w;
    u8_gather_crop_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(pool, (const long long*)index, out, h, w, crop_h, crop_w, total);
    MSTK_CHECK_LAUNCH("u8_gather_crop");
    return MSTK_OK;
}
int mstk_cast(const void* src, int sd, void* dst, int dd, size_t count, void* stream) {
    MSTK_REQUIRE(src && dst, "cast: null");
    if (count == 0) return MSTK_OK;
    cast_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(src, sd, dst, dd, count);
    MSTK_CHECK_LAUNCH("cast");
    return MSTK_OK;
}

int mstk_bn_apply(const mstk_bn_args* a, void* stream) {
    MSTK_REQUIRE(a && a->x && a->y && a->stats, "bn_apply: null");
    MSTK_REQUIRE(a->c > 0 && a->c <= 4096 && a->count > 0 && a->rows >= 0 && a->rows <= a->count, "bn_apply: bad dims");
    BnP p;
    p.x = a->x; p.y = a->y; p.stats = a->stats; p.gamma = a->gamma; p.beta = a->beta;
    p.running_mean = a->running_mean; p.running_var = a->running_var; p.save_mean = a->save_mean; p.save_invstd = a->save_invstd;
    p.count = a->count; p.rows = a->rows > 0 ? a->rows : a->count; p.c = a->c; p.eps = a->eps; p.momentum = a->momentum; p.act = a->act; p.dtype = a->dtype; p.out_dtype = a->out_dtype;
    const size_t total = (size_t)p.rows * a->c;
    p.vec = ((a->c & 7) == 0 && aligned16({a->x, a->y})) ? 1 : 0;
    const size_t items = p.vec ? total / 8 : total;
    bn_apply_kernel<<<grid_for(items, 256), 256, 2 * a->c * sizeof(float), (cudaStream_t)stream>>>(p);
    MSTK_CHECK_LAUNCH("bn_apply");
    return MSTK_OK;
}

int mstk_time_embed_train(const float* t, int b, const float* w1, const float* b1, const float* w2, const float* b2, float* out,
                          float* e_out, float* pre_out, float* h_out, int dim, void* stream);
int mstk_time_embed(const float* t, int b, const float* w1, const float* b1, const float* w2, const float* b2, float* out, int dim,
                    void* stream) {
    return mstk_time_embed_train(t, b, w1, b1, w2, b2, out, nullptr, nullptr, nullptr, dim, stream);
}
int mstk_time_embed_train(const float* t, int b, const float* w1, const float* b1, const float* w2, const float* b2, float* out,
                          float* e_out, float* pre_out, float* h_out, int dim, void* stream) {
    MSTK_REQUIRE(t && w1 && b1 && w2 && b2 && out && b > 0, "time_embed: null");
    MSTK_REQUIRE(dim >= 4 && dim <= 1024 && (dim % 2) == 0, "time_embed: bad dim");
    MSTK_REQUIRE(dim <= 192, "time_embed: dim %d too large for the shared-memory weight tile", dim);
    const size_t smem = ((size_t)dim * (dim + 1) + 2 * TE_SPB * dim) * sizeof(float);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(time_embed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return cuda_fail(e, "time_embed smem attr");
        smem_set = smem;
    }
    time_embed_kernel<<<(b + TE_SPB - 1) / TE_SPB, dim, smem, (cudaStream_t)stream>>>(t, b, w1, b1, w2, b2, out, e_out, pre_out, h_out, dim);
    MSTK_CHECK_LAUNCH("time_embed");
    return MSTK_OK;
}

int mstk_attention_fwd(const mstk_attn_args* a, void* stream) {
    MSTK_REQUIRE(a && a->q && a->k && a->v && a->out, "attention: null");
    MSTK_REQUIRE(a->b > 0 && a->l > 0 && a->s > 0 && a->heads > 0 && a->e % a->heads == 0, "attention: bad dims");
    AttnP p;
    p.q = a->q; p.k = a->k; p.v = a->v; p.out = a->out; p.b = a->b; p.l = a->l; p.s = a->s; p.e = a->e; p.heads = a->heads;
    p.qs = a->q_stride; p.ks = a->k_stride; p.vs = a->v_stride; p.os = a->out_stride; p.dtype = a->dtype;
    const int d = a->e / a->heads;
    {
        // block kernel: 16-byte staging loads need aligned rows; everything else takes the warp-per-query kernel below
        const int esz = a->dtype == MSTK_F32 ? 4 : 2, per16 = 16 / esz;
        const size_t smem2 = attn2_smem_floats(a->s, d) * sizeof(float);
        const bool aligned = (d % 8) == 0 && (a->q_stride % per16) == 0 && (a->k_stride % per16) == 0 && (a->v_stride % per16) == 0 &&
                             (a->out_stride % per16) == 0 &&
                             ((reinterpret_cast<uintptr_t>(a->q) | reinterpret_cast<uintptr_t>(a->k) | reinterpret_cast<uintptr_t>(a->v) |
                               reinterpret_cast<uintptr_t>(a->out)) & 15) == 0;
        if (aligned && smem2 <= 48 * 1024) {
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3(a->b * a->heads, (a->l + ATTN2_QPB - 1) / ATTN2_QPB);
            cfg.blockDim = dim3(ATTN2_THREADS, 1, 1);
            cfg.dynamicSmemBytes = smem2;
            cfg.stream = (cudaStream_t)stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            cudaError_t le = a->dtype == MSTK_F32 ? cudaLaunchKernelEx(&cfg, attention_block_kernel<float>, p)
                                                  : cudaLaunchKernelEx(&cfg, attention_block_kernel<__nv_bfloat16>, p);
            if (le != cudaSuccess) return cuda_fail(le, "attention launch");
            return MSTK_OK;
        }
    }
    const size_t smem = ((size_t)a->s * (d + 1) + (size_t)a->s * d + ATTN_WARPS * a->s + ATTN_WARPS * d) * sizeof(float);
    MSTK_REQUIRE(smem <= 200 * 1024, "attention: s*d too large for the shared-memory kernel (s=%d d=%d)", a->s, d);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "attention smem attr");
        smem_set = 200 * 1024;
    }
    dim3 grid(a->b * a->heads, (a->l + ATTN_QPB - 1) / ATTN_QPB);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = dim3(ATTN_WARPS * 32, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t le = cudaLaunchKernelEx(&cfg, attention_kernel, p);
    if (le != cudaSuccess) return cuda_fail(le, "attention launch");
    return MSTK_OK;
}

int mstk_q_sample(const float* x0, const float* eps, const float* ab, const int64_t* t, float* z, void* zlp, int b, int64_t inner,
                  void* stream) {
    MSTK_REQUIRE(x0 && eps && ab && t && z && b > 0 && inner > 0, "q_sample: bad args");
    const size_t total = (size_t)b * inner;
    q_sample_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(x0, eps, ab, (const long long*)t, z, (__nv_bfloat16*)zlp,
                                                                            inner, total);
    MSTK_CHECK_LAUNCH("q_sample");
    return MSTK_OK;
}
int mstk_predict_x0(const float* z, const float* eh, const float* ab, const int64_t* t, float* x0, void* x0lp, int b, int64_t inner,
                    void* stream) {
    MSTK_REQUIRE(z && eh && ab && t && x0 && b > 0 && inner > 0, "predict_x0: bad args");
    const size_t total = (size_t)b * inner;
    predict_x0_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(z, eh, ab, (const long long*)t, x0, (__nv_bfloat16*)x0lp,
                                                                              inner, total);
    MSTK_CHECK_LAUNCH("predict_x0");
    return MSTK_OK;
}
int mstk_ddim_step(const mstk_ddim_args* a, const float* eps, size_t count, void* stream) {
    MSTK_REQUIRE(a && a->x_cur && a->x_next && eps, "ddim_step: null");
    DdimP d;
    d.x_cur = a->x_cur; d.x_next = a->x_next; d.x_next_lp = (__nv_bfloat16*)a->x_next_lp; d.x0_pred = a->x0_pred;
    d.sab_t = a->sab_t; d.s1m_t = a->s1m_t; d.sab_n = a->sab_n; d.s1m_n = a->s1m_n; d.eta = a->eta;
    if (count == 0) return MSTK_OK;
    ddim_step_kernel<<<grid_for(count, 256), 256, 0, (cudaStream_t)stream>>>(d, eps, count);
    MSTK_CHECK_LAUNCH("ddim_step");
    return MSTK_OK;
}

int mstk_mse(const void* a, const void* b, int dtype, size_t count, double* scratch, float* out, void* stream) {
    MSTK_REQUIRE(a && b && scratch && out && count > 0, "mse: bad args");
    reduce_loss_kernel<<<grid_for(count, 256, 4), 256, 0, (cudaStream_t)stream>>>(a, b, dtype, count, 0, scratch, out);
    MSTK_CHECK_LAUNCH("mse");
    return MSTK_OK;
}
int mstk_kl(const float* z, size_t count, double* scratch, float* out, void* stream) {
    MSTK_REQUIRE(z && scratch && out && count > 0, "kl: bad args");
    reduce_loss_kernel<<<grid_for(count, 256, 4), 256, 0, (cudaStream_t)stream>>>(z, nullptr, MSTK_F32, count, 1, scratch, out);
    MSTK_CHECK_LAUNCH("kl");
    return MSTK_OK;
}
int mstk_feature_std_mse(const void* f, const void* g, int dtype, int b, int64_t inner, double* tmp, float weight, float* out,
                         void* stream) {
    MSTK_REQUIRE(f && g && tmp && out && b > 0 && inner > 1, "feature_std_mse: bad args");
    MSTK_REQUIRE(b <= 65535, "feature_std_mse: batch too large");
    if (feat_vec_ok(f, dtype, inner) && feat_vec_ok(g, dtype, inner)) {
        // vectorised path: per-sample statistics of both maps, then one fused normalise + MSE pass
        cudaStream_t st = (cudaStream_t)stream;
        int rc = launch_sample_stats(f, dtype, b, inner, tmp, st);
        if (rc != MSTK_OK) return rc;
        rc = launch_sample_stats(g, dtype, b, inner, tmp + 2 * (size_t)b, st);
        if (rc != MSTK_OK) return rc;
        return launch_feat_mse3(f, g, nullptr, dtype, b, inner, tmp, tmp + 2 * (size_t)b, nullptr, tmp + 4 * (size_t)b, weight, out, nullptr, st);
    }
    unsigned gx = grid_for((size_t)inner, 256, 4);
    const unsigned cap = (unsigned)((num_sms() * 8 + b - 1) / b);
    if (gx > cap) gx = cap < 1 ? 1 : cap;
    dim3 grid(gx, b);
    feat_stats_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(f, g, dtype, inner, tmp);
    MSTK_CHECK_LAUNCH("feature_std_mse(stats)");
    feat_mse_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(f, g, dtype, b, inner, tmp, tmp + (size_t)b * 4, weight, out);
    MSTK_CHECK_LAUNCH("feature_std_mse(mse)");
    return MSTK_OK;
}
int mstk_sample_stats(const void* f, int dtype, int samples, int64_t inner, double* stats, void* stream) {
    MSTK_REQUIRE(f && stats && samples > 0 && samples <= 65535 && inner > 1, "sample_stats: bad args");
    MSTK_REQUIRE(feat_vec_ok(f, dtype, inner), "sample_stats: inner must be a multiple of 8 (bf16) / 4 (fp32) and f 16-byte aligned");
    return launch_sample_stats(f, dtype, samples, inner, stats, (cudaStream_t)stream);
}
int mstk_feature_mse3(const void* f0, const void* f1, const void* f2, int dtype, int b, int64_t inner, double* st0, double* st1,
                      double* st2, double* scratch, float weight, float* out_a, float* out_c, void* stream) {
    MSTK_REQUIRE(f0 && f1 && st0 && st1 && scratch && out_a && b > 0 && b <= 65535 && inner > 1, "feature_mse3: bad args");
    MSTK_REQUIRE(!f2 || (st2 && out_c), "feature_mse3: the third map needs its statistics and an output");
    MSTK_REQUIRE(feat_vec_ok(f0, dtype, inner) && feat_vec_ok(f1, dtype, inner) && (!f2 || feat_vec_ok(f2, dtype, inner)),
                 "feature_mse3: inner must be a multiple of 8 (bf16) / 4 (fp32) and the maps 16-byte aligned");
    return launch_feat_mse3(f0, f1, f2, dtype, b, inner, st0, st1, st2, scratch, weight, out_a, out_c, (cudaStream_t)stream);
}
int mstk_feature_fused3(const void* f, void* pooled, int dtype, int b, int h, int w, int c, double* acc, double* scratch, float weight,
                        float* out_a, float* out_c, void* stream) {
    MSTK_REQUIRE(f && acc && scratch && out_a && out_c && b > 0 && b <= 65535 && h > 0 && w > 0 && c > 0, "feature_fused3: bad args");
    MSTK_REQUIRE(dtype == MSTK_BF16 || dtype == MSTK_F32, "feature_fused3: bad dtype");
    const int v = dtype == MSTK_BF16 ? 8 : 4;
    MSTK_REQUIRE((c % v) == 0 && (long long)h * w * (c / v) < (1ll << 31), "feature_fused3: c must be a multiple of %d", v);
    MSTK_REQUIRE(!(reinterpret_cast<uintptr_t>(f) & 15) && !(reinterpret_cast<uintptr_t>(pooled) & 15), "feature_fused3: 16-byte alignment");
    MSTK_REQUIRE(!pooled || (!(h & 1) && !(w & 1)), "feature_fused3: pooling needs even h, w");
    return launch_feat_fused3(f, pooled, dtype, b, h, w, c, acc, scratch, weight, out_a, out_c, (cudaStream_t)stream);
}
int mstk_feature_first3(const float* img, const float* wpacked, const float* bias, void* pooled, int pooled_dtype, int b, int h, int w,
                        double* acc, double* scratch, float weight, float* out_a, float* out_c, void* stream) {
    MSTK_REQUIRE(img && wpacked && pooled && acc && scratch && out_a && out_c && b > 0 && b <= 65535, "feature_first3: bad args");
    MSTK_REQUIRE(pooled_dtype == MSTK_BF16 || pooled_dtype == MSTK_F32, "feature_first3: bad dtype");
    MSTK_REQUIRE(h > 1 && w > 1 && !(h & 1) && !(w & 1) && (long long)(h / 2) * (w / 2) * 16 < (1ll << 31), "feature_first3: h, w must be even");
    MSTK_REQUIRE(!(reinterpret_cast<uintptr_t>(pooled) & 15), "feature_first3: 16-byte alignment");
    return launch_feat_first3(img, wpacked, bias, pooled, pooled_dtype, b, h, w, acc, scratch, weight, out_a, out_c, (cudaStream_t)stream);
}
int mstk_maxpool2x2(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream) {
    MSTK_REQUIRE(x && y && n > 0 && h > 1 && w > 1 && c > 0, "maxpool: bad args");
    {
        const int v = dtype == MSTK_BF16 ? 8 : 4;
        if ((c % v) == 0 && !(h & 1) && !(w & 1) && n <= 65535 && !(reinterpret_cast<uintptr_t>(x) & 15) && !(reinterpret_cast<uintptr_t>(y) & 15))
            return launch_maxpool_vec(x, y, n, h, w, c, dtype, (cudaStream_t)stream);
    }
    const size_t total = (size_t)n * (h / 2) * (w / 2) * c;
    maxpool2x2_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(x, y, n, h, w, c, dtype);
    MSTK_CHECK_LAUNCH("maxpool");
    return MSTK_OK;
}

}  // extern "C"
