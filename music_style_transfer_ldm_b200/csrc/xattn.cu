// Style cross-attention with both projections composed into the style side.
//
// Reference: CrossAttention.forward, models/model.py:126-160 = nn.MultiheadAttention(E, 4) with Q from the UNet
// features x [L tokens], K = V from the style map y [S tokens]; the output REPLACES the features.  Written out
// (SURVEY A.3), per head h with d = E/4, scale = 1/sqrt(d):
//     score_h[l,s] = scale (Wq_h x_l + bq_h) . (Wk_h y_s + bk_h)
//     out[l]       = bo + sum_h Wo[:, h] sum_s softmax_s(score_h[l, .])[s] (Wv_h y_s + bv_h)
// Everything that touches the style tokens is linear in y_s and independent of x, and the sampling loop evaluates
// the layer 49 times on the SAME style map.  So the style side is projected ONCE per loop by one k1 convolution
// whose weight is the composition computed here (xattn_compose_kernel):
//     K'_h[s] = scale Wq_h^T (Wk_h y_s + bk_h)   in R^E          score_h[l,s] = x_l . K'_h[s] + c_h[s]
//     c_h[s]  = scale bq_h . (Wk_h y_s + bk_h)   in R
//     V'_h[s] = Wo[:, h] (Wv_h y_s + bv_h)       in R^E          out[l] = bo + sum_{h,s} P_h[l,s] V'_h[s]
// and each of the 49 evaluations is ONE kernel (xattn_kernel) instead of q-projection GEMM + attention core +
// out-projection GEMM: per sample two small contractions over J = 4*S composed rows, [L x E] x [E x J] and
// [L x J] x [J x E] (0.5 MFLOP x 2 at E = 256, L = S = 16; less than the 2 x 16 x 256 x 256 of the two projections
// they replace).  Same function of (x, y, weights) as the reference; only the association order differs.
//
// Layout of one projected style token (ctot = 2*heads*E + 16 channels, activation dtype):
//     [h*E, (h+1)*E)                      K'_h        h = 0..heads-1
//     [heads*E + h*E, heads*E + (h+1)*E)  V'_h
//     [2*heads*E + h]                     c_h         (+ zero padding to a multiple of 16)
//
// xattn_kernel: one CTA (256 threads) per (sample, tile of LT query tokens).  fp32 accumulation on CUDA cores: per
// sample the B operand is private (J x E values used by LT <= 16 rows), so there is no M = 128 tcgen05 tile to fill;
// the kernel is bound by the latency of staging 2 x J x E values from L2, not by its 1 MFLOP.
//   pass 1  K' rows staged chunk-wise in shared memory (16-byte loads), scores[l][j] with x broadcast from smem
//   softmax per (l, head) over the S style tokens, 4 threads per row
//   pass 2  V' rows staged chunk-wise, out[l][e] accumulated in registers (thread = channel), + bo, stored
#include "common.cuh"
#include "tc_ptx.cuh"

namespace mstk {

constexpr int XA_THREADS = 256;

struct XattnP {
    const void* x; const void* kvp; const float* bias; void* out;
    int b, l, s, e, heads, ctot;
    int jc;         // rows of K' / V' staged per chunk
    int jcw;        // power of two >= jc, <= 256: pass-1 thread mapping (j = tid % jcw, query group = tid / jcw)
    int kv_static;  // the projected style map was written >= 2 launches ago: stage the first chunk before the grid dependency
};

template <typename T> struct XaVec;
template <> struct XaVec<float> { static constexpr int N = 4; };
template <> struct XaVec<__nv_bfloat16> { static constexpr int N = 8; };

template <typename T> __device__ __forceinline__ void xa_unpack(const uint4& q, float (&f)[XaVec<T>::N]);
template <> __device__ __forceinline__ void xa_unpack<float>(const uint4& q, float (&f)[4]) {
    f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
}
template <> __device__ __forceinline__ void xa_unpack<__nv_bfloat16>(const uint4& q, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(h[i]);
        f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
}

// stage rows [j0, j0 + n) of the composed K' (off = 0) or V' (off = heads*E) block of sample `bi` into ch[n][E + PAD]
template <typename T>
__device__ __forceinline__ void xa_stage(const XattnP& p, const int bi, const int off, const int j0, const int n, T* ch, const int pitch) {
    constexpr int V = XaVec<T>::N;
    const int vpr = p.e / V;                       // 16-byte vectors per row
    const T* base = reinterpret_cast<const T*>(p.kvp);
    for (int i = threadIdx.x; i < n * vpr; i += XA_THREADS) {
        const int jl = i / vpr, v = i - jl * vpr;
        const int j = j0 + jl;
        const int h = j / p.s, s = j - h * p.s;
        const uint4 q = *reinterpret_cast<const uint4*>(base + ((size_t)bi * p.s + s) * p.ctot + off + h * p.e + v * V);
        *reinterpret_cast<uint4*>(ch + (size_t)jl * pitch + v * V) = q;
    }
}

template <typename T, int LT, int EPT>
__global__ void __launch_bounds__(XA_THREADS) xattn_kernel(const XattnP p) {
    constexpr int V = XaVec<T>::N;
    extern __shared__ __align__(16) uint8_t xa_smem[];
    const int E = p.e, S = p.s, J = p.heads * p.s;
    const int pitch = E + V;                                   // 16 bytes of padding per staged row: conflict-free 16-byte row reads
    float* xs = reinterpret_cast<float*>(xa_smem);             // [LT][E]
    float* ps = xs + LT * E;                                   // [J][LT] scores, then probabilities
    float* cs = ps + (size_t)J * LT;                           // [J] score bias c_h[s]
    T* ch = reinterpret_cast<T*>(cs + ((J + 3) & ~3));         // [jc][pitch]
    const int tid = threadIdx.x;
    const int tiles = (p.l + LT - 1) / LT;
    const int bi = blockIdx.x / tiles, l0 = (blockIdx.x - bi * tiles) * LT;
    const int lvalid = min(LT, p.l - l0);

    const int n0 = min(p.jc, J);
    if (p.kv_static) xa_stage<T>(p, bi, 0, 0, n0, ch, pitch);
    pdl_wait();
    if (tid == 0) pdl_launch_dependents();
    if (!p.kv_static) xa_stage<T>(p, bi, 0, 0, n0, ch, pitch);
    {   // queries of this tile -> fp32 smem (rows past the end of the sample are zero), score biases
        const T* xg = reinterpret_cast<const T*>(p.x) + ((size_t)bi * p.l + l0) * E;
        const int vpr = E / V;
        for (int i = tid; i < LT * vpr; i += XA_THREADS) {
            const int l = i / vpr, v = i - l * vpr;
            float f[V];
            if (l < lvalid) {
                xa_unpack<T>(*reinterpret_cast<const uint4*>(xg + (size_t)l * E + v * V), f);
            } else {
#pragma unroll
                for (int k = 0; k < V; ++k) f[k] = 0.f;
            }
#pragma unroll
            for (int k = 0; k < V; k += 4) *reinterpret_cast<float4*>(xs + l * E + v * V + k) = make_float4(f[k], f[k + 1], f[k + 2], f[k + 3]);
        }
        const T* cg = reinterpret_cast<const T*>(p.kvp) + (size_t)bi * S * p.ctot + 2 * p.heads * E;
        for (int j = tid; j < J; j += XA_THREADS) {
            const int h = j / S, s = j - h * S;
            cs[j] = to_f(cg[(size_t)s * p.ctot + h]);
        }
    }
    __syncthreads();

    // ---- pass 1: scores[l][j] = x_l . K'_j + c_j ----
    {
        const int jl = tid & (p.jcw - 1), lg = tid / p.jcw, nlg = XA_THREADS / p.jcw;
        for (int j0 = 0; j0 < J; j0 += p.jc) {
            const int n = min(p.jc, J - j0);
            if (j0 > 0) {
                __syncthreads();
                xa_stage<T>(p, bi, 0, j0, n, ch, pitch);
                __syncthreads();
            }
            if (jl < n) {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                const T* kr = ch + (size_t)jl * pitch;
                for (int e0 = 0; e0 < E; e0 += V) {
                    float kf[V];
                    xa_unpack<T>(*reinterpret_cast<const uint4*>(kr + e0), kf);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int l = lg + nlg * i;
                        if (l < LT) {
                            const float* xr = xs + l * E + e0;
#pragma unroll
                            for (int k = 0; k < V; k += 4) {
                                const float4 xv = *reinterpret_cast<const float4*>(xr + k);
                                acc[i] = fmaf(xv.x, kf[k], acc[i]); acc[i] = fmaf(xv.y, kf[k + 1], acc[i]);
                                acc[i] = fmaf(xv.z, kf[k + 2], acc[i]); acc[i] = fmaf(xv.w, kf[k + 3], acc[i]);
                            }
                        }
                    }
                }
                const float c = cs[j0 + jl];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int l = lg + nlg * i;
                    if (l < LT) ps[(size_t)(j0 + jl) * LT + l] = acc[i] + c;
                }
            }
        }
    }
    __syncthreads();
    // V' chunk 0 can land while the softmax runs (ch is free: every pass-1 read is behind the barrier above)
    xa_stage<T>(p, bi, p.heads * E, 0, n0, ch, pitch);

    // ---- softmax over the S style tokens of each (query, head): 4 threads per row ----
    {
        const int nrows = LT * p.heads, sub = tid & 3;
        for (int r0 = 0; r0 < nrows; r0 += XA_THREADS / 4) {
            const int row = r0 + (tid >> 2);
            const bool ok = row < nrows;
            const int h = ok ? row / LT : 0, l = ok ? row - h * LT : 0;
            float* pr = ps + (size_t)h * S * LT + l;           // element s at pr[s * LT]
            float mx = -INFINITY;
            if (ok) for (int s = sub; s < S; s += 4) mx = fmaxf(mx, pr[(size_t)s * LT]);
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            float sum = 0.f;
            if (ok) for (int s = sub; s < S; s += 4) {
                const float ev = expf(pr[(size_t)s * LT] - mx);
                pr[(size_t)s * LT] = ev;
                sum += ev;
            }
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            const float inv = 1.f / sum;
            if (ok) for (int s = sub; s < S; s += 4) pr[(size_t)s * LT] *= inv;
        }
    }
    __syncthreads();

    // ---- pass 2: out[l][e] = bo[e] + sum_j P[l][j] V'_j[e]; thread = channel(s), all LT queries in registers ----
    float acc[LT][EPT];
#pragma unroll
    for (int l = 0; l < LT; ++l)
#pragma unroll
        for (int k = 0; k < EPT; ++k) acc[l][k] = 0.f;
    for (int j0 = 0; j0 < J; j0 += p.jc) {
        const int n = min(p.jc, J - j0);
        if (j0 > 0) {
            __syncthreads();
            xa_stage<T>(p, bi, p.heads * E, j0, n, ch, pitch);
            __syncthreads();
        }
        for (int jl = 0; jl < n; ++jl) {
            float v[EPT];
#pragma unroll
            for (int k = 0; k < EPT; ++k) v[k] = to_f(ch[(size_t)jl * pitch + tid + k * XA_THREADS]);
            const float* pj = ps + (size_t)(j0 + jl) * LT;
#pragma unroll
            for (int l4 = 0; l4 < LT; l4 += 4) {
                const float4 pv = *reinterpret_cast<const float4*>(pj + l4);
#pragma unroll
                for (int k = 0; k < EPT; ++k) {
                    acc[l4][k] = fmaf(pv.x, v[k], acc[l4][k]);
                    acc[l4 + 1][k] = fmaf(pv.y, v[k], acc[l4 + 1][k]);
                    acc[l4 + 2][k] = fmaf(pv.z, v[k], acc[l4 + 2][k]);
                    acc[l4 + 3][k] = fmaf(pv.w, v[k], acc[l4 + 3][k]);
                }
            }
        }
    }
    T* og = reinterpret_cast<T*>(p.out) + ((size_t)bi * p.l + l0) * E;
#pragma unroll
    for (int k = 0; k < EPT; ++k) {
        const int e = tid + k * XA_THREADS;
        const float bo = p.bias ? __ldg(p.bias + e) : 0.f;
#pragma unroll
        for (int l = 0; l < LT; ++l)
            if (l < lvalid) og[(size_t)l * E + e] = from_f<T>(acc[l][k] + bo);
    }
}

// Composition of the projection weights (once per parameter version; fp32 accumulation).  Packed conv weight
// [ctot][E] in `dtype` + fp32 bias [ctot]; thread = one (row, column).
__global__ void __launch_bounds__(256) xattn_compose_kernel(const float* __restrict__ w_in, const float* __restrict__ b_in,
                                                           const float* __restrict__ w_out, void* __restrict__ wp,
                                                           float* __restrict__ bias, const int E, const int heads, const int ctot,
                                                           const int dtype) {
    const int r = blockIdx.y;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= E) return;
    const int d = E / heads;
    const float scale = 1.0f / sqrtf((float)d);
    const float* Wq = w_in; const float* Wk = w_in + (size_t)E * E; const float* Wv = w_in + (size_t)2 * E * E;
    const float* bq = b_in; const float* bk = b_in + E; const float* bv = b_in + 2 * E;
    float acc = 0.f, bacc = 0.f;
    if (r < heads * E) {                                   // K'_h row e': scale * sum_j Wq[hd+j][e'] Wk[hd+j][:]
        const int h = r / E, ep = r - h * E;
        for (int j = 0; j < d; ++j) {
            const float a = Wq[(size_t)(h * d + j) * E + ep];
            acc = fmaf(a, Wk[(size_t)(h * d + j) * E + c], acc);
            bacc = fmaf(a, bk[h * d + j], bacc);
        }
        acc *= scale; bacc *= scale;
    } else if (r < 2 * heads * E) {                        // V'_h row e': sum_j Wo[e'][hd+j] Wv[hd+j][:]
        const int rr = r - heads * E, h = rr / E, ep = rr - h * E;
        for (int j = 0; j < d; ++j) {
            const float a = w_out[(size_t)ep * E + h * d + j];
            acc = fmaf(a, Wv[(size_t)(h * d + j) * E + c], acc);
            bacc = fmaf(a, bv[h * d + j], bacc);
        }
    } else if (r < 2 * heads * E + heads) {                // c_h: scale * sum_j bq[hd+j] Wk[hd+j][:]
        const int h = r - 2 * heads * E;
        for (int j = 0; j < d; ++j) {
            const float a = bq[h * d + j];
            acc = fmaf(a, Wk[(size_t)(h * d + j) * E + c], acc);
            bacc = fmaf(a, bk[h * d + j], bacc);
        }
        acc *= scale; bacc *= scale;
    }                                                       // padding rows stay zero
    store_from_f(wp, dtype, (size_t)r * E + c, acc);
    if (c == 0) bias[r] = bacc;
}

template <typename T, int LT, int EPT>
static int xattn_launch(const XattnP& p, size_t smem, int grid, cudaStream_t st) {
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(xattn_kernel<T, LT, EPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "xattn smem attr");
        smem_set = 200 * 1024;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(XA_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t le = cudaLaunchKernelEx(&cfg, xattn_kernel<T, LT, EPT>, p);
    if (le != cudaSuccess) return cuda_fail(le, "xattn launch");
    return MSTK_OK;
}

template <typename T>
static int xattn_dispatch(const XattnP& p, size_t smem, int lt, int grid, cudaStream_t st) {
    const int ept = p.e / XA_THREADS;
    if (lt == 4) return ept == 1 ? xattn_launch<T, 4, 1>(p, smem, grid, st) : xattn_launch<T, 4, 2>(p, smem, grid, st);
    return ept == 1 ? xattn_launch<T, 16, 1>(p, smem, grid, st) : xattn_launch<T, 16, 2>(p, smem, grid, st);
}

}  // namespace mstk

using namespace mstk;

extern "C" {

int mstk_xattn_ctot(int e, int heads) { return 2 * heads * e + 16; }

int mstk_xattn_compose(const float* in_proj_weight, const float* in_proj_bias, const float* out_proj_weight, void* w_packed,
                       float* bias, int e, int heads, int dtype, void* stream) {
    MSTK_REQUIRE(in_proj_weight && in_proj_bias && out_proj_weight && w_packed && bias, "xattn_compose: null");
    MSTK_REQUIRE(e > 0 && heads > 0 && heads <= 16 && e % heads == 0, "xattn_compose: bad dims (e=%d heads=%d)", e, heads);
    MSTK_REQUIRE(dtype == MSTK_F32 || dtype == MSTK_BF16, "xattn_compose: bad dtype");
    const int ctot = mstk_xattn_ctot(e, heads);
    dim3 grid((e + 255) / 256, ctot);
    xattn_compose_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in_proj_weight, in_proj_bias, out_proj_weight, w_packed, bias, e, heads,
                                                                ctot, dtype);
    MSTK_CHECK_LAUNCH("xattn_compose");
    return MSTK_OK;
}

int mstk_xattn_fwd(const mstk_xattn_args* a, void* stream) {
    MSTK_REQUIRE(a && a->x && a->kvp && a->out, "xattn: null");
    MSTK_REQUIRE(a->b > 0 && a->l > 0 && a->s > 0 && a->heads > 0, "xattn: bad dims");
    MSTK_REQUIRE(a->e == 256 || a->e == 512, "xattn: embed dim %d unsupported (256 or 512: models/model.py:184-185 hard-codes them)", a->e);
    MSTK_REQUIRE(a->ctot == mstk_xattn_ctot(a->e, a->heads), "xattn: ctot %d does not match the composed layout (%d)", a->ctot,
                 mstk_xattn_ctot(a->e, a->heads));
    MSTK_REQUIRE(a->dtype == MSTK_F32 || a->dtype == MSTK_BF16, "xattn: bad dtype");
    MSTK_REQUIRE(((reinterpret_cast<uintptr_t>(a->x) | reinterpret_cast<uintptr_t>(a->kvp) | reinterpret_cast<uintptr_t>(a->out)) & 15) == 0,
                 "xattn: tensors must be 16-byte aligned");
    XattnP p;
    p.x = a->x; p.kvp = a->kvp; p.bias = a->bias; p.out = a->out;
    p.b = a->b; p.l = a->l; p.s = a->s; p.e = a->e; p.heads = a->heads; p.ctot = a->ctot; p.kv_static = a->kv_static;
    const int J = a->heads * a->s;
    const int esz = a->dtype == MSTK_F32 ? 4 : 2, vec = 16 / esz;
    const int lt = a->l <= 4 ? 4 : 16;
    const int jc_max = a->dtype == MSTK_F32 ? 32 : 64;
    p.jc = J < jc_max ? J : jc_max;
    int jcw = 16;
    while (jcw < p.jc) jcw <<= 1;
    p.jcw = jcw;
    const size_t smem = ((size_t)lt * a->e + (size_t)J * lt + ((J + 3) & ~3)) * sizeof(float) + (size_t)p.jc * (a->e + vec) * esz;
    MSTK_REQUIRE(smem <= 200 * 1024, "xattn: %d style tokens x %d heads do not fit shared memory (%zu bytes)", a->s, a->heads, smem);
    const int tiles = (a->l + lt - 1) / lt;
    const int grid = a->b * tiles;
    return a->dtype == MSTK_F32 ? xattn_dispatch<float>(p, smem, lt, grid, (cudaStream_t)stream)
                                : xattn_dispatch<__nv_bfloat16>(p, smem, lt, grid, (cudaStream_t)stream);
}

}  // extern "C"
