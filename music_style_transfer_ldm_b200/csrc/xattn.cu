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
//   staging K' / V' rows chunk-wise into shared memory with cp.async (every 16-byte request of a chunk in flight at once:
//           a chunk costs one L2 round trip); the first chunk of both is requested BEFORE the grid dependency resolves
//           when the style side is loop-invariant (kv_static)
//   pass 1  scores[l][j] with x broadcast from smem
//   softmax per (l, head) over the S style tokens, 4 threads per row
//   pass 2  out[l][e] accumulated in registers (thread = channel), + bo, stored
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

__device__ __forceinline__ void xa_cp16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void xa_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void xa_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Asynchronous staging (cp.async, 16 bytes per request, all requests of a chunk in flight together: the kernel is bound
// by L2 round trips, so a chunk must cost ONE of them) of rows [j0, j0 + n) of the composed K' (off = 0) or V'
// (off = heads*E) block of sample `bi` into buf[n][pitch].  rowoff[j] = s*ctot + h*E for composed row j = h*S + s.
template <typename T>
__device__ __forceinline__ void xa_stage(const XattnP& p, const int bi, const int off, const int j0, const int n, T* buf, const int pitch,
                                         const int* rowoff, const int vshift) {
    constexpr int V = XaVec<T>::N;
    const T* base = reinterpret_cast<const T*>(p.kvp) + (size_t)bi * p.s * p.ctot + off;
    const int vmask = (1 << vshift) - 1;
    for (int i = threadIdx.x; i < (n << vshift); i += XA_THREADS) {
        const int jl = i >> vshift, v = i & vmask;
        xa_cp16(smem_u32(buf + (size_t)jl * pitch + v * V), base + rowoff[j0 + jl] + v * V);
    }
}

template <typename T, int LT, int EPT>
__global__ void __launch_bounds__(XA_THREADS) xattn_kernel(const XattnP p) {
    constexpr int V = XaVec<T>::N;
    extern __shared__ __align__(16) uint8_t xa_smem[];
    const int E = p.e, S = p.s, J = p.heads * p.s;
    const int pitch = E + V;                                   // 16 bytes of padding per staged row: conflict-free 16-byte row reads
    const int vshift = 31 - __clz(E / V);                      // E / V is a power of two (E in {256, 512})
    float* xs = reinterpret_cast<float*>(xa_smem);             // [LT][E] queries, fp32
    float* ps = xs + LT * E;                                   // [J][LT] scores, then probabilities
    float* cs = ps + (size_t)J * LT;                           // [J] score bias c_h[s]
    int* rowoff = reinterpret_cast<int*>(cs + ((J + 3) & ~3)); // [J]
    T* xraw = reinterpret_cast<T*>(rowoff + ((J + 3) & ~3));   // [LT][E] queries as stored
    T* bufk = xraw + LT * E;                                   // [jc][pitch] K' chunk
    T* bufv = bufk + (size_t)p.jc * pitch;                     // [jc][pitch] V' chunk
    const int tid = threadIdx.x;
    const int tiles = (p.l + LT - 1) / LT;
    const int bi = blockIdx.x / tiles, l0 = (blockIdx.x - bi * tiles) * LT;
    const int lvalid = min(LT, p.l - l0);
    const int n0 = min(p.jc, J);

    for (int j = tid; j < J; j += XA_THREADS) {
        const int h = j / S, s = j - h * S;
        rowoff[j] = s * p.ctot + h * E;
    }
    __syncthreads();
    if (p.kv_static) {      // the style side does not depend on the preceding launch: both first chunks fly before the grid dependency
        xa_stage<T>(p, bi, 0, 0, n0, bufk, pitch, rowoff, vshift);
        xa_stage<T>(p, bi, p.heads * E, 0, n0, bufv, pitch, rowoff, vshift);
    }
    pdl_wait();
    if (tid == 0) pdl_launch_dependents();
    if (!p.kv_static) {
        xa_stage<T>(p, bi, 0, 0, n0, bufk, pitch, rowoff, vshift);
        xa_stage<T>(p, bi, p.heads * E, 0, n0, bufv, pitch, rowoff, vshift);
    }
    {
        const T* xg = reinterpret_cast<const T*>(p.x) + ((size_t)bi * p.l + l0) * E;
        for (int i = tid; i < (lvalid << vshift); i += XA_THREADS) xa_cp16(smem_u32(xraw + (size_t)i * V), xg + (size_t)i * V);
        xa_commit();
        const T* cg = reinterpret_cast<const T*>(p.kvp) + (size_t)bi * S * p.ctot + 2 * p.heads * E;
        for (int j = tid; j < J; j += XA_THREADS) {
            const int h = j / S, s = j - h * S;
            cs[j] = to_f(cg[(size_t)s * p.ctot + h]);
        }
    }
    xa_wait_all();
    __syncthreads();
    for (int i = tid; i < (LT << vshift); i += XA_THREADS) {   // queries -> fp32 (rows past the end of the sample are zero)
        float f[V];
        if (i < (lvalid << vshift)) {
            xa_unpack<T>(*reinterpret_cast<const uint4*>(xraw + (size_t)i * V), f);
        } else {
#pragma unroll
            for (int k = 0; k < V; ++k) f[k] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < V; k += 4) *reinterpret_cast<float4*>(xs + (size_t)i * V + k) = make_float4(f[k], f[k + 1], f[k + 2], f[k + 3]);
    }
    __syncthreads();

    // ---- pass 1: scores[l][j] = x_l . K'_j + c_j ----
    {
        const int jl = tid & (p.jcw - 1), lg = tid / p.jcw, nlg = XA_THREADS / p.jcw;
        for (int j0 = 0; j0 < J; j0 += p.jc) {
            const int n = min(p.jc, J - j0);
            if (j0 > 0) {
                __syncthreads();
                xa_stage<T>(p, bi, 0, j0, n, bufk, pitch, rowoff, vshift);
                xa_commit();
                xa_wait_all();
                __syncthreads();
            }
            if (jl < n) {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                const T* kr = bufk + (size_t)jl * pitch;
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
            xa_stage<T>(p, bi, p.heads * E, j0, n, bufv, pitch, rowoff, vshift);
            xa_commit();
            xa_wait_all();
            __syncthreads();
        }
        for (int jl = 0; jl < n; ++jl) {
            float v[EPT];
#pragma unroll
            for (int k = 0; k < EPT; ++k) v[k] = to_f(bufv[(size_t)jl * pitch + tid + k * XA_THREADS]);
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

// ---------------------------------------------------------------------------------------------------------------
// bf16 execution on the tensor cores (mma.sync.m16n8k16, fp32 accumulate).  Per CTA M = one tile of 16 queries, so the
// tcgen05 tile (M >= 64, one B operand for all rows) does not apply: every sample brings its own K' / V'.  The CUDA-core
// kernel above spends ~5 k instructions per warp with two warps per scheduler (ncu: issue slots 25-34 % busy, stalled
// on shared-memory latency): 14-18 us.  Here the two contractions are 2 x 128 MMAs per CTA.
//   GEMM1  scores[16 x J]  = X[16 x E] K'^T : A = queries (row-major, ldmatrix), B = K' rows as staged ([j][e] = "col")
//          work unit = (8-column tile of the chunk, K split) so that all 8 warps have MMAs even at J = 16; partial sums
//          meet in shared memory (red.shared.add.f32 onto the score bias c).
//   softmax per (query, head), probabilities written as the bf16 A operand of GEMM2.
//   GEMM2  out[16 x E] = P[16 x J] V' : B = V' rows as staged ([j][e], e contiguous = "row") through ldmatrix.trans;
//          warp w owns output channels [w*E/8, (w+1)*E/8), accumulators stay in registers across chunks.
//   out + bo -> bf16 -> shared memory -> 16-byte coalesced stores.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t (&r)[2], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// stage rows [j0, j0+n) of K' / V' as in xa_stage and zero the rows up to the next multiple of 16 (MMA padding)
__device__ __forceinline__ void xm_stage(const XattnP& p, const int bi, const int off, const int j0, const int n, __nv_bfloat16* buf,
                                         const int pitch, const int* rowoff, const int vshift) {
    xa_stage<__nv_bfloat16>(p, bi, off, j0, n, buf, pitch, rowoff, vshift);
    const int npad = (n + 15) & ~15;
    for (int i = threadIdx.x; i < ((npad - n) << vshift); i += XA_THREADS) {
        const int jl = n + (i >> vshift), v = i & ((1 << vshift) - 1);
        *reinterpret_cast<uint4*>(buf + (size_t)jl * pitch + v * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
}

template <int NT>       // NT = E / 64: 8-channel output tiles per warp in GEMM2
__global__ void __launch_bounds__(XA_THREADS) xattn_mma_kernel(const XattnP p) {
    using T = __nv_bfloat16;
    extern __shared__ __align__(16) uint8_t xa_smem[];
    const int E = p.e, S = p.s, J = p.heads * p.s;
    const int pitch = E + 8;                                   // bf16 elements; 16-byte row skew: conflict-free ldmatrix
    const int vshift = 31 - __clz(E / 8);
    const int Jp = (J + 15) & ~15;                             // K extent of GEMM2 (padded probabilities are zero)
    const int pspitch = Jp + 4, papitch = Jp + 8;
    float* ps = reinterpret_cast<float*>(xa_smem);             // [16][pspitch] scores
    float* cs = ps + 16 * pspitch;                             // [Jp] score bias
    int* rowoff = reinterpret_cast<int*>(cs + Jp);             // [Jp]
    T* pa = reinterpret_cast<T*>(rowoff + Jp);                 // [16][papitch] probabilities (A operand of GEMM2)
    T* xq = pa + 16 * papitch;                                 // [16][pitch] queries (A operand of GEMM1), later the output tile
    T* bufk = xq + 16 * pitch;                                 // [jc][pitch]
    T* bufv = bufk + (size_t)p.jc * pitch;                     // [jc][pitch]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tiles = (p.l + 15) >> 4;
    const int bi = blockIdx.x / tiles, l0 = (blockIdx.x - bi * tiles) << 4;
    const int lvalid = min(16, p.l - l0);
    const int n0 = min(p.jc, J);

    for (int j = tid; j < J; j += XA_THREADS) {
        const int h = j / S, s = j - h * S;
        rowoff[j] = s * p.ctot + h * E;
    }
    __syncthreads();
    if (p.kv_static) {
        xm_stage(p, bi, 0, 0, n0, bufk, pitch, rowoff, vshift);
        xm_stage(p, bi, p.heads * E, 0, n0, bufv, pitch, rowoff, vshift);
    }
    pdl_wait();
    if (tid == 0) pdl_launch_dependents();
    if (!p.kv_static) {
        xm_stage(p, bi, 0, 0, n0, bufk, pitch, rowoff, vshift);
        xm_stage(p, bi, p.heads * E, 0, n0, bufv, pitch, rowoff, vshift);
    }
    {
        const T* xg = reinterpret_cast<const T*>(p.x) + ((size_t)bi * p.l + l0) * E;
        for (int i = tid; i < (16 << vshift); i += XA_THREADS) {
            const int l = i >> vshift, v = i & ((1 << vshift) - 1);
            if (l < lvalid) xa_cp16(smem_u32(xq + (size_t)l * pitch + v * 8), xg + (size_t)l * E + v * 8);
            else *reinterpret_cast<uint4*>(xq + (size_t)l * pitch + v * 8) = make_uint4(0u, 0u, 0u, 0u);
        }
        xa_commit();
        const T* cg = reinterpret_cast<const T*>(p.kvp) + (size_t)bi * S * p.ctot + 2 * p.heads * E;
        for (int j = tid; j < Jp; j += XA_THREADS) {
            float c = 0.f;
            if (j < J) { const int h = j / S, s = j - h * S; c = to_f(cg[(size_t)s * p.ctot + h]); }
            cs[j] = c;
        }
    }
    __syncthreads();                                           // cs visible
    for (int i = tid; i < 16 * Jp; i += XA_THREADS) {          // scores start at the bias
        const int l = i / Jp, j = i - l * Jp;
        ps[l * pspitch + j] = cs[j];
    }
    xa_wait_all();
    __syncthreads();

    // ---- GEMM1 ----
    const int ksteps = E >> 4;
    for (int j0 = 0; j0 < J; j0 += p.jc) {
        const int n = min(p.jc, J - j0);
        if (j0 > 0) {
            __syncthreads();
            xm_stage(p, bi, 0, j0, n, bufk, pitch, rowoff, vshift);
            xa_commit();
            xa_wait_all();
            __syncthreads();
        }
        const int nt = (n + 7) >> 3;                           // 8-column tiles of this chunk
        int ksplit = 1;
        while (nt * ksplit < 8 && ksplit < 8) ksplit <<= 1;    // spread over the 8 warps
        const int kper = ksteps / ksplit;                      // ksteps is 16 or 32
        for (int unit = warp; unit < nt * ksplit; unit += 8) {
            const int tile = unit % nt, kq = unit / nt;
            float c[4] = {0.f, 0.f, 0.f, 0.f};
            const uint32_t a_addr = smem_u32(xq + (size_t)(lane & 15) * pitch + ((lane >> 4) << 3));
            const uint32_t b_addr = smem_u32(bufk + (size_t)(tile * 8 + (lane & 7)) * pitch + (((lane >> 3) & 1) << 3));
            for (int ks = kq * kper; ks < (kq + 1) * kper; ++ks) {
                uint32_t a[4], b[2];
                ldsm_x4(a, a_addr + (uint32_t)ks * 32u);
                ldsm_x2(b, b_addr + (uint32_t)ks * 32u);
                mma_bf16_16816(c, a, b[0], b[1]);
            }
            const int row = lane >> 2, col = j0 + tile * 8 + ((lane & 3) << 1);
            atomicAdd(ps + row * pspitch + col, c[0]);
            atomicAdd(ps + row * pspitch + col + 1, c[1]);
            atomicAdd(ps + (row + 8) * pspitch + col, c[2]);
            atomicAdd(ps + (row + 8) * pspitch + col + 1, c[3]);
        }
    }
    __syncthreads();

    // ---- softmax per (query, head) over S, 4 threads per row; probabilities -> bf16 A operand ----
    {
        const int nrows = 16 * p.heads, sub = tid & 3;
        for (int r0 = 0; r0 < nrows; r0 += XA_THREADS / 4) {
            const int row = r0 + (tid >> 2);
            const bool ok = row < nrows;
            const int l = ok ? row / p.heads : 0, h = ok ? row - l * p.heads : 0;
            const float* pr = ps + l * pspitch + h * S;
            T* po = pa + l * papitch + h * S;
            float mx = -INFINITY;
            if (ok) for (int s = sub; s < S; s += 4) mx = fmaxf(mx, pr[s]);
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            float sum = 0.f;
            if (ok) for (int s = sub; s < S; s += 4) sum += expf(pr[s] - mx);
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            const float inv = 1.f / sum;
            if (ok) for (int s = sub; s < S; s += 4) po[s] = __float2bfloat16_rn(expf(pr[s] - mx) * inv);
        }
        for (int i = tid; i < 16 * (Jp - J); i += XA_THREADS) {            // K padding of GEMM2
            const int l = i / (Jp - J), j = J + i - l * (Jp - J);
            pa[l * papitch + j] = __float2bfloat16_rn(0.f);
        }
    }
    __syncthreads();

    // ---- GEMM2: warp w -> output channels [w*NT*8, (w+1)*NT*8) ----
    float acc[NT][4];
#pragma unroll
    for (int t = 0; t < NT; ++t) { acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f; }
    const int e_w = warp * NT * 8;
    for (int j0 = 0; j0 < J; j0 += p.jc) {
        const int n = min(p.jc, J - j0);
        if (j0 > 0) {
            __syncthreads();
            xm_stage(p, bi, p.heads * E, j0, n, bufv, pitch, rowoff, vshift);
            xa_commit();
            xa_wait_all();
            __syncthreads();
        }
        const int kk_n = (n + 15) >> 4;
        for (int kk = 0; kk < kk_n; ++kk) {
            uint32_t a[4];
            ldsm_x4(a, smem_u32(pa + (size_t)(lane & 15) * papitch + j0 + kk * 16 + ((lane >> 4) << 3)));
            // x4.trans: matrices (k 0-7, n 0-7), (k 8-15, n 0-7), (k 0-7, n 8-15), (k 8-15, n 8-15) of a 16 x 16 block of V'
            const uint32_t v_addr = smem_u32(bufv + (size_t)(kk * 16 + (lane & 15)) * pitch + e_w + ((lane >> 4) << 3));
#pragma unroll
            for (int t = 0; t < NT; t += 2) {
                uint32_t b[4];
                ldsm_x4_trans(b, v_addr + (uint32_t)t * 16u);
                mma_bf16_16816(acc[t], a, b[0], b[1]);
                mma_bf16_16816(acc[t + 1], a, b[2], b[3]);
            }
        }
    }
    // ---- + bo, bf16, through shared memory (the query tile is dead: every GEMM1 read is behind two barriers) ----
    {
        const int row = lane >> 2;
#pragma unroll
        for (int t = 0; t < NT; ++t) {
            const int e = e_w + t * 8 + ((lane & 3) << 1);
            const float b0 = p.bias ? __ldg(p.bias + e) : 0.f, b1 = p.bias ? __ldg(p.bias + e + 1) : 0.f;
            *reinterpret_cast<__nv_bfloat162*>(xq + (size_t)row * pitch + e) = __floats2bfloat162_rn(acc[t][0] + b0, acc[t][1] + b1);
            *reinterpret_cast<__nv_bfloat162*>(xq + (size_t)(row + 8) * pitch + e) = __floats2bfloat162_rn(acc[t][2] + b0, acc[t][3] + b1);
        }
    }
    __syncthreads();
    T* og = reinterpret_cast<T*>(p.out) + ((size_t)bi * p.l + l0) * E;
    for (int i = tid; i < (lvalid << vshift); i += XA_THREADS) {
        const int l = i >> vshift, v = i & ((1 << vshift) - 1);
        *reinterpret_cast<uint4*>(og + (size_t)l * E + v * 8) = *reinterpret_cast<const uint4*>(xq + (size_t)l * pitch + v * 8);
    }
}

template <int NT>
static int xattn_mma_launch(const XattnP& p, size_t smem, int grid, cudaStream_t st) {
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(xattn_mma_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "xattn_mma smem attr");
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
    cudaError_t le = cudaLaunchKernelEx(&cfg, xattn_mma_kernel<NT>, p);
    if (le != cudaSuccess) return cuda_fail(le, "xattn_mma launch");
    return MSTK_OK;
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
    if (a->dtype == MSTK_BF16) {
        // tensor-core execution: query tiles of 16, chunks of at most 64 composed rows (multiples of 16)
        const int Jp = (J + 15) & ~15;
        const size_t fixed_m = (size_t)16 * (Jp + 4) * 4 + (size_t)Jp * 8 + (size_t)16 * (Jp + 8) * 2 + (size_t)16 * (a->e + 8) * 2;
        const size_t row_m = (size_t)(a->e + 8) * 2;
        MSTK_REQUIRE(fixed_m + 2 * 16 * row_m <= 200 * 1024, "xattn: %d style tokens x %d heads do not fit shared memory", a->s, a->heads);
        int jcm = 64;
        while (jcm > 16 && fixed_m + 2 * (size_t)jcm * row_m > 160 * 1024) jcm >>= 1;
        p.jc = Jp < jcm ? Jp : jcm;  // the kernel stages min(jc, J - j0) rows and pads to 16 inside the buffer
        p.jcw = 0;
        const size_t smem_m = fixed_m + 2 * (size_t)p.jc * row_m;
        const int grid_m = a->b * ((a->l + 15) / 16);
        return a->e == 256 ? xattn_mma_launch<4>(p, smem_m, grid_m, (cudaStream_t)stream)
                           : xattn_mma_launch<8>(p, smem_m, grid_m, (cudaStream_t)stream);
    }
    const int esz = a->dtype == MSTK_F32 ? 4 : 2, vec = 16 / esz;
    const int lt = a->l <= 4 ? 4 : 16;
    // rows per staged chunk: as many as fit next to the query tile and the score matrix (two chunk buffers), at most 64
    const size_t fixed = ((size_t)lt * a->e + (size_t)J * lt + 2 * ((J + 3) & ~3)) * sizeof(float) + (size_t)lt * a->e * esz;
    const size_t row = (size_t)(a->e + vec) * esz;
    MSTK_REQUIRE(fixed + 2 * 16 * row <= 200 * 1024, "xattn: %d style tokens x %d heads do not fit shared memory", a->s, a->heads);
    int jc = 64;
    while (jc > 16 && fixed + 2 * (size_t)jc * row > 160 * 1024) jc >>= 1;
    p.jc = J < jc ? J : jc;
    int jcw = 16;
    while (jcw < p.jc) jcw <<= 1;
    p.jcw = jcw;
    const size_t smem = fixed + 2 * (size_t)p.jc * row;
    const int tiles = (a->l + lt - 1) / lt;
    const int grid = a->b * tiles;
    return a->dtype == MSTK_F32 ? xattn_dispatch<float>(p, smem, lt, grid, (cudaStream_t)stream)
                                : xattn_dispatch<__nv_bfloat16>(p, smem, lt, grid, (cudaStream_t)stream);
}

}  // extern "C"
