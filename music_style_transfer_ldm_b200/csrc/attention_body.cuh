// Softmax attention core of the stand-alone attention kernel (elementwise.cu).
#pragma once
#include "common.cuh"

namespace mstk {

// ---------------------------------------------------------------------------
// attention core: grid (b*heads, ceil(l / QPB)); K_h and V_h staged in shared memory
// as fp32; one warp per query row.
// ---------------------------------------------------------------------------
struct AttnP {
    const void* q; const void* k; const void* v; void* out;
    int b, l, s, e, heads, qs, ks, vs, os, dtype;
};
constexpr int ATTN_WARPS = 4;
constexpr int ATTN_QPB = 4;    // queries per CTA = one per warp: the UNet's L = 16 / 4 token maps are latency bound, so every
                               // query row gets its own warp (K/V staging is repeated per CTA from L2, 4 KB)

// One (batch*head, query block) unit, executed by all `nwarps` warps of the calling CTA; `sm` needs
// attn_smem_floats(a, nwarps) floats.  Ends with the CTA's threads past their last shared-memory access only after
// the caller synchronises.
__device__ __forceinline__ void attention_unit(const AttnP& a, const int bh, const int qblock, float* sm, const int nwarps) {
    const int d = a.e / a.heads;
    float* Ks = sm;                         // [s][d+1]
    float* Vs = Ks + (size_t)a.s * (d + 1); // [s][d]
    float* Ps = Vs + (size_t)a.s * d;       // [nwarps][s]
    float* Qs = Ps + nwarps * a.s;          // [nwarps][d]
    const int b = bh / a.heads, hd = bh - b * a.heads;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < a.s * d; i += blockDim.x) {
        const int si = i / d, di = i - si * d;
        Ks[si * (d + 1) + di] = load_as_f(a.k, a.dtype, ((size_t)b * a.s + si) * a.ks + hd * d + di);
        Vs[si * d + di] = load_as_f(a.v, a.dtype, ((size_t)b * a.s + si) * a.vs + hd * d + di);
    }
    __syncthreads();
    const float scaling = 1.0f / sqrtf((float)d);
    const int q_begin = qblock * ATTN_QPB;
    const int q_end = min(a.l, q_begin + ATTN_QPB);
    float* pw = Ps + warp * a.s;
    float* qw = Qs + warp * d;
    for (int qi = q_begin + warp; qi < q_end; qi += nwarps) {
        for (int di = lane; di < d; di += 32)
            qw[di] = load_as_f(a.q, a.dtype, ((size_t)b * a.l + qi) * a.qs + hd * d + di) * scaling;
        __syncwarp();
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
        for (int si = lane; si < a.s; si += 32) {
            const float ev = expf(pw[si] - mx);
            pw[si] = ev;
            sum += ev;
        }
        sum = warp_sum(sum);
        const float inv = 1.f / sum;
        __syncwarp();
        for (int di = lane; di < d; di += 32) {
            float acc = 0.f;
            for (int si = 0; si < a.s; ++si) acc = fmaf(pw[si], Vs[si * d + di], acc);
            store_from_f(a.out, a.dtype, ((size_t)b * a.l + qi) * a.os + hd * d + di, acc * inv);
        }
        __syncwarp();
    }
}


__host__ __device__ inline size_t attn_smem_floats(int s, int d, int nwarps) {
    return (size_t)s * (d + 1) + (size_t)s * d + (size_t)nwarps * s + (size_t)nwarps * d;
}

}  // namespace mstk
