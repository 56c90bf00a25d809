// CUDA-core tap-gather implicit GEMM (fp32 accumulate).  This is the exact-fp32 path
// ("fp32 mode", 1e-4 parity: a single-pass TF32 tensor-core product cannot reach it)
// and the kernel for the contractions the tcgen05 path does not take: Cin == 1 first
// layers and the Cout == 1 last decoder layer, which are HBM-bound, not MMA-bound.
//
// Tile: 64 GEMM rows (pixels) x 64 output channels per 256-thread CTA, K step 16,
// 4x4 register micro-tile, K flattened over (tap, cin) so Cin == 1 wastes nothing.
#include "common.cuh"

namespace mstk {

constexpr int BM = 64, BN = 64, BK = 16, LDS = 68;

template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&o)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&o)[4]) {
    float4 q = *reinterpret_cast<const float4*>(p);
    o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&o)[4]) {
    uint2 q = *reinterpret_cast<const uint2*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
    float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}

template <typename T>
__global__ void __launch_bounds__(256) conv_simt_kernel(const ConvP p) {
    __shared__ __align__(16) float As[BK][LDS];
    __shared__ __align__(16) float Bs[BK][LDS];

    const Phase& ph = p.ph[blockIdx.z];
    const int K = ph.ntaps * p.Cin;
    const int M = p.N * p.Hg * p.Wg;
    const int m0 = blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    const int tid = threadIdx.x;

    const T* __restrict__ X = reinterpret_cast<const T*>(p.x);
    const T* __restrict__ Wt = reinterpret_cast<const T*>(p.w);

    // loader mapping: one row (pixel / out-channel) and 4 consecutive k per thread
    const int lrow = tid >> 2;
    const int lk = (tid & 3) * 4;
    const int lm = m0 + lrow;
    const bool lm_ok = lm < M;
    int ln = 0, lh = 0, lw = 0;
    if (lm_ok) {
        ln = lm / (p.Hg * p.Wg);
        const int r = lm - ln * p.Hg * p.Wg;
        lh = r / p.Wg;
        lw = r - lh * p.Wg;
    }
    const int lco = n0 + lrow;
    const bool lco_ok = lco < p.Cout;
    const bool vec = (p.Cin & 3) == 0;

    const int ty = tid >> 4, tx = tid & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += BK) {
        float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
        const int kk = k0 + lk;
        if (vec) {
            if (kk < K) {
                const int tap = kk / p.Cin;
                const int ci = kk - tap * p.Cin;
                const Tap t = ph.taps[tap];
                if (lm_ok) {
                    const int ih = lh * p.s_in + t.dy, iw = lw * p.s_in + t.dx;
                    if (ih >= 0 && ih < p.H && iw >= 0 && iw < p.W)
                        load4<T>(X + (((size_t)ln * p.H + ih) * p.W + iw) * p.Cin + ci, av);
                }
                if (lco_ok) load4<T>(Wt + ((size_t)t.slice * p.Cout + lco) * p.Cin + ci, bv);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int kj = kk + j;
                if (kj < K) {
                    const int tap = kj / p.Cin;
                    const int ci = kj - tap * p.Cin;
                    const Tap t = ph.taps[tap];
                    if (lm_ok) {
                        const int ih = lh * p.s_in + t.dy, iw = lw * p.s_in + t.dx;
                        if (ih >= 0 && ih < p.H && iw >= 0 && iw < p.W)
                            av[j] = to_f(X[(((size_t)ln * p.H + ih) * p.W + iw) * p.Cin + ci]);
                    }
                    if (lco_ok) bv[j] = to_f(Wt[((size_t)t.slice * p.Cout + lco) * p.Cin + ci]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            As[lk + j][lrow] = av[j];
            Bs[lk + j][lrow] = bv[j];
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
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

    // ---- epilogue ----
    const int c0 = n0 + tx * 4;
    int nvalid = p.Cout - c0;
    nvalid = nvalid < 0 ? 0 : (nvalid > 4 ? 4 : nvalid);
    float bias[4] = {0.f, 0.f, 0.f, 0.f};
    if (p.bias)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < nvalid) bias[j] = p.bias[c0 + j];

    float ssum[4] = {0.f, 0.f, 0.f, 0.f}, ssq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = acc[i][j] + bias[j];
        if (m < M) {
            if (p.stats) {
#pragma unroll
                for (int j = 0; j < 4; ++j) { ssum[j] += v[j]; ssq[j] += v[j] * v[j]; }
            }
            if (nvalid > 0) {
                const int n_img = m / (p.Hg * p.Wg);
                const int r = m - n_img * p.Hg * p.Wg;
                const int hg = r / p.Wg;
                const int wg = r - hg * p.Wg;
                epilogue_row<4>(p, ph, n_img, hg, wg, c0, nvalid, v);
            }
        }
    }
    if (p.stats) {
        // reduce the 16 row-groups of the CTA in shared memory, then one double
        // atomic per channel and CTA
        float* rs = &As[0][0];   // [16][64]  (As has 16*68 floats)
        float* rq = &Bs[0][0];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            rs[ty * BN + tx * 4 + j] = ssum[j];
            rq[ty * BN + tx * 4 + j] = ssq[j];
        }
        __syncthreads();
        if (tid < BN) {
            const int c = n0 + tid;
            if (c < p.Cout) {
                double s = 0.0, q = 0.0;
#pragma unroll
                for (int r = 0; r < 16; ++r) { s += (double)rs[r * BN + tid]; q += (double)rq[r * BN + tid]; }
                atomicAdd(p.stats + c, s);
                atomicAdd(p.stats + p.Cout + c, q);
            }
        }
    }
}

int launch_conv_simt(const ConvP& p, cudaStream_t st) {
    const long long M = (long long)p.N * p.Hg * p.Wg;
    MSTK_REQUIRE(M < (1ll << 31), "conv(simt): too many GEMM rows");
    dim3 grid((unsigned)((M + BM - 1) / BM), (unsigned)((p.Cout + BN - 1) / BN), (unsigned)p.nphases);
    MSTK_REQUIRE(grid.y <= 65535, "conv(simt): cout too large");
    if (p.dtype == MSTK_F32) conv_simt_kernel<float><<<grid, 256, 0, st>>>(p);
    else conv_simt_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p);
    MSTK_CHECK_LAUNCH("conv_simt launch");
    return MSTK_OK;
}

}  // namespace mstk
