// Shared device/host definitions for libmstk (sm_100a only).
#pragma once
#include <initializer_list>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/mstk.h"

namespace mstk {

// ---------------------------------------------------------------------------
// error plumbing (thread-local message, never throws across the ABI)
// ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define MSTK_REQUIRE(cond, ...)                          \
    do {                                                 \
        if (!(cond)) {                                   \
            ::mstk::set_error(__VA_ARGS__);              \
            return MSTK_EINVAL;                          \
        }                                                \
    } while (0)

// true when every pointer is 16-byte aligned (null counts as aligned): the gate of the kernels' 128-bit load / store paths
inline bool aligned16(std::initializer_list<const void*> ps) {
    for (const void* q : ps)
        if (reinterpret_cast<uintptr_t>(q) & 15) return false;
    return true;
}

#define MSTK_CHECK_LAUNCH(what)                                         \
    do {                                                                \
        cudaError_t e__ = cudaGetLastError();                           \
        if (e__ != cudaSuccess) return ::mstk::cuda_fail(e__, what);    \
    } while (0)

// ---------------------------------------------------------------------------
// Tap-gather implicit GEMM description shared by the CUDA-core and the tcgen05
// convolution kernels.
//
// The GEMM is D[m, co] = sum_{tap, ci} X[pixel(m) shifted by tap, ci] * W[tap.slice, co, ci]
// over a "GEMM pixel grid" of Hg x Wg positions per image:
//   conv  k3 s1/s2 p1 : grid = output pixels, input pixel = g*s_in + (dy,dx)
//   conv  k1          : one tap (0,0)
//   convT s2          : one phase per output parity (py,px); grid = input pixels,
//                       output pixel = g*2 + (py,px); each phase uses the 1/2/2/4 (k3)
//                       or 4 (k4) taps that land on that parity.
// ---------------------------------------------------------------------------
struct Tap {
    int16_t dy, dx;   // input-pixel offset
    int32_t slice;    // index of the [cout][cin] weight slice (ky*k+kx)
};
struct Phase {
    int32_t ntaps;
    int32_t oy, ox;   // output offset of this phase
    Tap taps[16];
};

struct DdimP {
    const float* x_cur;
    float* x_next;
    __nv_bfloat16* x_next_lp;
    float* x0_pred;
    float sab_t, s1m_t, sab_n, s1m_n, eta;
};

struct ConvP {
    const void* x;
    const void* w;
    const float* bias;
    void* y;
    const void* add;
    const float* add_nc;
    double* stats;
    DdimP ddim;
    int32_t has_ddim;
    int32_t N, H, W, Cin, Cout;   // input tensor dims, output channels
    int32_t Hg, Wg;               // GEMM pixel grid per image (per phase)
    int32_t Ho, Wo;               // output tensor dims
    int32_t s_in, s_out;
    int32_t nphases;
    int32_t act;
    int32_t affine;
    float out_scale, out_shift;
    int32_t dtype, out_dtype;
    Phase ph[4];
};

// builds ConvP (plan of phases/taps) from the ABI struct; returns MSTK_* code
int build_conv_plan(const mstk_conv_args* a, ConvP* p);

int launch_conv_simt(const ConvP& p, cudaStream_t st);
// HBM-streaming kernels for Cin == 1 (k3) and Cout == 1 (convT k4) layers
bool conv_edge_supported(const ConvP& p);
int launch_conv_edge(const ConvP& p, cudaStream_t st);
// returns MSTK_EUNSUPPORTED when the shape is outside the tcgen05 kernel's domain
int launch_conv_tc(const ConvP& p, void* workspace, size_t workspace_bytes, cudaStream_t st);
bool conv_tc_supported(const ConvP& p);
// halo-resident persistent kernel for the throughput-bound 3x3 stride-1 layers (conv_halo.cu)
bool conv_halo_supported(const ConvP& p);
int launch_conv_halo(const ConvP& p, cudaStream_t st);
size_t conv_tc_workspace_bytes(const ConvP& p);
bool tc_init();
// vectorised feature-loss kernels (feature_loss.cu)
bool feat_vec_ok(const void* f, int dtype, long long inner);
int launch_sample_stats(const void* f, int dtype, int samples, long long inner, double* stats, cudaStream_t st);
int launch_feat_mse3(const void* f0, const void* f1, const void* f2, int dtype, int b, long long inner, double* st0, double* st1,
                     double* st2, double* scratch, float weight, float* out_a, float* out_c, cudaStream_t st);
int launch_feat_fused3(const void* f, void* pooled, int dtype, int b, int h, int w, int c, double* acc, double* scratch, float weight,
                       float* out_a, float* out_c, cudaStream_t st);
int launch_feat_first3(const float* img, const float* wgt, const float* bias, void* pooled, int dtype, int b, int h, int w, double* acc,
                       double* scratch, float weight, float* out_a, float* out_c, cudaStream_t st);
int launch_maxpool_vec(const void* x, void* y, int n, int h, int w, int c, int dtype, cudaStream_t st);
// tcgen05 weight gradient (wgrad_tc.cu); `p` = forward plan with p.x = input, p.y = dy
bool wgrad_tc_supported(const ConvP& p, int x_dtype, int dy_dtype);
int launch_wgrad_tc(const ConvP& p, float* dw, long long s_co, long long s_ci, cudaStream_t st);
void tc_set_debug(long long* p);
void tc_set_force(int bn, int sk, int stages, int pdl);
void tc_set_persist(int on);
const int* tc_last_config();

// ---------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float load_as_f(const void* p, int dtype, size_t i) {
    return dtype == MSTK_F32 ? reinterpret_cast<const float*>(p)[i]
                             : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
}
__device__ __forceinline__ void store_from_f(void* p, int dtype, size_t i, float v) {
    if (dtype == MSTK_F32) reinterpret_cast<float*>(p)[i] = v;
    else reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
}

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == MSTK_ACT_RELU) return fmaxf(v, 0.f);
    if (act == MSTK_ACT_TANH) return tanhf(v);
    return v;
}

// Channel index of element `first`, `first + stride`, ... of a [rows, c] tensor without a 64-bit modulo per
// iteration (a 64-bit div/mod is ~100 dependent instructions; the streaming kernels do 8-72 FMAs per item).
struct ChanWalk {
    int ch, step, c;
    __device__ __forceinline__ ChanWalk(size_t first_elem, size_t stride_elems, int c_) : c(c_) {
        ch = (int)(first_elem % (size_t)c_);
        step = (int)(stride_elems % (size_t)c_);
    }
    __device__ __forceinline__ void next() {
        ch += step;
        if (ch >= c) ch -= c;
    }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---------------------------------------------------------------------------
// The epilogue shared by both conv kernels: NV contiguous output channels of ONE
// GEMM row (one output pixel).  `v` holds acc (+ nothing else yet).
//   n_img/hg/wg : GEMM pixel of the row; c0: first channel; nvalid: channels < Cout
// Statistics are NOT handled here (callers reduce them across rows first).
// ---------------------------------------------------------------------------
template <int NV>
__device__ __forceinline__ void epilogue_row(const ConvP& p, const Phase& ph, int n_img, int hg, int wg,
                                             int c0, int nvalid, float (&v)[NV]) {
    const int oh = hg * p.s_out + ph.oy;
    const int ow = wg * p.s_out + ph.ox;
    const size_t off = (((size_t)n_img * p.Ho + oh) * p.Wo + ow) * (size_t)p.Cout + c0;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        float x = apply_act(v[j], p.act);
        if (p.affine) x = x * p.out_scale + p.out_shift;
        v[j] = x;
    }
    if (p.add_nc) {
        const float* t = p.add_nc + (size_t)n_img * p.Cout + c0;
#pragma unroll
        for (int j = 0; j < NV; ++j)
            if (j < nvalid) v[j] += t[j];
    }
    if (p.add) {
        if (p.dtype == MSTK_F32) {
            const float* t = reinterpret_cast<const float*>(p.add) + off;
            if (NV % 4 == 0 && nvalid == NV && (off & 3) == 0) {
#pragma unroll
                for (int j = 0; j < NV; j += 4) {
                    float4 q = *reinterpret_cast<const float4*>(t + j);
                    v[j] += q.x; v[j + 1] += q.y; v[j + 2] += q.z; v[j + 3] += q.w;
                }
            } else {
#pragma unroll
                for (int j = 0; j < NV; ++j)
                    if (j < nvalid) v[j] += t[j];
            }
        } else {
            const __nv_bfloat16* t = reinterpret_cast<const __nv_bfloat16*>(p.add) + off;
            if (NV % 8 == 0 && nvalid == NV && (off & 7) == 0) {
#pragma unroll
                for (int j = 0; j < NV; j += 8) {
                    uint4 q = *reinterpret_cast<const uint4*>(t + j);
                    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float2 f = __bfloat1622float2(h[e]);
                        v[j + 2 * e] += f.x; v[j + 2 * e + 1] += f.y;
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < NV; ++j)
                    if (j < nvalid) v[j] += __bfloat162float(t[j]);
            }
        }
    }
    if (p.has_ddim) {
        const DdimP& d = p.ddim;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            if (j < nvalid) {
                const float eps = v[j];
                const float xc = d.x_cur[off + j];
                const float dt = d.s1m_t * eps;
                const float x0 = (xc - dt) / d.sab_t;
                const float dn = d.s1m_n * eps;
                const float xn = d.sab_n * x0 + dn + d.eta * (dn - dt);
                d.x_next[off + j] = xn;
                if (d.x_next_lp) d.x_next_lp[off + j] = __float2bfloat16_rn(xn);
                if (d.x0_pred) d.x0_pred[off + j] = x0;
            }
        }
    }
    if (p.y) {
        if (p.out_dtype == MSTK_F32) {
            float* t = reinterpret_cast<float*>(p.y) + off;
            if (NV % 4 == 0 && nvalid == NV && (off & 3) == 0) {
#pragma unroll
                for (int j = 0; j < NV; j += 4)
                    *reinterpret_cast<float4*>(t + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
#pragma unroll
                for (int j = 0; j < NV; ++j)
                    if (j < nvalid) t[j] = v[j];
            }
        } else {
            __nv_bfloat16* t = reinterpret_cast<__nv_bfloat16*>(p.y) + off;
            if (NV % 8 == 0 && nvalid == NV && (off & 7) == 0) {
#pragma unroll
                for (int j = 0; j < NV; j += 8) {
                    uint4 q;
                    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
                    for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[j + 2 * e], v[j + 2 * e + 1]);
                    *reinterpret_cast<uint4*>(t + j) = q;
                }
            } else {
#pragma unroll
                for (int j = 0; j < NV; ++j)
                    if (j < nvalid) t[j] = __float2bfloat16_rn(v[j]);
            }
        }
    }
}

}  // namespace mstk
