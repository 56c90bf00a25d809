// mma.sync helpers of the skinny-GEMM kernel of the first VGGish tap (feature_loss.cu):
// m16n8k16 bf16 x bf16 -> fp32, and the hi + lo split that carries fp32 operands with ~16 mantissa bits.
#pragma once
#include "common.cuh"

namespace mstk {

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], const uint32_t b0, const uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16x2(const float lo, const float hi) {
    const __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&t);
}
__device__ __forceinline__ void split_pack(const float x0, const float x1, uint32_t& hi, uint32_t& lo) {
    const __nv_bfloat16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
    const __nv_bfloat162 hh = __halves2bfloat162(h0, h1);
    hi = *reinterpret_cast<const uint32_t*>(&hh);
    lo = pack_bf16x2(x0 - __bfloat162float(h0), x1 - __bfloat162float(h1));
}

}  // namespace mstk
