// Development probe (not part of libmstk): can a tcgen05 K-major SWIZZLE_128B operand START at a row that is not a
// multiple of 8 rows (1024 B) of a TMA-written tile, and can its 8-row groups be spaced by an SBO other than 1024?
// That is what a halo-resident A operand needs: ONE box with the halo per tile, every tap an (dy * rowpitch + dx) * 128 B
// shift of the descriptor's start address.  Tests row offsets 0..9 x SBO {1024, 1280, 2304} x base-offset field {0, (addr>>7)&7}.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I music_style_transfer_ldm_b200/csrc -o tools/halo_probe.bin tools/halo_probe.cu
#include "tc_ptx.cuh"
#include <stdlib.h>
#include <vector>
using namespace mstk;

constexpr int ROWS = 512, NB = 64;
__host__ __device__ inline int a_val(int r, int k) { return ((r * 7 + k * 3) % 17) - 8; }
__host__ __device__ inline int b_val(int n, int k) { return ((n * 5 + k) % 13) - 6; }

struct Test { int j, sbo, bo_mode; };

__global__ void __launch_bounds__(128, 1) probe(const Test* tests, int ntests, float* out) {
    extern __shared__ uint8_t raw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    uint8_t* smem = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
    uint8_t* sa = smem;                       // ROWS x 128 B, SW128 pattern anchored at the 1024-aligned base
    uint8_t* sb = smem + ROWS * 128;          // NB x 128 B
    for (int i = threadIdx.x; i < ROWS * 64; i += 128) {
        const int r = i / 64, k = i % 64;
        const uint32_t off = (uint32_t)r * 128 + (uint32_t)(((k >> 3) ^ (r & 7)) * 16) + (k & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(sa + off) = __float2bfloat16((float)a_val(r, k));
    }
    for (int i = threadIdx.x; i < NB * 64; i += 128) {
        const int r = i / 64, k = i % 64;
        const uint32_t off = (uint32_t)r * 128 + (uint32_t)(((k >> 3) ^ (r & 7)) * 16) + (k & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(sb + off) = __float2bfloat16((float)b_val(r, k));
    }
    if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    fence_async_smem();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(64u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base_s;
    const uint32_t idesc = make_idesc(NB);
    for (int t = 0; t < ntests; ++t) {
        const Test ts = tests[t];
        if (threadIdx.x == 0) {
            const uint32_t start = smem_u32(sa) + (uint32_t)ts.j * 128u;
            uint64_t ad = 0;
            ad |= (uint64_t)((start & 0x3FFFFu) >> 4);
            ad |= (uint64_t)(((uint32_t)ts.sbo >> 4) & 0x3FFF) << 32;
            ad |= 1ull << 46;
            if (ts.bo_mode) ad |= (uint64_t)((start >> 7) & 7u) << 49;
            ad |= 2ull << 61;
            const uint64_t bd = make_smem_desc(smem_u32(sb), 64);
            for (int k = 0; k < 4; ++k) umma_bf16(tm, ad + 2u * k, bd + 2u * k, idesc, k ? 1u : 0u);
            umma_commit(&bar);
        }
        mbar_wait(&bar, (uint32_t)(t & 1));
        tc_fence_after();
        const int row = warp * 32 + lane;
        for (int c = 0; c < NB; c += 16) {
            float v[16];
            tmem_ld16(tm + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
            for (int i = 0; i < 16; ++i) out[((size_t)t * 128 + row) * NB + c + i] = v[i];
        }
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(64u) : "memory");
}

int main() {
    std::vector<Test> tests;
    const int sbos[3] = {1024, 1280, 2304};
    for (int bo = 0; bo < 2; ++bo)
        for (int s = 0; s < 3; ++s)
            for (int j = 0; j < 10; ++j) tests.push_back({j, sbos[s], bo});
    const int nt = (int)tests.size();
    Test* dt; float* dout;
    cudaMalloc(&dt, nt * sizeof(Test)); cudaMalloc(&dout, (size_t)nt * 128 * NB * sizeof(float));
    cudaMemcpy(dt, tests.data(), nt * sizeof(Test), cudaMemcpyHostToDevice);
    const int smem = ROWS * 128 + NB * 128 + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe<<<1, 128, smem>>>(dt, nt, dout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<float> out((size_t)nt * 128 * NB);
    cudaMemcpy(out.data(), dout, out.size() * sizeof(float), cudaMemcpyDeviceToHost);
    for (int t = 0; t < nt; ++t) {
        int bad = 0, firstbad = -1;
        for (int r = 0; r < 128; ++r) {
            const int ar = tests[t].j + (r / 8) * (tests[t].sbo / 128) + (r % 8);
            for (int n = 0; n < NB; ++n) {
                float ref = 0;
                for (int k = 0; k < 64; ++k) ref += (float)(a_val(ar, k) * b_val(n, k));
                if (out[((size_t)t * 128 + r) * NB + n] != ref) { if (!bad) firstbad = r; ++bad; }
            }
        }
        printf("j=%d sbo=%4d base_offset=%s : %s (%d mismatches, first bad row %d)\n", tests[t].j, tests[t].sbo, tests[t].bo_mode ? "(addr>>7)&7" : "0",
               bad ? "WRONG" : "ok", bad, firstbad);
    }
    return 0;
}
