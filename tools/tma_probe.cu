// Development probe (not part of libmstk): how many cycles does one 4-D TMA box load of the conv kernel's A
// operand cost per SM, as a function of the box geometry and of the number of loads in flight?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_probe tools/tma_probe.cu -lcuda && /tmp/tma_probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cudaTypedefs.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(b) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    const uint32_t a = smem_u32(bar);
    long long t0 = clock64();
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (clock64() - t0 > 2000000000ll) __trap();
    }
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, uint32_t dst, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
                 "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}

struct Probe {
    int stages, iters, box_bytes, n_tiles, w_off, h_off, multi_thread;
};

// thread 0 (or thread s for stage s when multi_thread) keeps `stages` loads in flight; reports cycles per load
__global__ void __launch_bounds__(128, 1) probe_kernel(const __grid_constant__ CUtensorMap map, const Probe p, long long* out, long long* out_issue) {
    extern __shared__ uint8_t raw[];
    __shared__ __align__(8) uint64_t bar[8];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    if (threadIdx.x == 0) {
        for (int s = 0; s < 8; ++s) mbar_init(&bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int nt0 = (blockIdx.x * 7) % p.n_tiles;
    long long t0 = 0, t1 = 0, issue = 0;
    if (!p.multi_thread) {
        if (threadIdx.x == 0) {
            t0 = clock64();
            for (int it = 0; it < p.iters + p.stages; ++it) {
                const int s = it % p.stages;
                if (it >= p.stages) mbar_wait(&bar[s], ((it / p.stages) - 1) & 1);
                if (it < p.iters) {
                    mbar_expect_tx(&bar[s], (uint32_t)p.box_bytes);
                    const long long i0 = clock64();
                    tma_load_4d(&map, &bar[s], smem_u32(smem) + s * ((p.box_bytes + 1023) & ~1023), 0, p.w_off, p.h_off, ((nt0 + it) % p.n_tiles));
                    issue += clock64() - i0;
                }
            }
            t1 = clock64();
            if (out_issue) out_issue[blockIdx.x] = issue;
        }
    } else if (p.multi_thread == 3) {
        // lock-step warp (the conv kernel's producer): every lane runs the loop, lane == stage issues
        if (threadIdx.x < 32) {
            t0 = clock64();
            for (int it = 0; it < p.iters + p.stages; ++it) {
                const int s = it % p.stages;
                if (it >= p.stages) mbar_wait(&bar[s], ((it / p.stages) - 1) & 1);
                if (it < p.iters && (int)threadIdx.x == s) {
                    mbar_expect_tx(&bar[s], (uint32_t)p.box_bytes);
                    tma_load_4d(&map, &bar[s], smem_u32(smem) + s * ((p.box_bytes + 1023) & ~1023), 0, p.w_off, p.h_off, ((nt0 + it) % p.n_tiles));
                }
            }
            t1 = clock64();
        }
    } else {
        // multi_thread 1: one warp per stage, lane 0 issues; 2: lane s of warp 0 owns stage s
        const int s = p.multi_thread == 1 ? (threadIdx.x >> 5) : (int)threadIdx.x;
        if ((p.multi_thread == 1 ? (threadIdx.x & 31) == 0 : threadIdx.x < 32) && s < p.stages) {
            t0 = clock64();
            int k = 0;
            for (int it = s; it < p.iters; it += p.stages, ++k) {
                if (k > 0) mbar_wait(&bar[s], (k - 1) & 1);
                mbar_expect_tx(&bar[s], (uint32_t)p.box_bytes);
                tma_load_4d(&map, &bar[s], smem_u32(smem) + s * ((p.box_bytes + 1023) & ~1023), 0, p.w_off, p.h_off, ((nt0 + it) % p.n_tiles));
            }
            if (k > 0) mbar_wait(&bar[s], (k - 1) & 1);
            t1 = clock64();
        }
    }
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

static PFN_cuTensorMapEncodeTiled_v12000 g_encode;

static void run(const char* name, int N, int H, int W, int C, int bc, int bw, int bh, int bn, int sw_stride, int w_off, int h_off, int grid) {
    // tensor [N, H, W, C] bf16; view with element stride sw_stride in w and h (parity view when 2)
    const size_t elems = (size_t)N * H * W * C;
    __nv_bfloat16* d;
    cudaMalloc(&d, elems * 2);
    cudaMemset(d, 0, elems * 2);
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)(W / sw_stride), (cuuint64_t)(H / sw_stride), (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2 * sw_stride, (cuuint64_t)W * C * 2 * sw_stride, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUtensorMap map;
    CUresult r = g_encode(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          bc * 2 == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("%s: encode failed %d\n", name, (int)r); return; }
    long long* out;
    cudaMalloc(&out, 148 * sizeof(long long));
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    Probe p;
    p.iters = 64; p.box_bytes = bc * bw * bh * bn * 2; p.n_tiles = N / bn; p.w_off = w_off; p.h_off = h_off;
    printf("%-44s box %5d B (%3d rows):", name, p.box_bytes, bw * bh * bn);
    long long* out_issue;
    cudaMalloc(&out_issue, 148 * sizeof(long long));
    for (int mt = 0; mt < 4; ++mt)
        for (int st = 1; st <= 8; st *= 2) {
            if (mt == 1 && st > 4) continue;
            if (mt >= 2 && st < 2) continue;
            p.stages = st; p.multi_thread = mt;
            for (int rep = 0; rep < 2; ++rep) probe_kernel<<<grid, 128, 8 * ((p.box_bytes + 1023) & ~1023) + 1024>>>(map, p, out, out_issue);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf(" [%s]", cudaGetErrorString(e)); break; }
            long long h[148];
            cudaMemcpy(h, out, grid * sizeof(long long), cudaMemcpyDeviceToHost);
            double avg = 0;
            for (int i = 0; i < grid; ++i) avg += (double)h[i];
            avg /= grid;
            printf("  %s%d:%6.0f", mt == 0 ? "st" : mt == 1 ? "mt" : mt == 2 ? "ml" : "ls", st, avg / p.iters);
            if (mt == 0) {
                cudaMemcpy(h, out_issue, grid * sizeof(long long), cudaMemcpyDeviceToHost);
                printf("(issue %4.0f)", (double)h[0] / p.iters);
            }
        }
    printf("  cyc/load (grid %d)\n", grid);
    cudaFree(d);
    cudaFree(out);
}

int main() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    g_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    for (int grid : {1, 148}) {
        run("vgg   {64,128,1,1} of 128x128x64", 768, 128, 128, 64, 64, 128, 1, 1, 1, 0, 5, grid);
        run("enc1  {32,16,8,1} of 16x16x32 (SW64)", 64, 16, 16, 32, 32, 16, 8, 1, 1, -1, -1, grid);
        run("dec1  {64,16,8,1} of 16x16x64", 64, 16, 16, 64, 64, 16, 8, 1, 1, -1, 0, grid);
        run("enc2  {64,8,8,2} parity view of 16x16x64", 64, 16, 16, 64, 64, 8, 8, 2, 2, 0, 0, grid);
        run("enc3  {64,4,4,8} parity view of 8x8x128", 64, 8, 8, 128, 64, 4, 4, 8, 2, 0, 0, grid);
        run("q2    {64,4,4,8} of 4x4x256", 64, 4, 4, 256, 64, 4, 4, 8, 1, 0, 0, grid);
        run("bott  {64,2,2,32} of 2x2x512", 64, 2, 2, 512, 64, 2, 2, 32, 1, -1, 0, grid);
        run("flat  {64,128,1,1} of 1x128x64 (2-D like)", 4096, 1, 128, 64, 64, 128, 1, 1, 1, 0, 0, grid);
    }
    return 0;
}
