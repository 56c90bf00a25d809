// Development probe (not part of libmstk): cost of one kernel node in a CUDA graph as a function of the size of its
// parameter block (the tcgen05 conv kernel passes ~3.4 KB: 13 tensor maps + the convolution plan).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/launch_probe.bin tools/launch_probe.cu && tools/launch_probe.bin
#include <cuda_runtime.h>
#include <stdio.h>
template <int N> struct Blob { int v[N / 4]; };
template <int N> __global__ void k(const __grid_constant__ Blob<N> b, int* out) { if (b.v[0] == 12345 && threadIdx.x == 0) out[0] = b.v[N / 4 - 1]; }
template <int N> void run(int grid, bool serial_chain) {
    Blob<N> b; for (int i = 0; i < N / 4; ++i) b.v[i] = i;
    int* out; cudaMalloc(&out, 4);
    cudaStream_t s; cudaStreamCreate(&s);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s, cudaStreamCaptureModeGlobal);
    for (int i = 0; i < 400; ++i) k<N><<<grid, 128, 0, s>>>(b, out);
    cudaStreamEndCapture(s, &g);
    cudaGraphInstantiate(&ge, g, 0);
    for (int i = 0; i < 3; ++i) cudaGraphLaunch(ge, s);
    cudaStreamSynchronize(s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, s);
    for (int i = 0; i < 5; ++i) cudaGraphLaunch(ge, s);
    cudaEventRecord(e1, s); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("params %5d B  grid %4d: %6.2f us per kernel node\n", N, grid, 1e3 * ms / (5 * 400));
}
int main() {
    for (int grid : {1, 148}) {
        run<64>(grid, true); run<512>(grid, true); run<1024>(grid, true); run<2048>(grid, true); run<3456>(grid, true); run<4000>(grid, true);
    }
    return 0;
}
