// One-shot all-reduce of small fp64 vectors over peer-mapped memory (NVLink / NVSwitch P2P stores), for the BatchNorm
// statistic exchanges of the data-parallel training step (5 forward + 4 backward vectors of 2C <= 256 doubles
// per step, each a dependent link of the step's critical path).  An NCCL all-reduce of 2 KB costs a collective launch plus
// its protocol latency (>= 20 us measured); here every rank's single CTA
//   1. PUSHES its vector into slot [rank] of every peer's exchange buffer (plain stores through the peer mapping),
//   2. fences (system scope) and raises a per-(slot, source rank) epoch flag in every peer's buffer,
//   3. spins until the flags of all ranks for this epoch have arrived in its OWN buffer, and
//   4. sums the world's vectors in rank order out of local memory (same order on every rank: bit-identical replicas).
// Buffers are cudaMalloc'ed once per rank and exchanged as CUDA IPC handles (one process per GPU); data slots are
// double-buffered by epoch parity (a rank can run at most one exchange of a slot ahead of its slowest peer).  Epochs come
// from a device-side counter, so the launch is CUDA-graph replayable.  A missing peer traps after ~2 s instead of hanging.
#include "common.cuh"

namespace mstk {

constexpr int PEER_MAX_WORLD = MSTK_PEER_MAX_WORLD, PEER_SLOTS = 32, PEER_MAXN = 512;
constexpr size_t PEER_OFF_FLAGS = 1024, PEER_OFF_DATA = 4096;
constexpr size_t PEER_BYTES = PEER_OFF_DATA + (size_t)PEER_SLOTS * 2 * PEER_MAX_WORLD * PEER_MAXN * sizeof(double);

__device__ __forceinline__ unsigned int* peer_seq(void* b) { return reinterpret_cast<unsigned int*>(b); }
__device__ __forceinline__ unsigned int* peer_flags(void* b, int slot) {
    return reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(b) + PEER_OFF_FLAGS) + slot * PEER_MAX_WORLD;
}
__device__ __forceinline__ double* peer_data(void* b, int slot, int parity, int src) {
    return reinterpret_cast<double*>(reinterpret_cast<uint8_t*>(b) + PEER_OFF_DATA) +
           (((size_t)slot * 2 + parity) * PEER_MAX_WORLD + src) * PEER_MAXN;
}

__global__ void __launch_bounds__(256) peer_allreduce_kernel(double* __restrict__ data, int n, int slot, const mstk_peer_ctx ctx) {
    __shared__ unsigned int s_epoch;
    void* const local = ctx.bufs[ctx.rank];
    if (threadIdx.x == 0) {
        const unsigned int e = peer_seq(local)[slot] + 1u;
        peer_seq(local)[slot] = e;
        s_epoch = e;
    }
    __syncthreads();
    const unsigned int e = s_epoch;
    const int parity = (int)(e & 1u);
    // 1. push
    for (int i = threadIdx.x; i < n * ctx.world; i += blockDim.x) {
        const int p = i / n, j = i - p * n;
        peer_data(ctx.bufs[p], slot, parity, ctx.rank)[j] = data[j];
    }
    __threadfence_system();
    __syncthreads();
    // 2. announce, 3. wait
    if (threadIdx.x < ctx.world) {
        volatile unsigned int* f = peer_flags(ctx.bufs[threadIdx.x], slot) + ctx.rank;
        *f = e;
        volatile unsigned int* mine = peer_flags(local, slot) + threadIdx.x;
        const long long t0 = clock64();
        while ((int)(*mine - e) < 0) {
            if (clock64() - t0 > 4000000000ll) __trap();      // a peer never arrived: fail the launch, never hang the GPU
        }
        __threadfence_system();
    }
    __syncthreads();
    // 4. sum in rank order, reading past L1
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        double s = 0.0;
        for (int p = 0; p < ctx.world; ++p) s += __ldcg(peer_data(local, slot, parity, p) + j);
        data[j] = s;
    }
}

}  // namespace mstk

using namespace mstk;

extern "C" {

size_t mstk_peer_buffer_bytes(void) { return PEER_BYTES; }

int mstk_peer_alloc(void** dev_ptr, void* ipc_handle_out) {
    MSTK_REQUIRE(dev_ptr && ipc_handle_out, "peer_alloc: null");
    static_assert(sizeof(cudaIpcMemHandle_t) == MSTK_PEER_HANDLE_BYTES, "IPC handle size");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, PEER_BYTES);
    if (e != cudaSuccess) return cuda_fail(e, "peer_alloc: cudaMalloc");
    e = cudaMemset(p, 0, PEER_BYTES);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return cuda_fail(e, "peer_alloc: cudaIpcGetMemHandle"); }
    memcpy(ipc_handle_out, &h, sizeof(h));
    *dev_ptr = p;
    return MSTK_OK;
}

int mstk_peer_open(const void* ipc_handle, void** dev_ptr) {
    MSTK_REQUIRE(ipc_handle && dev_ptr, "peer_open: null");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle, sizeof(h));
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return cuda_fail(e, "peer_open: cudaIpcOpenMemHandle");
    *dev_ptr = p;
    return MSTK_OK;
}

int mstk_peer_close(void* dev_ptr) {
    if (!dev_ptr) return MSTK_OK;
    cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
    if (e != cudaSuccess) return cuda_fail(e, "peer_close");
    return MSTK_OK;
}

int mstk_peer_free(void* dev_ptr) {
    if (!dev_ptr) return MSTK_OK;
    cudaError_t e = cudaFree(dev_ptr);
    if (e != cudaSuccess) return cuda_fail(e, "peer_free");
    return MSTK_OK;
}

int mstk_peer_allreduce_f64(double* data, int n, int slot, const mstk_peer_ctx* ctx, void* stream) {
    MSTK_REQUIRE(data && ctx, "peer_allreduce: null");
    MSTK_REQUIRE(n > 0 && n <= PEER_MAXN, "peer_allreduce: n %d outside 1..%d", n, PEER_MAXN);
    MSTK_REQUIRE(slot >= 0 && slot < PEER_SLOTS, "peer_allreduce: slot %d outside 0..%d", slot, PEER_SLOTS - 1);
    MSTK_REQUIRE(ctx->world >= 1 && ctx->world <= PEER_MAX_WORLD && ctx->rank >= 0 && ctx->rank < ctx->world, "peer_allreduce: bad rank / world");
    for (int p = 0; p < ctx->world; ++p) MSTK_REQUIRE(ctx->bufs[p] != nullptr, "peer_allreduce: buffer of rank %d not mapped", p);
    peer_allreduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(data, n, slot, *ctx);
    MSTK_CHECK_LAUNCH("peer_allreduce");
    return MSTK_OK;
}

}  // extern "C"
