// Persistent layer-program kernel: a recorded sequence of tcgen05 convolutions and attention cores (e.g. the
// 49 x 17 layers of a DDIM sampling loop, reference models/model.py:503-559) executed by ONE cooperative launch of
// one CTA per SM.  Every op is tiled over the resident CTAs; ops are separated by a device-wide barrier instead of
// a kernel boundary, so a layer costs its tiles + ~1 us of barrier instead of launch + prologue + drain.
//
// Recording: between mstk_program_begin() and mstk_program_end() the calling thread's mstk_conv2d_fwd (tcgen05
// path) and mstk_attention_fwd calls append their fully resolved launch (tensor maps, tile configuration, epilogue
// arguments, grid) to a caller-owned host table instead of launching.  The caller copies the table to device memory
// and runs any contiguous range of it with mstk_program_run().  No allocation happens inside the library.
//
// Memory ordering between ops: every CTA drains its TMA stores (cp.async.bulk.wait_group 0), fences the async
// proxy and the generic proxy, then arrives on a self-resetting counter; the last arriver bumps a generation word
// the others spin on (bounded: a lost arrival traps instead of hanging the GPU).
#include "conv_tc_body.cuh"
#include "attention_body.cuh"

namespace mstk {

enum : int32_t { PROG_CONV = 1, PROG_ATTN = 2 };

struct alignas(128) ProgOp {
    TcMaps maps;
    TcParams q;
    AttnP attn;
    int32_t kind, flags, gx, gy, gz, smem_bytes, pad0, pad1;
};

__device__ __forceinline__ void grid_sync(unsigned int* sync) {
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // this thread's TMA stores have landed
    asm volatile("fence.proxy.async;" ::: "memory");
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned int* gen_p = sync + 1;
        const unsigned int gen = *gen_p;                             // read BEFORE arriving
        __threadfence();
        if (atomicInc(sync, gridDim.x - 1) == gridDim.x - 1) {
            __threadfence();
            atomicAdd(sync + 1, 1u);
        } else {
            const long long t0 = clock64();
            while (*gen_p == gen) {
                if (clock64() - t0 > 4000000000ll) __trap();
            }
        }
        __threadfence();
    }
    __syncthreads();
    asm volatile("fence.proxy.async;" ::: "memory");
}

__device__ __forceinline__ long long gtimer() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(TC_THREADS, 1)
program_kernel(const ProgOp* __restrict__ ops, const int nops, unsigned int* sync, const int tmem_cols, long long* dbg) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) TcShared sh;
    __shared__ __align__(16) TcParams qs;
    const int warp = threadIdx.x >> 5;

    if (threadIdx.x == 0) {
        for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(&sh.full_bar[s], 2); mbar_init(&sh.empty_bar[s], 1); }
        mbar_init(&sh.accum_bar, 1); mbar_init(&sh.add_bar, 1); mbar_init(&sh.part_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sh.tmem_base)),
                     "r"((uint32_t)tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    // development aid (mstk_debug_program_timeline): CTA 0 and the last CTA stamp %globaltimer at op start, after
    // their tiles, and after the barrier: dbg[(cta01 * nops + op) * 3 + {0,1,2}]
    const bool stamp = dbg != nullptr && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1);
    long long* my_dbg = dbg ? dbg + (size_t)(blockIdx.x == 0 ? 0 : 1) * nops * 3 : nullptr;
    for (int op = 0; op < nops; ++op) {
        const ProgOp& o = ops[op];
        if (stamp) my_dbg[op * 3 + 0] = gtimer();
        if (o.kind == PROG_CONV) {
            // per-op parameters: global -> shared once, then every tile reads them at shared-memory latency
            {
                const uint32_t* src = reinterpret_cast<const uint32_t*>(&o.q);
                uint32_t* dst = reinterpret_cast<uint32_t*>(&qs);
                for (int i = threadIdx.x; i < (int)(sizeof(TcParams) / 4); i += TC_THREADS) dst[i] = __ldg(src + i);
            }
            __syncthreads();
            const int gx = o.gx, gy = o.gy, ntiles = gx * gy * o.gz;
            const uint32_t fl = (uint32_t)o.flags;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                const int bx = t % gx, by = (t / gx) % gy, bz = t / (gx * gy);
                // ONE generic instantiation for every op: inside a persistent kernel its code stays hot in the
                // instruction cache across layers (per-feature instantiations are 10 x 43 KB of cold code)
                (void)fl;
                conv_tc_body<F_ALL, true>(o.maps, qs, bx, by, bz, gx, gy, sh, smem_raw);
            }
        } else if (o.kind == PROG_ATTN) {
            const AttnP a = o.attn;
            const int bh = a.b * a.heads;
            const int units = bh * ((a.l + ATTN_QPB - 1) / ATTN_QPB);
            uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 15) & ~(uintptr_t)15);
            for (int u = blockIdx.x; u < units; u += gridDim.x) {
                attention_unit(a, u % bh, u / bh, reinterpret_cast<float*>(smem), TC_THREADS / 32);
                __syncthreads();
            }
        }
        if (stamp) my_dbg[op * 3 + 1] = gtimer();
        if (op + 1 < nops) grid_sync(sync);
        if (stamp) my_dbg[op * 3 + 2] = gtimer();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(sh.tmem_base), "r"((uint32_t)tmem_cols) : "memory");
    }
}

// ---------------------------------------------------------------------------
// host side: thread-local recorder
// ---------------------------------------------------------------------------
struct Recorder {
    ProgOp* ops = nullptr;
    int max_ops = 0, nops = 0, smem_bytes = 0, tmem_cols = 32, error = 0;
};
static thread_local Recorder g_rec;
static long long* g_prog_dbg = nullptr;

bool prog_recording() { return g_rec.ops != nullptr; }

static ProgOp* prog_next() {
    if (g_rec.nops >= g_rec.max_ops) {
        g_rec.error = 1;
        set_error("program: more than %d ops recorded", g_rec.max_ops);
        return nullptr;
    }
    ProgOp* o = &g_rec.ops[g_rec.nops++];
    memset(static_cast<void*>(o), 0, sizeof(ProgOp));
    return o;
}

int prog_record_conv(const TcMaps& maps, const TcParams& q, uint32_t flags, int gx, int gy, int gz, size_t smem_bytes) {
    ProgOp* o = prog_next();
    if (!o) return MSTK_EINVAL;
    o->kind = PROG_CONV; o->flags = (int32_t)flags; o->gx = gx; o->gy = gy; o->gz = gz; o->smem_bytes = (int32_t)smem_bytes;
    o->maps = maps;
    o->q = q;
    o->q.dbg = q.dbg ? q.dbg + 16 * (size_t)(g_rec.nops - 1) : nullptr;      // development aid: 16 clock64 stamps per op
    if ((int)smem_bytes > g_rec.smem_bytes) g_rec.smem_bytes = (int)smem_bytes;
    if (q.tmem_cols > g_rec.tmem_cols) g_rec.tmem_cols = q.tmem_cols;
    return MSTK_OK;
}

int prog_record_attn(const AttnP& a) {
    const int d = a.e / a.heads;
    const size_t smem = attn_smem_floats(a.s, d, TC_THREADS / 32) * sizeof(float) + 16;
    if (smem > 200 * 1024) {
        g_rec.error = 1;
        set_error("program: attention s*d too large for shared memory (s=%d d=%d)", a.s, d);
        return MSTK_EUNSUPPORTED;
    }
    ProgOp* o = prog_next();
    if (!o) return MSTK_EINVAL;
    o->kind = PROG_ATTN; o->attn = a; o->smem_bytes = (int32_t)smem;
    if ((int)smem > g_rec.smem_bytes) g_rec.smem_bytes = (int)smem;
    return MSTK_OK;
}

}  // namespace mstk

using namespace mstk;

extern "C" {

size_t mstk_program_op_bytes(void) { return sizeof(ProgOp); }
/* development aid (not in mstk.h): per-op %globaltimer stamps of the first and last CTA, 2 * nops * 3 int64 */
int mstk_debug_program_timeline(long long* dev_buf) { g_prog_dbg = dev_buf; return MSTK_OK; }

int mstk_program_begin(void* host_ops, int max_ops) {
    MSTK_REQUIRE(host_ops && max_ops > 0, "program_begin: bad args");
    MSTK_REQUIRE((reinterpret_cast<uintptr_t>(host_ops) & 127) == 0, "program_begin: the host table must be 128-byte aligned");
    MSTK_REQUIRE(!prog_recording(), "program_begin: already recording on this thread");
    MSTK_REQUIRE(tc_init(), "program_begin: tcgen05 path unavailable");
    g_rec = Recorder();
    g_rec.ops = reinterpret_cast<ProgOp*>(host_ops);
    g_rec.max_ops = max_ops;
    return MSTK_OK;
}

int mstk_program_end(mstk_program_info* info) {
    MSTK_REQUIRE(prog_recording(), "program_end: not recording");
    const Recorder r = g_rec;
    g_rec = Recorder();
    MSTK_REQUIRE(info != nullptr, "program_end: null info");
    MSTK_REQUIRE(!r.error, "program_end: recording failed (unsupported op or table overflow)");
    info->nops = r.nops; info->smem_bytes = r.smem_bytes; info->tmem_cols = r.tmem_cols;
    return MSTK_OK;
}

int mstk_program_run(const void* dev_ops, const mstk_program_info* info, int first_op, int nops, unsigned int* dev_sync, void* stream) {
    MSTK_REQUIRE(dev_ops && info && dev_sync, "program_run: null");
    MSTK_REQUIRE(first_op >= 0 && nops > 0 && first_op + nops <= info->nops, "program_run: op range [%d, %d) outside the program of %d ops",
                 first_op, first_op + nops, info->nops);
    MSTK_REQUIRE((reinterpret_cast<uintptr_t>(dev_ops) & 127) == 0, "program_run: the op table must be 128-byte aligned");
    MSTK_REQUIRE(tc_init(), "program_run: tcgen05 path unavailable");
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(program_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(program_kernel)");
        attr_set = true;
    }
    const ProgOp* ops = reinterpret_cast<const ProgOp*>(dev_ops) + first_op;
    int tmem_cols = info->tmem_cols;
    int n = nops;
    long long* dbg = g_prog_dbg;
    void* args[] = {(void*)&ops, (void*)&n, (void*)&dev_sync, (void*)&tmem_cols, (void*)&dbg};
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)tc_num_sms(), 1, 1);
    cfg.blockDim = dim3(TC_THREADS, 1, 1);
    cfg.dynamicSmemBytes = (size_t)info->smem_bytes + 1024;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;      // all CTAs co-resident, or the launch fails (never a deadlock)
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelExC(&cfg, (const void*)program_kernel, args);
    if (e != cudaSuccess) return cuda_fail(e, "program launch");
    return MSTK_OK;
}

}  // extern "C"
