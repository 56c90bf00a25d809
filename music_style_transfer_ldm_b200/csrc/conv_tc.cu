// tcgen05 / TMEM / TMA tap-gather implicit GEMM for sm_100a (bf16 in, fp32 accumulate).
//
// One CTA computes a 128-row x BN-column output tile:
//   rows    = 128 GEMM pixels = a (Wt x Ht x Nt) box of the NHWC pixel grid, so that the
//             A operand of tap (dy,dx) and channel block c is ONE 4-D TMA box load
//             {c*KC.., w0+dx.., h0+dy.., n0..} with hardware zero fill for the padding halo;
//             stride-2 convolutions read one of four parity views (h%2, w%2) of the input,
//             each its own tensor map, so the box stays dense.
//   columns = BN output channels; B operand = [BN x KC] box of the packed weights
//             [tap][cout][cin] (K-major).
//   K loop  = (tap, channel block) pairs, optionally split across gridDim.z (split-K) so
//             that the small-M layers of the UNet still spread over all 148 SMs; partial
//             tiles are reduced with fp32 RED into a zeroed workspace and the last CTA to
//             arrive runs the epilogue and re-zeroes it.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer
// (single thread issues tcgen05.mma, accumulator in TMEM), warps 2..5 = epilogue
// (tcgen05.ld -> bias/act/adds/DDIM -> global).
#include <cuda.h>
#include <cudaTypedefs.h>
#include "common.cuh"

namespace mstk {

constexpr int TC_THREADS = 192;
constexpr int TILE_M = 128;
constexpr int MAX_STAGES = 8;

struct TcParams {
    ConvP p;
    int32_t Wt, Ht, Nt;            // tile box; Wt*Ht*Nt == 128
    int32_t tiles_w, tiles_h, tiles_n;
    int32_t KC;                    // channels per k-block: 64 (SWIZZLE_128B) or 32 (SWIZZLE_64B)
    int32_t BN;                    // output channels per CTA (16..256, multiple of 16)
    int32_t kb_per_tap;            // Cin / KC
    int32_t splitk;
    int32_t stages;
    int32_t tmem_cols;
    float* ws;                     // split-K workspace
    unsigned int* counters;        // split-K tile counters
    long long* dbg;                // optional timeline of CTA (0,0,0): 8 clock64 stamps
};
#define DBG_STAMP(i) do { if (q.dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) q.dbg[i] = clock64(); } while (0)

// ---------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok;
}
// Bounded wait: a pipeline bug must surface as a launch failure (trap), never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    if (mbar_try_wait(a, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(a, parity)) {
        if (clock64() - t0 > 4000000000ll) __trap();   // ~2 s at 2 GHz
    }
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// K-major operand tile in shared memory, rows of KC bf16 (64 -> 128 B rows / SWIZZLE_128B,
// 32 -> 64 B rows / SWIZZLE_64B), 8-row groups SBO bytes apart.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, int kc) {
    const uint32_t sbo = (uint32_t)(8 * kc * 2);                 // 1024 or 512
    const uint64_t layout = (kc == 64) ? 2ull : 4ull;            // SWIZZLE_128B : SWIZZLE_64B
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);                    // start address
    d |= (uint64_t)0 << 16;                                      // LBO (unused for swizzled K-major)
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;                  // SBO
    d |= 1ull << 46;                                             // descriptor version (sm_100)
    d |= layout << 61;
    return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major, M=128, N=n
__device__ __forceinline__ uint32_t make_idesc(int n) {
    uint32_t d = 0;
    d |= 1u << 4;                       // D format f32
    d |= 1u << 7;                       // A format bf16
    d |= 1u << 10;                      // B format bf16
    d |= (uint32_t)(n >> 3) << 17;      // N >> 3
    d |= (uint32_t)(TILE_M >> 4) << 24; // M >> 4
    return d;
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// ---------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
               const __grid_constant__ CUtensorMap mapA2, const __grid_constant__ CUtensorMap mapA3,
               const __grid_constant__ CUtensorMap mapB, const __grid_constant__ TcParams q) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ uint32_t tmem_base_smem;
    __shared__ uint32_t is_last_smem;

    const ConvP& p = q.p;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // tile coordinates
    const int mt = blockIdx.x;
    const int tw = mt % q.tiles_w;
    const int th = (mt / q.tiles_w) % q.tiles_h;
    const int tn = mt / (q.tiles_w * q.tiles_h);
    const int w0 = tw * q.Wt, h0 = th * q.Ht, n0 = tn * q.Nt;
    const int co0 = blockIdx.y * q.BN;
    const int phase_idx = blockIdx.z / q.splitk;
    const int ks = blockIdx.z - phase_idx * q.splitk;
    const Phase& ph = p.ph[phase_idx];
    const int total_kb = ph.ntaps * q.kb_per_tap;
    const int kb_begin = (int)(((long long)ks * total_kb) / q.splitk);
    const int kb_end = (int)(((long long)(ks + 1) * total_kb) / q.splitk);
    const int nkb = kb_end - kb_begin;

    // shared-memory carve-up (1024-byte aligned for the 128B swizzle atom)
    const uint32_t a_bytes = TILE_M * q.KC * 2;
    const uint32_t b_bytes = q.BN * q.KC * 2;
    const uint32_t b_bytes_al = (b_bytes + 1023u) & ~1023u;
    const uint32_t stage_bytes = a_bytes + b_bytes_al;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);

    if (threadIdx.x == 0) {
        DBG_STAMP(0);
        prefetch_tmap(&mapA0);
        prefetch_tmap(&mapB);
        for (int s = 0; s < q.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&accum_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        // TMEM allocation (whole warp), base address lands in shared memory
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)),
                     "r"((uint32_t)q.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    if (threadIdx.x == 0) DBG_STAMP(1);

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            for (int it = 0; it < nkb; ++it) {
                const int s = it % q.stages;
                const uint32_t par = (uint32_t)((it / q.stages) & 1);
                mbar_wait(&empty_bar[s], par ^ 1u);
                const int kb = kb_begin + it;
                const int tap_i = kb / q.kb_per_tap;
                const int cb = kb - tap_i * q.kb_per_tap;
                const Tap t = ph.taps[tap_i];
                uint8_t* sa = smem + (size_t)s * stage_bytes;
                uint8_t* sb = sa + a_bytes;
                mbar_expect_tx(&full_bar[s], a_bytes + b_bytes);
                int dy = t.dy, dx = t.dx;
                const CUtensorMap* mA = &mapA0;
                if (p.s_in == 2) {
                    const int py = dy & 1, px = dx & 1;        // parity view (works for negative offsets)
                    dy = (dy - py) / 2;
                    dx = (dx - px) / 2;
                    const int v = py * 2 + px;
                    mA = (v == 0) ? &mapA0 : (v == 1) ? &mapA1 : (v == 2) ? &mapA2 : &mapA3;
                }
                tma_load_4d(mA, &full_bar[s], sa, cb * q.KC, w0 + dx, h0 + dy, n0);
                tma_load_2d(&mapB, &full_bar[s], sb, cb * q.KC, t.slice * p.Cout + co0);
                if (it == 0) DBG_STAMP(2);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = make_idesc(q.BN);
            for (int it = 0; it < nkb; ++it) {
                const int s = it % q.stages;
                const uint32_t par = (uint32_t)((it / q.stages) & 1);
                mbar_wait(&full_bar[s], par);
                if (it == 0) DBG_STAMP(3);
                if (it == nkb - 1) DBG_STAMP(4);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
                const uint32_t sb = sa + a_bytes;
                const uint64_t adesc = make_smem_desc(sa, q.KC);
                const uint64_t bdesc = make_smem_desc(sb, q.KC);
                const int ksteps = q.KC / 16;
                for (int k = 0; k < ksteps; ++k) {
                    // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in the (>>4) address field
                    umma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[s]);     // frees the smem slot when these MMAs retire
            }
            umma_commit(&accum_bar);            // accumulator complete
        }
    } else {
        // ===================== epilogue (warps 2..5) =====================
        const int qd = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = qd * 32 + lane;          // GEMM row inside the tile
        const int w_l = row % q.Wt;
        const int h_l = (row / q.Wt) % q.Ht;
        const int n_l = row / (q.Wt * q.Ht);
        const int n_img = n0 + n_l, hg = h0 + h_l, wg = w0 + w_l;
        const bool row_ok = n_img < p.N;
        const int et = threadIdx.x - 64;         // 0..127

        mbar_wait(&accum_bar, 0);
        if (et == 0) DBG_STAMP(5);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16);

        if (q.splitk == 1) {
            for (int c = 0; c < q.BN; c += 16) {
                float v[16];
                tmem_ld16(taddr + (uint32_t)c, v);
                const int cg = co0 + c;
                if (p.bias) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += __ldg(p.bias + cg + j);
                }
                if (p.stats) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float x = row_ok ? v[j] : 0.f;
                        const float s1 = warp_sum(x);
                        const float s2 = warp_sum(x * x);
                        if (lane == j) {
                            atomicAdd(p.stats + cg + j, (double)s1);
                            atomicAdd(p.stats + p.Cout + cg + j, (double)s2);
                        }
                    }
                }
                if (row_ok) epilogue_row<16>(p, ph, n_img, hg, wg, cg, 16, v);
            }
        } else {
            // ---- split-K: RED partials into the workspace; last CTA finishes the tile ----
            const size_t tile_rows = (size_t)(phase_idx * gridDim.x + mt) * TILE_M;
            float* wrow = q.ws + (tile_rows + row) * (size_t)p.Cout + co0;
            for (int c = 0; c < q.BN; c += 16) {
                float v[16];
                tmem_ld16(taddr + (uint32_t)c, v);
#pragma unroll
                for (int j = 0; j < 16; ++j) atomicAdd(wrow + c + j, v[j]);
            }
            __threadfence();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            unsigned int* ctr = q.counters + ((size_t)phase_idx * gridDim.x + mt) * gridDim.y + blockIdx.y;
            if (et == 0) {
                const unsigned int old = atomicAdd(ctr, 1u);
                is_last_smem = (old == (unsigned int)(q.splitk - 1)) ? 1u : 0u;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (is_last_smem) {
                __threadfence();
                for (int c = 0; c < q.BN; c += 16) {
                    float v[16];
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        const float4 t = __ldcg(reinterpret_cast<const float4*>(wrow + c + j));
                        v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
                        __stcg(reinterpret_cast<float4*>(wrow + c + j), make_float4(0.f, 0.f, 0.f, 0.f));
                    }
                    const int cg = co0 + c;
                    if (p.bias) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] += __ldg(p.bias + cg + j);
                    }
                    if (p.stats) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float x = row_ok ? v[j] : 0.f;
                            const float s1 = warp_sum(x);
                            const float s2 = warp_sum(x * x);
                            if (lane == j) {
                                atomicAdd(p.stats + cg + j, (double)s1);
                                atomicAdd(p.stats + p.Cout + cg + j, (double)s2);
                            }
                        }
                    }
                    if (row_ok) epilogue_row<16>(p, ph, n_img, hg, wg, cg, 16, v);
                }
                if (et == 0) *ctr = 0u;
            }
        }
        if (et == 0) DBG_STAMP(6);
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)q.tmem_cols) : "memory");
        if (lane == 0) DBG_STAMP(7);
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static int g_tc_state = 0;   // 0 unknown, 1 ok, -1 failed
static int g_num_sms = 148;
static long long* g_dbg = nullptr;
void tc_set_debug(long long* p) { g_dbg = p; }

bool tc_init() {
    if (g_tc_state != 0) return g_tc_state > 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || fn == nullptr) {
        (void)cudaGetLastError();
        set_error("tcgen05 path: cuTensorMapEncodeTiled entry point not available");
        g_tc_state = -1;
        return false;
    }
    g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0)
        g_num_sms = sms;
    g_tc_state = 1;
    return true;
}

static int pow2_divisor(int x, int cap) {
    int d = 1;
    while (d * 2 <= cap && (x % (d * 2)) == 0) d *= 2;
    return d;
}

bool conv_tc_supported(const ConvP& p) {
    if (p.dtype != MSTK_BF16) return false;
    if (p.Cin % 32 != 0 || p.Cout % 16 != 0) return false;
    if (p.s_in == 2 && ((p.H & 1) || (p.W & 1))) return false;
    if ((reinterpret_cast<uintptr_t>(p.x) & 15) || (reinterpret_cast<uintptr_t>(p.w) & 15)) return false;
    return true;
}

struct TcConfig {
    int Wt, Ht, Nt, tiles_w, tiles_h, tiles_n, KC, BN, kb_per_tap, splitk, stages, tmem_cols;
    size_t smem_bytes;
};

static TcConfig choose_config(const ConvP& p) {
    TcConfig c;
    c.Wt = pow2_divisor(p.Wg, TILE_M);
    c.Ht = pow2_divisor(p.Hg, TILE_M / c.Wt);
    c.Nt = TILE_M / (c.Wt * c.Ht);
    c.tiles_w = p.Wg / c.Wt;
    c.tiles_h = p.Hg / c.Ht;
    c.tiles_n = (p.N + c.Nt - 1) / c.Nt;
    c.KC = (p.Cin % 64 == 0) ? 64 : 32;
    c.kb_per_tap = p.Cin / c.KC;
    const long long mtiles = (long long)c.tiles_w * c.tiles_h * c.tiles_n * p.nphases;
    // widest N tile that still yields roughly one CTA per SM
    int bn = 128;
    while (bn > 32 && (p.Cout % bn != 0 || mtiles * (p.Cout / bn) < g_num_sms)) bn >>= 1;
    while (p.Cout % bn != 0) bn >>= 1;       // cout is a multiple of 16
    c.BN = bn;
    long long ctas = mtiles * (p.Cout / bn);
    int min_kb = 1 << 30;
    for (int i = 0; i < p.nphases; ++i) {
        const int kb = p.ph[i].ntaps * c.kb_per_tap;
        if (kb < min_kb) min_kb = kb;
    }
    int sk = 1;
    // split K while the grid is under one wave and every split keeps >= 4 k-blocks
    while (ctas * sk * 2 <= 2 * g_num_sms && min_kb / (sk * 2) >= 4 && sk < 16) sk *= 2;
    c.splitk = sk;
    const size_t a_bytes = (size_t)TILE_M * c.KC * 2;
    const size_t b_bytes = (((size_t)c.BN * c.KC * 2) + 1023) & ~(size_t)1023;
    const size_t stage = a_bytes + b_bytes;
    int st = (int)((200 * 1024) / stage);
    if (st > MAX_STAGES) st = MAX_STAGES;
    if (st < 2) st = 2;
    c.stages = st;
    c.smem_bytes = stage * st + 1024;
    int cols = 32;
    while (cols < c.BN) cols <<= 1;
    c.tmem_cols = cols;
    return c;
}

size_t conv_tc_workspace_bytes(const ConvP& p) {
    if (!tc_init()) return 0;
    const TcConfig c = choose_config(p);
    if (c.splitk == 1) return 0;
    const size_t mtiles = (size_t)c.tiles_w * c.tiles_h * c.tiles_n * p.nphases;
    const size_t ws = mtiles * TILE_M * (size_t)p.Cout * sizeof(float);
    const size_t ctr = mtiles * (size_t)(p.Cout / c.BN) * sizeof(unsigned int);
    return ws + ((ctr + 255) & ~(size_t)255);
}

static int encode_map(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_bytes,
                      const cuuint32_t* box, CUtensorMapSwizzle sw) {
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_bytes, box,
                          estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]", (int)r, rank,
                  (unsigned long long)dims[0], (unsigned long long)dims[1], rank > 2 ? (unsigned long long)dims[2] : 0ull,
                  rank > 3 ? (unsigned long long)dims[3] : 0ull, box[0], box[1], rank > 2 ? box[2] : 0u, rank > 3 ? box[3] : 0u);
        return MSTK_ECUDA;
    }
    return MSTK_OK;
}

int launch_conv_tc(const ConvP& p, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (!tc_init()) return MSTK_EUNSUPPORTED;
    TcConfig c = choose_config(p);
    TcParams q;
    memset(&q, 0, sizeof(q));
    q.p = p;
    const size_t mtiles = (size_t)c.tiles_w * c.tiles_h * c.tiles_n * p.nphases;
    if (c.splitk > 1) {
        const size_t ws = mtiles * TILE_M * (size_t)p.Cout * sizeof(float);
        const size_t ctr = mtiles * (size_t)(p.Cout / c.BN) * sizeof(unsigned int);
        if (workspace == nullptr || workspace_bytes < ws + ctr) {
            c.splitk = 1;   // caller gave no scratch: run un-split (still correct, fewer CTAs)
        } else {
            q.ws = reinterpret_cast<float*>(workspace);
            q.counters = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(workspace) + ws);
        }
    }
    q.Wt = c.Wt; q.Ht = c.Ht; q.Nt = c.Nt;
    q.tiles_w = c.tiles_w; q.tiles_h = c.tiles_h; q.tiles_n = c.tiles_n;
    q.dbg = g_dbg;
    q.KC = c.KC; q.BN = c.BN; q.kb_per_tap = c.kb_per_tap; q.splitk = c.splitk; q.stages = c.stages; q.tmem_cols = c.tmem_cols;

    const CUtensorMapSwizzle sw = (c.KC == 64) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    CUtensorMap mA[4], mB;
    memset(mA, 0, sizeof(mA));
    const int s = p.s_in;
    const int nviews = (s == 2) ? 4 : 1;
    for (int v = 0; v < nviews; ++v) {
        const int py = v >> 1, px = v & 1;
        const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(p.x) + ((size_t)py * p.W + px) * p.Cin;
        cuuint64_t dims[4] = {(cuuint64_t)p.Cin, (cuuint64_t)(p.W / s), (cuuint64_t)(p.H / s), (cuuint64_t)p.N};
        cuuint64_t strides[3] = {(cuuint64_t)p.Cin * 2 * s, (cuuint64_t)p.W * p.Cin * 2 * s, (cuuint64_t)p.H * p.W * p.Cin * 2};
        cuuint32_t box[4] = {(cuuint32_t)c.KC, (cuuint32_t)c.Wt, (cuuint32_t)c.Ht, (cuuint32_t)c.Nt};
        int rc = encode_map(&mA[v], base, 4, dims, strides, box, sw);
        if (rc != MSTK_OK) return rc;
    }
    for (int v = nviews; v < 4; ++v) mA[v] = mA[0];
    {
        int nslices = 0;
        for (int i = 0; i < p.nphases; ++i)
            for (int t = 0; t < p.ph[i].ntaps; ++t)
                if (p.ph[i].taps[t].slice + 1 > nslices) nslices = p.ph[i].taps[t].slice + 1;
        cuuint64_t dims[2] = {(cuuint64_t)p.Cin, (cuuint64_t)nslices * p.Cout};
        cuuint64_t strides[1] = {(cuuint64_t)p.Cin * 2};
        cuuint32_t box[2] = {(cuuint32_t)c.KC, (cuuint32_t)c.BN};
        int rc = encode_map(&mB, p.w, 2, dims, strides, box, sw);
        if (rc != MSTK_OK) return rc;
    }
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(conv_tc_kernel)");
        attr_set = true;
    }
    dim3 grid((unsigned)(c.tiles_w * c.tiles_h * c.tiles_n), (unsigned)(p.Cout / c.BN), (unsigned)(p.nphases * c.splitk));
    conv_tc_kernel<<<grid, TC_THREADS, c.smem_bytes, st>>>(mA[0], mA[1], mA[2], mA[3], mB, q);
    MSTK_CHECK_LAUNCH("conv_tc launch");
    return MSTK_OK;
}

}  // namespace mstk
