// tcgen05 / TMEM / TMA weight gradient of the convolution family (bf16 in, fp32 accumulate).
//
//   dW[tap][co][ci] = sum over GEMM pixels of dy[out pixel, co] * x[in pixel shifted by tap, ci]
//
// is a GEMM whose K dimension is the PIXEL index.  Both operands are NHWC, i.e. channel-contiguous:
// they are "MN-major" UMMA operands, so no transposition is needed -- a 4-D TMA box
// {64 channels, Wt, Ht, Nt} (128-byte rows, SWIZZLE_128B) of dy or of the tap-shifted x lands in
// shared memory exactly in the canonical MN-major SW128 layout (8-pixel groups 1024 B apart = SBO,
// 64-channel chunks KP*128 B apart = LBO).  Padding halos and ragged batch tiles are zero-filled by
// TMA and contribute nothing.
//
// One CTA owns a 128-channel slice of the M side, a BNt-channel slice of the N side, a GROUP of up
// to `tpc` taps of one phase (one TMEM accumulator per tap, so the tap-independent operand is loaded
// once per group) and a range of pixel tiles (split-K).  M side = dy (Cout) unless only Cin is a
// multiple of 128 ("swap": the accumulator is then [ci][co]).  The epilogue adds the fp32
// accumulators into the gradient tensor in the parameter's own layout with RED.ADD.F32.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer,
// warps 2..5 = epilogue.
#include <cuda.h>
#include <stdlib.h>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace mstk {

constexpr int WG_THREADS = 192;
constexpr int WG_KP = 64;            // pixels per k-block
constexpr int WG_MAX_STAGES = 8;
constexpr int WG_MAX_GROUPS = 32;

struct WgMaps {
    CUtensorMap dy[4];   // per phase (strided views of dy for conv-transpose)
    CUtensorMap x[4];    // per stride-2 parity view of x
};

struct WgGroup { int16_t phase, tap0, ntaps, pad; };

struct WgParams {
    ConvP p;                       // forward plan (phases / taps / geometry)
    float* dw;
    long long s_m, s_n;            // dw strides (elements) of the M-side / N-side channel
    int32_t swap;                  // 0: M = dy channels (cout), N = x channels (cin); 1: the other way round
    int32_t CM;                    // channels on the M side (rows >= CM of the last tile are zero padding)
    int32_t BNt;                   // N-side channels per CTA (32, 64, 128 or 256)
    int32_t n_cc;                  // channels per N-side TMA chunk (64 -> SW128, 32 -> SW64)
    int32_t n_chunks;              // BNt / n_cc
    int32_t m_tiles, n_tiles, ngroups, tpc;
    int32_t Wt, Ht, Nt, tiles_w, tiles_h, ptiles;
    int32_t ksplit, stages, tmem_cols;
    int32_t tpm;                   // taps served by one MMA (their B tiles are one MN-major operand): 1 when swapped
    uint32_t m_tile_bytes, n_tile_bytes, stage_bytes;
    WgGroup groups[WG_MAX_GROUPS];
};

// MN-major operand tile: rows of `rowbytes` (128 -> SWIZZLE_128B, 64 -> SWIZZLE_64B) per pixel,
// 8-pixel groups `8*rowbytes` apart (SBO), channel chunks `lbo` bytes apart (LBO).
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t saddr, int rowbytes, uint32_t lbo) {
    const uint32_t sbo = (uint32_t)(8 * rowbytes);
    const uint64_t layout = (rowbytes == 128) ? 2ull : 4ull;
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= 1ull << 46;
    d |= layout << 61;
    return d;
}
// kind::f16 instruction descriptor, D = f32, A = B = bf16, BOTH MN-major, M = 128, N = n
__device__ __forceinline__ uint32_t make_idesc_mn(int n) {
    uint32_t d = 0;
    d |= 1u << 4;
    d |= 1u << 7;
    d |= 1u << 10;
    d |= 1u << 15;                      // A MN-major
    d |= 1u << 16;                      // B MN-major
    d |= (uint32_t)(n >> 3) << 17;
    d |= (uint32_t)(TILE_M >> 4) << 24;
    return d;
}

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_tc_kernel(const __grid_constant__ WgMaps maps, const __grid_constant__ WgParams q) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[WG_MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[WG_MAX_STAGES];
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ uint32_t tmem_base_smem;

    const ConvP& p = q.p;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // work decomposition: blockIdx.x = group + ngroups * (m_tile + m_tiles * n_tile)  (taps of the same
    // pixel range are neighbours in launch order, so the shared operand is an L2 hit), blockIdx.y = k-split
    const int g_idx = blockIdx.x % q.ngroups;
    const int rest = blockIdx.x / q.ngroups;
    const int mt = rest % q.m_tiles;
    const int nt = rest / q.m_tiles;
    const WgGroup grp = q.groups[g_idx];
    const Phase& ph = p.ph[grp.phase];
    const int ntaps = grp.ntaps;
    const int m0 = mt * TILE_M;
    const int n0c = nt * q.BNt;
    const int pt_begin = (int)(((long long)blockIdx.y * q.ptiles) / q.ksplit);
    const int pt_end = (int)(((long long)(blockIdx.y + 1) * q.ptiles) / q.ksplit);
    const int nkb = pt_end - pt_begin;

    // pointer arithmetic (not an integer round trip) so that the compiler keeps the shared address space: LDS/STS/ATOMS
    // instead of generic LD/ST/ATOM in the epilogue
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    const int n_rowbytes = q.n_cc * 2;
    const uint32_t m_chunk_bytes = (uint32_t)(WG_KP * 128);
    const uint32_t n_chunk_bytes = (uint32_t)(WG_KP * n_rowbytes);
    // stage layout: [M tiles][N tiles]; the tap-dependent side (x) holds `tpc` tiles, the other one tile
    const uint32_t off_n = (q.swap ? (uint32_t)q.tpc : 1u) * q.m_tile_bytes;

    if (threadIdx.x == 0) {
        prefetch_tmap(&maps.dy[grp.phase]);
        prefetch_tmap(&maps.x[0]);
        for (int s = 0; s < q.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&accum_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)),
                     "r"((uint32_t)q.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    // ===================== TMA producers =====================
    // One stage takes n_dy + ntaps * n_x box loads.  Issuing a TMA load occupies its warp for ~200 cycles (measured,
    // tools/tma_probe.cu), so each load of the stage is a job owned by one lane, jobs dealt round-robin over the five
    // non-MMA warps (the epilogue warps have nothing else to do until the accumulators are complete); every job lane
    // walks all k-blocks, waiting on the stage's empty barrier itself; job 0 posts the stage's byte count.
    if (warp != 1 && nkb > 0) {
        const int n_dy = q.swap ? q.n_chunks : 2;
        const int n_x = q.swap ? 2 : q.n_chunks;
        const int njobs = n_dy + ntaps * n_x;
        const int slot = warp == 0 ? 0 : warp - 1;                 // warps 0,2,3,4,5 -> 0..4
        const int j = lane * 5 + slot;
        if (j < njobs) {
            const uint32_t tx_bytes = (q.swap ? (uint32_t)ntaps : 1u) * q.m_tile_bytes + (q.swap ? 1u : (uint32_t)ntaps) * q.n_tile_bytes;
            const int dy_c0 = q.swap ? n0c : m0;        // first dy channel of this CTA
            const int x_c0 = q.swap ? m0 : n0c;         // first x channel
            const CUtensorMap* map;
            uint32_t dst;
            int c0, dx = 0, dy = 0;
            if (j < n_dy) {
                map = &maps.dy[grp.phase];
                dst = q.swap ? off_n + (uint32_t)j * n_chunk_bytes : (uint32_t)j * m_chunk_bytes;
                c0 = q.swap ? dy_c0 + q.n_cc * j : dy_c0 + 64 * j;
            } else {
                const int tj = (j - n_dy) / n_x, cj = (j - n_dy) - tj * n_x;
                const Tap tap = ph.taps[grp.tap0 + tj];
                dy = tap.dy; dx = tap.dx;
                map = &maps.x[0];
                if (p.s_in == 2) {
                    const int py = dy & 1, px = dx & 1;
                    dy = (dy - py) / 2;
                    dx = (dx - px) / 2;
                    map = &maps.x[py * 2 + px];
                }
                dst = q.swap ? (uint32_t)tj * q.m_tile_bytes + (uint32_t)cj * m_chunk_bytes
                             : off_n + (uint32_t)tj * q.n_tile_bytes + (uint32_t)cj * n_chunk_bytes;
                c0 = q.swap ? x_c0 + 64 * cj : x_c0 + q.n_cc * cj;
            }
            int s = 0;
            uint32_t par = 1;                            // parity of the empty barrier's PREVIOUS phase use
            for (int it = 0; it < nkb; ++it) {
                if (it >= q.stages) mbar_wait(&empty_bar[s], par);
                const int pt = pt_begin + it;
                const int tw = pt % q.tiles_w;
                const int th = (pt / q.tiles_w) % q.tiles_h;
                const int tn = pt / (q.tiles_w * q.tiles_h);
                if (j == 0) mbar_expect_tx(&full_bar[s], tx_bytes);
                tma_load_4d(map, &full_bar[s], smem_base + (uint32_t)s * q.stage_bytes + dst, c0, tw * q.Wt + dx, th * q.Ht + dy, tn * q.Nt);
                if (++s == q.stages) { s = 0; par ^= 1u; }
            }
        }
    }
    if (warp == 1) {
        // ===================== MMA issuer =====================
        // whole warp converged, one elected lane issues (see conv_tc_body.cuh); invariants pinned in registers
        if (nkb > 0) {
            const uint32_t idesc = keep_reg(make_idesc_mn(q.BNt));
            const int tpm = keep_reg(q.tpm);                                   // taps per MMA (1 when swapped)
            const uint32_t idesc_m[3] = {make_idesc_mn(2 * q.BNt <= 256 ? 2 * q.BNt : q.BNt), make_idesc_mn(3 * q.BNt <= 256 ? 3 * q.BNt : q.BNt),
                                         make_idesc_mn(4 * q.BNt <= 256 ? 4 * q.BNt : q.BNt)};
            const uint32_t kstep_a = (uint32_t)(16 * 128) >> 4;            // 16 pixels further along K
            const uint32_t kstep_b = keep_reg((uint32_t)(16 * n_rowbytes) >> 4);
            const int stages = keep_reg(q.stages), nk = keep_reg(nkb), nt_ = keep_reg(ntaps);
            const uint32_t full0 = keep_reg(smem_u32(&full_bar[0])), empty0 = keep_reg(smem_u32(&empty_bar[0]));
            const uint32_t accum_a = keep_reg(smem_u32(&accum_bar));
            const uint32_t a_tap = keep_reg(q.swap ? q.m_tile_bytes >> 4 : 0u), b_tap = keep_reg(q.swap ? 0u : q.n_tile_bytes >> 4);
            const uint64_t adesc0 = make_desc_mn(smem_base, 128, m_chunk_bytes);
            const uint64_t bdesc0 = make_desc_mn(smem_base + off_n, n_rowbytes, n_chunk_bytes);
            const uint32_t alo0 = keep_reg((uint32_t)adesc0), blo0 = keep_reg((uint32_t)bdesc0);
            const uint32_t ahi = keep_reg((uint32_t)(adesc0 >> 32)), bhi = keep_reg((uint32_t)(bdesc0 >> 32));
            const uint32_t dstage = keep_reg(q.stage_bytes >> 4);
            const uint32_t bnt = keep_reg((uint32_t)q.BNt), tm = keep_reg(tmem_base);
            int s = 0;
            uint32_t par = 0, soff = 0, acc = 0;
            for (int it = 0; it < nk; ++it) {
                mbar_wait_addr(full0 + 8u * (uint32_t)s, par);
                tc_fence_after();
                if (elect_one()) {
                    uint32_t alo = alo0 + soff, blo = blo0 + soff, tacc = tm;
                    // Not swapped: the taps' x tiles follow each other in the stage exactly one channel-chunk stride apart, so
                    // `tpm` taps are ONE MN-major B operand of N = tpm * BNt (<= 256) and their accumulators adjacent TMEM
                    // columns: one MMA instead of tpm (an MMA costs ~110 cycles to issue whatever its N; with N = 64 per tap the
                    // K loop was bound by that, tools/wgrad_tune.py)
                    for (int t = 0; t < nt_; t += tpm) {
                        const int nn = nt_ - t < tpm ? nt_ - t : tpm;
                        const uint32_t ide = nn == 1 ? idesc : idesc_m[nn - 2];
#pragma unroll
                        for (int k = 0; k < WG_KP / 16; ++k)
                            umma_bf16(tacc, ((uint64_t)ahi << 32) | (alo + k * kstep_a), ((uint64_t)bhi << 32) | (blo + k * kstep_b), ide,
                                      (acc | (uint32_t)k) ? 1u : 0u);
                        alo += a_tap * (uint32_t)nn; blo += b_tap * (uint32_t)nn; tacc += bnt * (uint32_t)nn;
                    }
                    umma_commit_addr(empty0 + 8u * (uint32_t)s);
                    if (it == nk - 1) umma_commit_addr(accum_a);
                }
                acc = 1u;
                __syncwarp();
                if (++s == stages) { s = 0; par ^= 1u; soff = 0; }
                else soff += dstage;
            }
        }
    }
    if (warp >= 2 && nkb > 0) {
        // ===================== epilogue (warps 2..5) =====================
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        mbar_wait(&accum_bar, 0);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16);
        float* drow = q.dw + (long long)(m0 + row) * q.s_m + (long long)n0c * q.s_n;
        for (int t = 0; t < ntaps && m0 + row < q.CM; ++t) {
            float* dt = drow + ph.taps[grp.tap0 + t].slice;
            for (int c = 0; c < q.BNt; c += 16) {
                float v[16];
                tmem_ld16(taddr + (uint32_t)(t * q.BNt + c), v);
#pragma unroll
                for (int j = 0; j < 16; ++j) atomicAdd(dt + (long long)(c + j) * q.s_n, v[j]);
            }
        }
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)q.tmem_cols) : "memory");
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int g_wg_state = 0;     // 0 unknown, 1 ok, -1 failed / disabled

static bool wgrad_tc_init() {
    if (g_wg_state != 0) return g_wg_state > 0;
    if (const char* s = getenv("MSTK_WGRAD_TC")) {
        if (atoi(s) == 0) { g_wg_state = -1; return false; }
    }
    if (!tc_init()) { g_wg_state = -1; return false; }
    cudaError_t e = cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    if (e != cudaSuccess) { (void)cudaGetLastError(); g_wg_state = -1; return false; }
    g_wg_state = 1;
    return true;
}

static int wg_pow2_divisor(int x, int cap) {
    int d = 1;
    while (d * 2 <= cap && (x % (d * 2)) == 0) d *= 2;
    return d;
}

// M side (128 TMEM lanes): a channel count that is a multiple of 128, else a multiple of 64 (the upper 64 rows
// are then an out-of-range TMA chunk = zeros).  0: M = cout, 1: M = cin ("swap"), -1: unsupported.
static int wg_pick_swap(const ConvP& p) {
    if ((p.Cout % 128) == 0) return 0;
    if ((p.Cin % 128) == 0) return 1;
    if ((p.Cout % 64) == 0) return 0;
    if ((p.Cin % 64) == 0) return 1;
    return -1;
}

// `p` is the FORWARD plan with p.x = forward input and p.y = dy.
bool wgrad_tc_supported(const ConvP& p, int x_dtype, int dy_dtype) {
    if (x_dtype != MSTK_BF16 || dy_dtype != MSTK_BF16) return false;
    const int sw = wg_pick_swap(p);
    if (sw < 0) return false;
    const int cn = sw ? p.Cout : p.Cin;
    if (!(cn % 64 == 0 || cn == 32)) return false;
    if (p.s_in == 2 && ((p.H & 1) || (p.W & 1))) return false;
    if ((reinterpret_cast<uintptr_t>(p.x) & 15) || (reinterpret_cast<uintptr_t>(p.y) & 15)) return false;
    int groups = 0;
    for (int i = 0; i < p.nphases; ++i) groups += p.ph[i].ntaps;     // worst case tpc = 1
    if (groups > WG_MAX_GROUPS) return false;
    return wgrad_tc_init();
}

int launch_wgrad_tc(const ConvP& p, float* dw, long long s_co, long long s_ci, cudaStream_t st) {
    if (!wgrad_tc_init()) return MSTK_EUNSUPPORTED;
    WgParams q;
    memset(&q, 0, sizeof(q));
    q.p = p;
    q.dw = dw;
    q.swap = wg_pick_swap(p);
    if (q.swap < 0) return MSTK_EUNSUPPORTED;
    const int CM = q.swap ? p.Cin : p.Cout;
    const int CN = q.swap ? p.Cout : p.Cin;
    q.s_m = q.swap ? s_ci : s_co;
    q.s_n = q.swap ? s_co : s_ci;
    if (CN == 32) { q.BNt = 32; q.n_cc = 32; }
    else { q.n_cc = 64; q.BNt = (CN % 256 == 0) ? 256 : (CN % 128 == 0) ? 128 : 64; }
    q.n_chunks = q.BNt / q.n_cc;
    q.m_tiles = (CM + TILE_M - 1) / TILE_M;
    q.CM = CM;
    q.n_tiles = CN / q.BNt;
    q.m_tile_bytes = (uint32_t)(TILE_M * WG_KP * 2);
    q.n_tile_bytes = (uint32_t)(q.BNt * WG_KP * 2);
    // taps per CTA: bounded by TMEM columns (512) and a 64 KB stage
    const uint32_t budget = 64 * 1024;
    int tmax = 512 / q.BNt;
    if (tmax > 4) tmax = 4;
    {
        const uint32_t dep = q.swap ? q.m_tile_bytes : q.n_tile_bytes;
        const uint32_t fix = q.swap ? q.n_tile_bytes : q.m_tile_bytes;
        const int fit = (int)((budget - fix) / dep);
        if (tmax > fit) tmax = fit;
        if (tmax < 1) tmax = 1;
    }
    q.ngroups = 0;
    q.tpc = 1;
    for (int i = 0; i < p.nphases; ++i) {
        const int nt = p.ph[i].ntaps;
        const int ng = (nt + tmax - 1) / tmax;
        const int t = (nt + ng - 1) / ng;                 // balanced group size
        if (t > q.tpc) q.tpc = t;
        for (int t0 = 0; t0 < nt; t0 += t) {
            if (q.ngroups >= WG_MAX_GROUPS) { set_error("wgrad(tcgen05): too many tap groups"); return MSTK_EUNSUPPORTED; }
            WgGroup& g = q.groups[q.ngroups++];
            g.phase = (int16_t)i; g.tap0 = (int16_t)t0; g.ntaps = (int16_t)((nt - t0) < t ? (nt - t0) : t); g.pad = 0;
        }
    }
    q.stage_bytes = (q.swap ? (uint32_t)q.tpc : 1u) * q.m_tile_bytes + (q.swap ? 1u : (uint32_t)q.tpc) * q.n_tile_bytes;
    q.stages = (int)((200u * 1024u) / q.stage_bytes);
    if (q.stages > WG_MAX_STAGES) q.stages = WG_MAX_STAGES;
    if (q.stages < 2) q.stages = 2;
    int cols = 32;
    while (cols < q.tpc * q.BNt) cols <<= 1;
    q.tmem_cols = cols;
    // pixel tiles
    q.Wt = wg_pow2_divisor(p.Wg, WG_KP);
    q.Ht = wg_pow2_divisor(p.Hg, WG_KP / q.Wt);
    q.Nt = WG_KP / (q.Wt * q.Ht);
    q.tiles_w = p.Wg / q.Wt;
    q.tiles_h = p.Hg / q.Ht;
    const int tiles_n = (p.N + q.Nt - 1) / q.Nt;
    q.ptiles = q.tiles_w * q.tiles_h * tiles_n;
    const long long base = (long long)q.ngroups * q.m_tiles * q.n_tiles;
    // one CTA per SM (the stages take ~200 KB): aim at two FULL waves, never a third one holding a handful of CTAs
    long long ks = (2LL * tc_num_sms()) / base;
    const long long ks_max = (q.ptiles + 3) / 4;          // at least ~4 k-blocks per CTA
    if (ks > ks_max) ks = ks_max;
    if (ks < 1) ks = 1;
    if (ks > 65535) ks = 65535;
    {
        static int ks_cap = -1;                               // development: MSTK_WG_KSMAX caps the pixel split
        if (ks_cap < 0) { const char* e = getenv("MSTK_WG_KSMAX"); ks_cap = e ? atoi(e) : 0; }
        if (ks_cap > 0 && ks > ks_cap) ks = ks_cap;
    }
    q.ksplit = (int)ks;
    {
        static int merge = -1;                                // MSTK_WG_MERGE=0: one MMA per tap (development)
        if (merge < 0) { const char* e = getenv("MSTK_WG_MERGE"); merge = e ? atoi(e) : 1; }
        q.tpm = (!q.swap && merge) ? 256 / q.BNt : 1;
        if (q.tpm < 1) q.tpm = 1;
        if (q.tpm > 4) q.tpm = 4;
    }

    WgMaps maps;
    memset(&maps, 0, sizeof(maps));
    const int dy_cc = q.swap ? q.n_cc : 64;
    const int x_cc = q.swap ? 64 : q.n_cc;
    for (int i = 0; i < p.nphases; ++i) {
        const Phase& ph = p.ph[i];
        const __nv_bfloat16* b = reinterpret_cast<const __nv_bfloat16*>(p.y) + ((size_t)ph.oy * p.Wo + ph.ox) * p.Cout;
        const uint64_t s = (uint64_t)p.s_out;
        uint64_t dims[4] = {(uint64_t)p.Cout, (uint64_t)p.Wg, (uint64_t)p.Hg, (uint64_t)p.N};
        uint64_t strides[3] = {s * p.Cout * 2, s * p.Wo * p.Cout * 2, (uint64_t)p.Ho * p.Wo * p.Cout * 2};
        uint32_t box[4] = {(uint32_t)dy_cc, (uint32_t)q.Wt, (uint32_t)q.Ht, (uint32_t)q.Nt};
        int rc = tc_encode_map(&maps.dy[i], 0, b, 4, dims, strides, box, dy_cc * 2);
        if (rc != MSTK_OK) return rc;
    }
    for (int i = p.nphases; i < 4; ++i) maps.dy[i] = maps.dy[0];
    const int s = p.s_in;
    const int nviews = (s == 2) ? 4 : 1;
    for (int v = 0; v < nviews; ++v) {
        const int py = v >> 1, px = v & 1;
        const __nv_bfloat16* b = reinterpret_cast<const __nv_bfloat16*>(p.x) + ((size_t)py * p.W + px) * p.Cin;
        uint64_t dims[4] = {(uint64_t)p.Cin, (uint64_t)(p.W / s), (uint64_t)(p.H / s), (uint64_t)p.N};
        uint64_t strides[3] = {(uint64_t)p.Cin * 2 * s, (uint64_t)p.W * p.Cin * 2 * s, (uint64_t)p.H * p.W * p.Cin * 2};
        uint32_t box[4] = {(uint32_t)x_cc, (uint32_t)q.Wt, (uint32_t)q.Ht, (uint32_t)q.Nt};
        int rc = tc_encode_map(&maps.x[v], 0, b, 4, dims, strides, box, x_cc * 2);
        if (rc != MSTK_OK) return rc;
    }
    for (int v = nviews; v < 4; ++v) maps.x[v] = maps.x[0];

    dim3 grid((unsigned)base, (unsigned)q.ksplit, 1);
    const size_t smem = (size_t)q.stages * q.stage_bytes + 1024;
    wgrad_tc_kernel<<<grid, WG_THREADS, smem, st>>>(maps, q);
    MSTK_CHECK_LAUNCH("wgrad_tc launch");
    return MSTK_OK;
}

}  // namespace mstk
