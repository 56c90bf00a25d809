// Halo-resident tcgen05 convolution for the throughput-bound 3x3 stride-1 layers (the frozen VGGish stack of the training
// step: 3.2 of 8.9 ms).  conv_tc.cu loads one A box per (tap, channel block): nine reads of every activation tile through
// L2, and a fresh B box per CTA.  Here:
//   * ONE 4-D TMA box per channel block holds a 16 x 16 pixel block WITH its halo (18 x 18 pixels x 64 channels, 128-byte
//     rows, SWIZZLE_128B, hardware zero fill outside the image).  A tap (dy, dx) is not another load but a shift of the
//     operand descriptor's start address by (dy * 18 + dx) rows; the 8-row core-matrix groups of the MMA are spaced by one
//     halo row (SBO = 18 * 128 B), so group g reads the 8 pixels of image row g.  tools/halo_probe.cu established on B200 that
//     a K-major SWIZZLE_128B descriptor may start at any 128-byte row of a TMA-written tile and take any SBO (the swizzle is a
//     function of the absolute shared-memory address; the descriptor's base-offset field stays 0).
//   * the 16 x 16 block is TWO 128-row MMA tiles (columns 0-7 and 8-15 of every image row) that share every B (weight) box:
//     per channel block a CTA moves 41 KB of activations + 9 x 16 KB of weights for 72 MMAs (conv_tc: 9 x 2 x 32 KB).
//   * persistent CTAs (one per SM) loop over (pixel block, 128-channel output block) items; the four accumulators
//     (2 tiles x 2 buffers x 128 columns = all 512 TMEM columns) are double-buffered, so the epilogue of item i (bias,
//     activation, bf16 / fp32 store straight from registers, 256 contiguous bytes per pixel) runs under the MMAs of item i+1.
// Warps: 0 = TMA producers (lane 0: A; lanes 1..S*G: one B sub-slot each -- loads issued by one thread do not overlap,
// tools/tma_probe.cu), 1 = MMA issuer (converged warp, elect.sync), 2..9 = epilogue.
#include <cuda.h>
#include <cudaTypedefs.h>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace mstk {

constexpr int CH_THREADS = 320;
constexpr int CH_BW = 16, CH_BH = 16;                         // pixel block of one item
constexpr int CH_HW = CH_BW + 2, CH_HH = CH_BH + 2;           // with halo
constexpr uint32_t CH_A_BYTES = CH_HW * CH_HH * 128;          // 41 472
constexpr uint32_t CH_A_STAGE = (CH_A_BYTES + 1023u) & ~1023u;
constexpr uint32_t CH_B_BYTES = 128 * 128;                    // one tap, 128 output channels x 64 input channels
constexpr int CH_MAX_SUB = 6;                                 // B sub-slots (stages x taps per stage)
constexpr int CH_BN = 128;

struct HaloParams {
    const float* bias;
    void* y;
    int32_t N, H, W, Cin, Cout;
    int32_t blocks_w, blocks_h;        // pixel blocks per image
    int32_t nblocks;                   // Cout / 128
    int32_t nitems;
    int32_t KB;                        // Cin / 64
    int32_t G, S;                      // taps per B stage, B stages (G * S <= CH_MAX_SUB)
    int32_t act, affine, out_f32;
    float out_scale, out_shift;
    uint32_t inv_nblocks, inv_bw, inv_bwh;
    int32_t dbg;                       // development: 1 = no output stores, 2 = no TMEM reads either
};

struct HaloShared {
    uint64_t a_full[2], a_empty[2];
    uint64_t b_full[CH_MAX_SUB], b_empty[CH_MAX_SUB];
    uint64_t acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ int hdiv(int n, uint32_t inv) { return inv ? (int)__umulhi((uint32_t)n, inv) : n; }

__global__ void __launch_bounds__(CH_THREADS, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_o,
                 const __grid_constant__ HaloParams q) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) HaloShared sh;
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t off_b = 2u * CH_A_STAGE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = q.G, S = q.S;
    const int R = (9 + G - 1) / G;                  // B rounds per channel block

    if (threadIdx.x == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        prefetch_tmap(&map_o);
        for (int s = 0; s < 2; ++s) { mbar_init(&sh.a_full[s], 1); mbar_init(&sh.a_empty[s], 1); }
        for (int s = 0; s < CH_MAX_SUB; ++s) { mbar_init(&sh.b_full[s], (uint32_t)q.G); mbar_init(&sh.b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&sh.acc_full[s], 1); mbar_init(&sh.acc_empty[s], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sh.tmem_base)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = sh.tmem_base;
    const int first = (int)blockIdx.x, step = (int)gridDim.x;

    if (warp == 0) {
        // ===================== producers =====================
        pdl_wait();
        if (lane == 0) {
            int ka = 0;
            for (int it = first; it < q.nitems; it += step) {
                const int blk = hdiv(it, q.inv_nblocks);
                const int n = hdiv(blk, q.inv_bwh);
                const int rem = blk - n * (q.blocks_w * q.blocks_h);
                const int bh = hdiv(rem, q.inv_bw), bw = rem - bh * q.blocks_w;
                const int w0 = bw * CH_BW - 1, h0 = bh * CH_BH - 1;
                for (int kb = 0; kb < q.KB; ++kb, ++ka) {
                    const int s = ka & 1;
                    if (ka >= 2) mbar_wait(&sh.a_empty[s], (uint32_t)(((ka >> 1) - 1) & 1));
                    mbar_expect_tx(&sh.a_full[s], CH_A_BYTES);
                    tma_load_4d(&map_a, &sh.a_full[s], smem_base + (uint32_t)s * CH_A_STAGE, kb * 64, w0, h0, n);
                }
            }
        } else if (lane <= S * G) {
            const int j = lane - 1;
            const int s = j / G, g = j - s * G;
            const uint32_t dst = smem_base + off_b + (uint32_t)j * CH_B_BYTES;
            int rr = 0;                              // global round counter (all items, all channel blocks)
            for (int it = first; it < q.nitems; it += step) {
                const int blk = hdiv(it, q.inv_nblocks);
                const int nb = it - blk * q.nblocks;
                for (int kb = 0; kb < q.KB; ++kb) {
                    for (int r = 0; r < R; ++r, ++rr) {
                        if (rr % S != s) continue;
                        const int use = rr / S;
                        const int cnt = min(G, 9 - r * G);
                        // EVERY lane of the stage observes EVERY release and arrives on the full barrier (count = G), also
                        // in a short round it does not load for: a lane that skipped a release could not tell release #u
                        // from #u+2 by parity, and a round that could complete without one of its lanes could run two
                        // releases ahead of it -- the same aliasing from the other side.
                        if (use > 0) mbar_wait(&sh.b_empty[s], (uint32_t)((use - 1) & 1));
                        if (g == 0) mbar_expect_tx(&sh.b_full[s], CH_B_BYTES * (uint32_t)cnt);
                        else mbar_arrive(&sh.b_full[s]);
                        if (g >= cnt) continue;
                        tma_load_2d(&map_b, &sh.b_full[s], dst, kb * 64, (r * G + g) * q.Cout + nb * CH_BN);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint32_t idesc = keep_reg(make_idesc(CH_BN));
        const uint32_t a_full0 = keep_reg(smem_u32(&sh.a_full[0])), a_empty0 = keep_reg(smem_u32(&sh.a_empty[0]));
        const uint32_t b_full0 = keep_reg(smem_u32(&sh.b_full[0])), b_empty0 = keep_reg(smem_u32(&sh.b_empty[0]));
        const uint32_t acc_full0 = keep_reg(smem_u32(&sh.acc_full[0])), acc_empty0 = keep_reg(smem_u32(&sh.acc_empty[0]));
        // descriptor high words: SBO = one halo row (A) / 1024 (B), version 1, SWIZZLE_128B
        const uint32_t ahi = keep_reg((uint32_t)((((uint64_t)((CH_HW * 128) >> 4)) | (1ull << 14) | (2ull << 29))));
        const uint32_t bhi = keep_reg((uint32_t)((((uint64_t)(1024 >> 4)) | (1ull << 14) | (2ull << 29))));
        const uint32_t a_lo0 = keep_reg((smem_base & 0x3FFFFu) >> 4);
        const uint32_t b_lo0 = keep_reg(((smem_base + off_b) & 0x3FFFFu) >> 4);
        int ka = 0, rr = 0, ia = 0;
        for (int it = first; it < q.nitems; it += step, ++ia) {
            const int buf = ia & 1;
            if (ia >= 2) mbar_wait_addr(acc_empty0 + 8u * (uint32_t)buf, (uint32_t)(((ia >> 1) - 1) & 1));
            tc_fence_after();
            const uint32_t d0 = tmem_base + (uint32_t)buf * 256u;
            uint32_t accf = 0;
            for (int kb = 0; kb < q.KB; ++kb, ++ka) {
                const int sa = ka & 1;
                mbar_wait_addr(a_full0 + 8u * (uint32_t)sa, (uint32_t)((ka >> 1) & 1));
                const uint32_t a_lo = a_lo0 + (uint32_t)sa * (CH_A_STAGE >> 4);
                for (int r = 0; r < R; ++r, ++rr) {
                    const int s = rr % S;
                    mbar_wait_addr(b_full0 + 8u * (uint32_t)s, (uint32_t)((rr / S) & 1));
                    tc_fence_after();
                    const int cnt = min(G, 9 - r * G);
                    if (elect_one()) {
                        for (int g = 0; g < cnt; ++g) {
                            const int tap = r * G + g;
                            const int dy = tap / 3, dx = tap - dy * 3;
                            const uint32_t at = a_lo + (uint32_t)(dy * CH_HW + dx) * 8u;            // 128 B per halo row = 8 units of 16 B
                            const uint32_t bt = b_lo0 + (uint32_t)(s * G + g) * (CH_B_BYTES >> 4);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                umma_bf16(d0, ((uint64_t)ahi << 32) | (at + 2u * k), ((uint64_t)bhi << 32) | (bt + 2u * k), idesc, accf);
                                umma_bf16(d0 + 128u, ((uint64_t)ahi << 32) | (at + 64u + 2u * k), ((uint64_t)bhi << 32) | (bt + 2u * k), idesc, accf);
                                accf = 1u;
                            }
                        }
                        umma_commit_addr(b_empty0 + 8u * (uint32_t)s);
                        if (r == R - 1) {
                            umma_commit_addr(a_empty0 + 8u * (uint32_t)sa);
                            if (kb == q.KB - 1) umma_commit_addr(acc_full0 + 8u * (uint32_t)buf);
                        }
                    }
                    accf = 1u;
                    __syncwarp();
                }
            }
        }
        if (lane == 0) pdl_launch_dependents();
    } else {
        // ===================== epilogue: warps 2..9 =====================
        const int ew = warp - 2;
        const int qd = warp & 3;                    // TMEM lane quarter this warp may access
        const int half = ew >> 2;                   // column half: [0,64) or [64,128)
        const int row = qd * 32 + lane;             // GEMM row of the tile = pixel (row >> 3, row & 7) of the tile
        const int esz = q.out_f32 ? 4 : 2;
        uint8_t* const stg = smem + off_b + (uint32_t)CH_MAX_SUB * CH_B_BYTES + (uint32_t)ew * 4096u;      // this warp's staging tile
        pdl_wait();
        int ia = 0;
        for (int it = first; it < q.nitems; it += step, ++ia) {
            const int buf = ia & 1;
            const int blk = hdiv(it, q.inv_nblocks);
            const int nb = it - blk * q.nblocks;
            const int n = hdiv(blk, q.inv_bwh);
            const int rem = blk - n * (q.blocks_w * q.blocks_h);
            const int bh = hdiv(rem, q.inv_bw), bw = rem - bh * q.blocks_w;
            const int co0 = nb * CH_BN + half * 64;
            const float* bias = q.bias ? q.bias + co0 : nullptr;
            mbar_wait(&sh.acc_full[buf], (uint32_t)((ia >> 1) & 1));
            tc_fence_after();
            // Both tiles' 64 columns are fetched with ONE tcgen05.ld each, both in flight before the first is used: with
            // 16-column loads issued one at a time the epilogue of the 64-channel layer (one channel block per item) took
            // 15 k cycles per item under the concurrent MMAs and bounded the kernel (727 -> 1426 TFLOP/s without it).
            const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(buf * 256 + half * 64);
            float va[64], vb[64];
            if (!(q.dbg & 2)) {
                tmem_ld64_issue(taddr, va);
                tmem_ld64_issue(taddr + 128u, vb);
            }
            if (!(q.dbg & 2)) { tmem_ld64_wait(va); tmem_ld64_wait(vb); }
            // both tiles are in registers: hand the accumulator buffer back before the arithmetic and the stores
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sh.acc_empty[buf]);
#pragma unroll
            for (int tile = 0; tile < 2; ++tile) {
                if (q.dbg & 2) continue;
                float* v = tile ? vb : va;
                const int h = bh * CH_BH + (row >> 3), w = bw * CH_BW + tile * 8 + (row & 7);
                uint8_t* orow = reinterpret_cast<uint8_t*>(q.y) + ((((size_t)n * q.H + h) * q.W + w) * (size_t)q.Cout + (size_t)co0) * esz;
                if (!q.out_f32) {
                    // the previous TMA store of this warp has finished READING the staging tile
                    if (lane == 0) bulk_wait_read();
                    __syncwarp();
                }
#pragma unroll
                for (int c = 0; c < 64; c += 16) {
#pragma unroll
                    for (int j4 = 0; j4 < 16; j4 += 4) {
                        const float4 bq = bias ? __ldg(reinterpret_cast<const float4*>(bias + c + j4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                        const float bb[4] = {bq.x, bq.y, bq.z, bq.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float x = v[c + j4 + e] + bb[e];
                            if (q.act == MSTK_ACT_RELU) x = fmaxf(x, 0.f);
                            else if (q.act != MSTK_ACT_NONE) x = apply_act(x, q.act);
                            if (q.affine) x = x * q.out_scale + q.out_shift;
                            v[c + j4 + e] = x;
                        }
                    }
                    if (q.dbg & 1) { if (v[c + 3] == 123.456f) reinterpret_cast<float*>(orow)[0] = v[c + 5]; continue; }
                    if (q.out_f32) {
                        float4* o4 = reinterpret_cast<float4*>(orow + (size_t)c * 4);
#pragma unroll
                        for (int j = 0; j < 4; ++j) o4[j] = make_float4(v[c + 4 * j], v[c + 4 * j + 1], v[c + 4 * j + 2], v[c + 4 * j + 3]);
                    } else {
                        // bf16: this warp's 32 pixels x 64 channels go through its own 4 KB staging tile (128-byte rows in the
                        // TMA swizzle: conflict-free 16-byte stores) and leave as ONE TMA box store; 32-byte register stores to
                        // 32 different pixels per instruction cost the 64-channel layer 195 us of 581
                        uint4 pk[2];
                        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                        for (int e = 0; e < 8; ++e) hh[e] = __floats2bfloat162_rn(v[c + 2 * e], v[c + 2 * e + 1]);
                        const int ch = c >> 3;                                  // 16-byte chunk index of the row
                        *reinterpret_cast<uint4*>(stg + lane * 128 + (((ch) ^ (lane & 7)) << 4)) = pk[0];
                        *reinterpret_cast<uint4*>(stg + lane * 128 + (((ch + 1) ^ (lane & 7)) << 4)) = pk[1];
                    }
                }
                if (!q.out_f32 && !(q.dbg & 1)) {
                    fence_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_4d(&map_o, smem_u32(stg), co0, bw * CH_BW + tile * 8, bh * CH_BH + qd * 4, n);
                        bulk_commit();
                    }
                }
            }
        }
        if (lane == 0) bulk_wait_read();          // the staging tile must outlive the last store's read
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int g_halo = -1, g_halo_g = 3, g_halo_s = 2, g_halo_dbg = 0;
static bool g_halo_attr = false;

bool conv_halo_supported(const ConvP& p) {
    if (g_halo < 0) {
        g_halo = 1;
        if (const char* s = getenv("MSTK_CONV_HALO")) g_halo = atoi(s);
        if (const char* s = getenv("MSTK_HALO_DBG")) g_halo_dbg = atoi(s);
        if (const char* s = getenv("MSTK_HALO_GS")) sscanf(s, "%d,%d", &g_halo_g, &g_halo_s);
        if (g_halo_g < 1 || g_halo_s < 2 || g_halo_g * g_halo_s > CH_MAX_SUB) { g_halo_g = 3; g_halo_s = 2; }
    }
    if (!g_halo) return false;
    if (p.dtype != MSTK_BF16 || !p.y || p.add || p.add_nc || p.has_ddim || p.stats) return false;
    if (p.nphases != 1 || p.ph[0].ntaps != 9 || p.s_in != 1 || p.s_out != 1) return false;
    for (int t = 0; t < 9; ++t) {
        const Tap& tp = p.ph[0].taps[t];
        if (tp.dy != t / 3 - 1 || tp.dx != t % 3 - 1 || tp.slice != t) return false;
    }
    if (p.Cin % 64 != 0 || p.Cout % CH_BN != 0) return false;
    if (p.Hg % CH_BH != 0 || p.Wg % CH_BW != 0 || p.Hg != p.H || p.Wg != p.W) return false;
    if ((reinterpret_cast<uintptr_t>(p.x) & 15) || (reinterpret_cast<uintptr_t>(p.w) & 15) || (reinterpret_cast<uintptr_t>(p.y) & 15)) return false;
    if (p.bias && (reinterpret_cast<uintptr_t>(p.bias) & 15)) return false;
    const long long items = (long long)p.N * (p.H / CH_BH) * (p.W / CH_BW) * (p.Cout / CH_BN);
    // throughput-bound grids only: several items per SM, otherwise the one-tile kernel's finer grid wins
    return items >= 4ll * tc_num_sms() && items < 0x7fffffffll;
}

int launch_conv_halo(const ConvP& p, cudaStream_t st) {
    if (!tc_init()) return MSTK_EUNSUPPORTED;
    HaloParams q;
    memset(&q, 0, sizeof(q));
    q.bias = p.bias; q.y = p.y;
    q.N = p.N; q.H = p.H; q.W = p.W; q.Cin = p.Cin; q.Cout = p.Cout;
    q.blocks_w = p.W / CH_BW; q.blocks_h = p.H / CH_BH;
    q.nblocks = p.Cout / CH_BN;
    q.nitems = p.N * q.blocks_w * q.blocks_h * q.nblocks;
    q.KB = p.Cin / 64;
    q.G = g_halo_g; q.S = g_halo_s; q.dbg = g_halo_dbg;
    q.act = p.act; q.affine = p.affine; q.out_f32 = p.out_dtype == MSTK_F32 ? 1 : 0;
    q.out_scale = p.out_scale; q.out_shift = p.out_shift;
    auto inv = [](int d) -> uint32_t { return d <= 1 ? 0u : (uint32_t)((0x100000000ull + (uint64_t)d - 1) / (uint64_t)d); };
    q.inv_nblocks = inv(q.nblocks); q.inv_bw = inv(q.blocks_w); q.inv_bwh = inv(q.blocks_w * q.blocks_h);
    MSTK_REQUIRE((long long)q.nitems * (q.blocks_w * q.blocks_h > q.nblocks ? q.blocks_w * q.blocks_h : q.nblocks) < (1ll << 32),
                 "conv(halo): item grid outside the range of the division-free index math");
    CUtensorMap map_a, map_b;
    {
        const uint64_t dims[4] = {(uint64_t)p.Cin, (uint64_t)p.W, (uint64_t)p.H, (uint64_t)p.N};
        const uint64_t strides[3] = {(uint64_t)p.Cin * 2, (uint64_t)p.W * p.Cin * 2, (uint64_t)p.H * p.W * p.Cin * 2};
        const uint32_t box[4] = {64u, (uint32_t)CH_HW, (uint32_t)CH_HH, 1u};
        int rc = tc_encode_map(&map_a, 0, p.x, 4, dims, strides, box, 128);
        if (rc != MSTK_OK) return rc;
    }
    {
        const uint64_t dims[2] = {(uint64_t)p.Cin, (uint64_t)9 * p.Cout};
        const uint64_t strides[1] = {(uint64_t)p.Cin * 2};
        const uint32_t box[2] = {64u, (uint32_t)CH_BN};
        int rc = tc_encode_map(&map_b, 0, p.w, 2, dims, strides, box, 128);
        if (rc != MSTK_OK) return rc;
    }
    CUtensorMap map_o = map_a;
    if (!q.out_f32) {
        const uint64_t dims[4] = {(uint64_t)p.Cout, (uint64_t)p.W, (uint64_t)p.H, (uint64_t)p.N};
        const uint64_t strides[3] = {(uint64_t)p.Cout * 2, (uint64_t)p.W * p.Cout * 2, (uint64_t)p.H * p.W * p.Cout * 2};
        const uint32_t box[4] = {64u, 8u, 4u, 1u};
        int rc = tc_encode_map(&map_o, 0, p.y, 4, dims, strides, box, 128);
        if (rc != MSTK_OK) return rc;
    }
    const size_t smem = 2 * (size_t)CH_A_STAGE + (size_t)CH_MAX_SUB * CH_B_BYTES + 8 * 4096 + 1024;
    if (!g_halo_attr) {
        cudaError_t ea = cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ea != cudaSuccess) return cuda_fail(ea, "cudaFuncSetAttribute(conv_halo_kernel)");
        g_halo_attr = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    const int sms = tc_num_sms();
    cfg.gridDim = dim3((unsigned)(q.nitems < sms ? q.nitems : sms), 1, 1);
    cfg.blockDim = dim3(CH_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_halo_kernel, map_a, map_b, map_o, q);
    if (e != cudaSuccess) return cuda_fail(e, "conv_halo launch");
    return MSTK_OK;
}

}  // namespace mstk
