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
#include "conv_tc_body.cuh"      // colsum16 / colsum_index (BatchNorm column sums)

namespace mstk {

constexpr int CH_THREADS = 320;
constexpr int CH_BH = 16;                                     // pixel rows of one item (its width: 8 per MMA tile)
constexpr int CH_HH = CH_BH + 2;                              // with halo
constexpr uint32_t CH_B_BYTES = 128 * 128;                    // one tap, 128 output channels x 64 input channels
constexpr int CH_MAX_SUB = 6;                                 // B sub-slots (stages x taps per stage)
constexpr int CH_BN = 128;

struct HaloParams {
    const float* bias;
    void* y;
    double* stats;                     // optional [2 * Cout] BatchNorm sums of (conv + bias), conv-transpose kernel only
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
    // conv-transpose k4 s2 (CONVT kernel): 16 (phase, tap) operations per channel block, all on the SAME 18 x 18 halo box
    // Four rounds of four 8 KB weight boxes per channel block (op_slice[round * 4 + slot]); parities that read the SAME input offset
    // sit in adjacent boxes and adjacent accumulator columns ([P00 | P01 | P11 | P10]), so one MMA serves them together:
    // 10 operations per channel block (1 x N = 256, 3 x N = 128, 6 x N = 64) instead of 16 of N = 64 -- the kernel was bound by
    // the MMA issue rate (~110 cycles per MMA whatever its N).
    int16_t op_slice[16];              // weight slice ky * 4 + kx of box (round, slot)
    int16_t t_aoff[10];                // halo row of the operation's A start: (dy + 1) * (halo width) + (dx + 1)
    int8_t t_slot[10], t_nsel[10], t_dcol[10];   // first box of the round, N = 64 << nsel, first accumulator column / 64
    int8_t t_first[5];                 // operations of round r: [t_first[r], t_first[r + 1])
    int32_t nboxes;                    // 16 (k4) or 9 (k3)
    // stride-2 3x3 convolution (S2 kernel): a stride-1 2x2 convolution over the four parity planes of the input, one plane per
    // "channel block": plane (py, px) contributes (1 + py) * (1 + px) taps, all on the plane's own 18 x 18 halo box
    int32_t KBC;                       // channel blocks per plane (Cin / 64)
    int8_t s2_ntaps[4];                // taps of plane py * 2 + px
    int8_t s2_slice[4][4];             // weight slice ky * 3 + kx of the plane's tap
    int16_t s2_aoff[4][4];             // halo row of its A start
    uint32_t off_stg;                  // byte offset of the epilogue warps' staging tiles
};

struct HaloOutMaps { CUtensorMap m[4]; };
struct HaloInMaps { CUtensorMap m[4]; };    // input tensor map(s): [0], or one per parity plane for the stride-2 convolution   // output tensor map(s): [0] for the convolution, one per output parity for conv-transpose

struct HaloShared {
    uint64_t a_full[2], a_empty[2];
    uint64_t b_full[CH_MAX_SUB], b_empty[CH_MAX_SUB];
    uint64_t acc_full[2], acc_empty[2];
    uint32_t tmem_base;
    float sstat[2 * 64];               // conv-transpose + BatchNorm statistics: per-CTA column sums (sum, sum of squares)
};

__device__ __forceinline__ int hdiv(int n, uint32_t inv) { return inv ? (int)__umulhi((uint32_t)n, inv) : n; }

// CONVT = false: 3x3 stride-1 convolution, 9 taps, BN = 128 output channels per item, two accumulator buffers of 2 x 128 columns.
// CONVT = true : ConvTranspose2d(k4, s2, p1) with Cout = 64, or (k3, s2, p1, output_padding 1) with Cout % 64 == 0 (the data
//                gradient of the stride-2 convolutions): ONE 16 x 8 tile per item, its four output parities four accumulators
//                of 64 columns ([P00 | P01 | P11 | P10], double-buffered); parities that read the same input offset share one
//                MMA (N = 256 / 128 / 64, host-built operation table); each epilogue warp stores (parity) boxes through the
//                parity's strided output map.
template <bool CONVT, int TILES, bool STATS, bool S2>
__global__ void __launch_bounds__(CH_THREADS, 1)
conv_halo_kernel(const __grid_constant__ HaloInMaps maps_a, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ HaloOutMaps maps_o,
                 const __grid_constant__ HaloParams q) {
    const CUtensorMap& map_a = maps_a.m[0];
    const int NOPS = CONVT ? q.nboxes : 9;                  // weight boxes per channel block
    constexpr int BN = CONVT ? 64 : 128;                    // MMA N
    constexpr uint32_t B_BYTES = (uint32_t)BN * 128u;       // one operation's weight box
    constexpr uint32_t TILE_COLS = CONVT ? 256u : 128u;     // TMEM columns of one tile's accumulator(s)
    constexpr int NBUF = 512 / (TILES * (int)TILE_COLS);   // accumulator buffers: all 512 TMEM columns are used
    constexpr int BW = 8 * TILES;                           // pixel columns of one item
    constexpr int HW = BW + 2;                              // with halo
    constexpr uint32_t A_BYTES = (uint32_t)(HW * CH_HH * 128);
    constexpr uint32_t A_STAGE = (A_BYTES + 1023u) & ~1023u;
    static_assert(NBUF >= 1 && NBUF <= 2, "accumulator budget");
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) HaloShared sh;
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t off_b = 2u * A_STAGE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = q.G, S = q.S;

    if (threadIdx.x == 0) {
        prefetch_tmap(&map_a);
        prefetch_tmap(&map_b);
        prefetch_tmap(&maps_o.m[0]);
        for (int s = 0; s < 2; ++s) { mbar_init(&sh.a_full[s], 1); mbar_init(&sh.a_empty[s], 1); }
        for (int s = 0; s < CH_MAX_SUB; ++s) { mbar_init(&sh.b_full[s], (uint32_t)q.G); mbar_init(&sh.b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&sh.acc_full[s], 1); mbar_init(&sh.acc_empty[s], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x >= 64 && threadIdx.x < 64 + 128) sh.sstat[threadIdx.x - 64] = 0.f;
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
                const int w0 = bw * BW - 1, h0 = bh * CH_BH - 1;
                for (int kb = 0; kb < q.KB; ++kb, ++ka) {
                    const int s = ka & 1;
                    if (ka >= 2) mbar_wait(&sh.a_empty[s], (uint32_t)(((ka >> 1) - 1) & 1));
                    mbar_expect_tx(&sh.a_full[s], A_BYTES);
                    const int plane = S2 ? kb / q.KBC : 0, cb = S2 ? kb - plane * q.KBC : kb;
                    tma_load_4d(&maps_a.m[plane], &sh.a_full[s], smem_base + (uint32_t)s * A_STAGE, cb * 64, w0, h0, n);
                }
            }
        } else if (lane <= S * G) {
            const int j = lane - 1;
            const int s = j / G, g = j - s * G;
            const uint32_t dst = smem_base + off_b + (uint32_t)j * B_BYTES;
            int rr = 0;                              // global round counter (all items, all channel blocks)
            for (int it = first; it < q.nitems; it += step) {
                const int blk = hdiv(it, q.inv_nblocks);
                const int nb = it - blk * q.nblocks;
                for (int kb = 0; kb < q.KB; ++kb) {
                    const int plane = S2 ? kb / q.KBC : 0, cb = S2 ? kb - plane * q.KBC : kb;
                    const int nops = S2 ? (int)q.s2_ntaps[plane] : NOPS, R = (nops + G - 1) / G;      // B rounds of this channel block
                    for (int r = 0; r < R; ++r, ++rr) {
                        if (rr % S != s) continue;
                        const int use = rr / S;
                        const int cnt = min(G, nops - r * G);
                        // EVERY lane of the stage observes EVERY release and arrives on the full barrier (count = G), also
                        // in a short round it does not load for: a lane that skipped a release could not tell release #u
                        // from #u+2 by parity, and a round that could complete without one of its lanes could run two
                        // releases ahead of it -- the same aliasing from the other side.
                        if (use > 0) mbar_wait(&sh.b_empty[s], (uint32_t)((use - 1) & 1));
                        if (g == 0) mbar_expect_tx(&sh.b_full[s], B_BYTES * (uint32_t)cnt);
                        else mbar_arrive(&sh.b_full[s]);
                        if (g >= cnt) continue;
                        const int slice = CONVT ? (int)q.op_slice[r * G + g] : S2 ? (int)q.s2_slice[plane][r * G + g] : r * G + g;
                        tma_load_2d(&map_b, &sh.b_full[s], dst, cb * 64, slice * q.Cout + nb * BN);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint32_t idesc = keep_reg(make_idesc(BN));
        const uint32_t idesc_n[3] = {make_idesc(64), make_idesc(128), make_idesc(256)};
        const uint32_t a_full0 = keep_reg(smem_u32(&sh.a_full[0])), a_empty0 = keep_reg(smem_u32(&sh.a_empty[0]));
        const uint32_t b_full0 = keep_reg(smem_u32(&sh.b_full[0])), b_empty0 = keep_reg(smem_u32(&sh.b_empty[0]));
        const uint32_t acc_full0 = keep_reg(smem_u32(&sh.acc_full[0])), acc_empty0 = keep_reg(smem_u32(&sh.acc_empty[0]));
        // descriptor high words: SBO = one halo row (A) / 1024 (B), version 1, SWIZZLE_128B
        const uint32_t ahi = keep_reg((uint32_t)((((uint64_t)((HW * 128) >> 4)) | (1ull << 14) | (2ull << 29))));
        const uint32_t bhi = keep_reg((uint32_t)((((uint64_t)(1024 >> 4)) | (1ull << 14) | (2ull << 29))));
        const uint32_t a_lo0 = keep_reg((smem_base & 0x3FFFFu) >> 4);
        const uint32_t b_lo0 = keep_reg(((smem_base + off_b) & 0x3FFFFu) >> 4);
        int ka = 0, rr = 0, ia = 0;
        for (int it = first; it < q.nitems; it += step, ++ia) {
            const int buf = ia % NBUF, use = ia / NBUF;
            if (use >= 1) mbar_wait_addr(acc_empty0 + 8u * (uint32_t)buf, (uint32_t)((use - 1) & 1));
            tc_fence_after();
            const uint32_t d0 = tmem_base + (uint32_t)buf * (uint32_t)(TILES * TILE_COLS);
            uint32_t accf = 0;
            for (int kb = 0; kb < q.KB; ++kb, ++ka) {
                const int sa = ka & 1;
                mbar_wait_addr(a_full0 + 8u * (uint32_t)sa, (uint32_t)((ka >> 1) & 1));
                const uint32_t a_lo = a_lo0 + (uint32_t)sa * (A_STAGE >> 4);
                const int plane = S2 ? kb / q.KBC : 0;
                const int nops = S2 ? (int)q.s2_ntaps[plane] : NOPS, R = (nops + G - 1) / G;
                for (int r = 0; r < R; ++r, ++rr) {
                    const int s = rr % S;
                    mbar_wait_addr(b_full0 + 8u * (uint32_t)s, (uint32_t)((rr / S) & 1));
                    tc_fence_after();
                    const int cnt = min(G, nops - r * G);
                    if (elect_one()) {
                        if (CONVT) {
                            // round 0 is the one N = 256 operation on the centre offset: it opens all four parity accumulators
                            for (int o = q.t_first[r]; o < q.t_first[r + 1]; ++o) {
                                const uint32_t at = a_lo + (uint32_t)q.t_aoff[o] * 8u;
                                const uint32_t bt = b_lo0 + (uint32_t)(s * G + q.t_slot[o]) * (B_BYTES >> 4);
                                const uint32_t dcol = (uint32_t)q.t_dcol[o] * 64u, ide = idesc_n[q.t_nsel[o]];
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    umma_bf16(d0 + dcol, ((uint64_t)ahi << 32) | (at + 2u * k), ((uint64_t)bhi << 32) | (bt + 2u * k), ide, accf);
                                    accf = 1u;
                                }
                            }
                        }
                        for (int g = 0; g < (CONVT ? 0 : cnt); ++g) {
                            const int op = r * G + g;
                            const int dy = op / 3, dx = op - dy * 3;
                            const uint32_t aoff = S2 ? (uint32_t)q.s2_aoff[plane][op] : (uint32_t)(dy * HW + dx), dcol = 0u;
                            const uint32_t at = a_lo + aoff * 8u;                                   // 128 B per halo row = 8 units of 16 B
                            const uint32_t bt = b_lo0 + (uint32_t)(s * G + g) * (B_BYTES >> 4);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                umma_bf16(d0 + dcol, ((uint64_t)ahi << 32) | (at + 2u * k), ((uint64_t)bhi << 32) | (bt + 2u * k), idesc, accf);
                                if (TILES == 2)
                                    umma_bf16(d0 + TILE_COLS + dcol, ((uint64_t)ahi << 32) | (at + 64u + 2u * k), ((uint64_t)bhi << 32) | (bt + 2u * k), idesc, accf);
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
        uint8_t* const stg = smem + q.off_stg + (uint32_t)ew * 4096u;      // this warp's staging tile
        pdl_wait();
        int ia = 0;
        for (int it = first; it < q.nitems; it += step, ++ia) {
            const int buf = ia % NBUF, use = ia / NBUF;
            const int blk = hdiv(it, q.inv_nblocks);
            const int nb = it - blk * q.nblocks;
            const int n = hdiv(blk, q.inv_bwh);
            const int rem = blk - n * (q.blocks_w * q.blocks_h);
            const int bh = hdiv(rem, q.inv_bw), bw = rem - bh * q.blocks_w;
            const int co0 = CONVT ? nb * 64 : nb * BN + half * 64;
            const float* bias = q.bias ? q.bias + co0 : nullptr;
            mbar_wait(&sh.acc_full[buf], (uint32_t)(use & 1));
            tc_fence_after();
            // 64 accumulator columns per tcgen05.ld, two loads in flight before the first is used: with 16-column loads issued
            // one at a time the epilogue of the 64-channel layer (one channel block per item) took 15 k cycles per item under
            // the concurrent MMAs and bounded the kernel (727 -> 1426 TFLOP/s without it).
            // Convolution: warp = (lane quarter, column half); the pair is (tile 0, tile 1).
            // Conv-transpose: warp = (lane quarter, output row parity py); pair p = tile p, its two members the column parities.
            constexpr int NPAIR = CONVT ? TILES : 1;
#pragma unroll 1
            for (int pair = 0; pair < NPAIR; ++pair) {
                const uint32_t col0 = (uint32_t)buf * (uint32_t)(TILES * TILE_COLS);
                // conv-transpose accumulator columns: [P00 | P01 | P11 | P10] (row parity = half, column parity = pair member)
                const uint32_t col_a = col0 + (CONVT ? (uint32_t)(pair * 256 + (half ? 192 : 0)) : (uint32_t)(half * 64));
                const uint32_t col_b = CONVT ? col0 + (uint32_t)(pair * 256 + (half ? 128 : 64)) : col_a + 128u;
                const uint32_t tlane = tmem_base + ((uint32_t)(qd * 32) << 16);
                float va[64], vb[64];
                if (!(q.dbg & 2)) {
                    tmem_ld64_issue(tlane + col_a, va);
                    tmem_ld64_issue(tlane + col_b, vb);
                    tmem_ld64_wait(va); tmem_ld64_wait(vb);
                }
                if (pair == NPAIR - 1) {
                    // everything this warp needs is in registers: hand the accumulators back before the arithmetic and the stores
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sh.acc_empty[buf]);
                }
#pragma unroll
                for (int sub = 0; sub < 2; ++sub) {
                    if (q.dbg & 2) continue;
                    float* v = sub ? vb : va;
                    const int tile = CONVT ? pair : sub;
                    const int h = bh * CH_BH + (row >> 3), w = bw * BW + tile * 8 + (row & 7);
                    uint8_t* orow = reinterpret_cast<uint8_t*>(q.y) + ((((size_t)n * q.H + h) * q.W + w) * (size_t)q.Cout + (size_t)co0) * esz;
                    const bool direct = !CONVT && q.out_f32;
                    if (!direct) {
                        // the previous TMA store of this warp has finished READING the staging tile
                        if (lane == 0) bulk_wait_read();
                        __syncwarp();
                    }
#pragma unroll
                    for (int c = 0; c < 64; c += 16) {
                        if (CONVT && STATS) {
                            // BatchNorm batch statistics of (conv + bias), before the activation: column sums over this warp's 32
                            // pixels (halving butterfly), then shared-memory atomics; flushed once per CTA after the item loop
                            float a1[16], a2[16];
#pragma unroll
                            for (int j4 = 0; j4 < 16; j4 += 4) {
                                const float4 bq = bias ? __ldg(reinterpret_cast<const float4*>(bias + c + j4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                                const float bb[4] = {bq.x, bq.y, bq.z, bq.w};
#pragma unroll
                                for (int e = 0; e < 4; ++e) { const float x = v[c + j4 + e] + bb[e]; a1[j4 + e] = x; a2[j4 + e] = x * x; }
                            }
                            const float s1 = colsum16(a1, lane), s2 = colsum16(a2, lane);
                            if (!(lane & 1)) {
                                const int ci = c + colsum_index(lane);
                                atomicAdd(&sh.sstat[ci], s1);
                                atomicAdd(&sh.sstat[64 + ci], s2);
                            }
                        }
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
                        if (direct) {
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
                    if (!direct && !(q.dbg & 1)) {
                        fence_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            // conv-transpose: the parity's map is the output seen with doubled strides, indexed by INPUT pixel
                            tma_store_4d(&maps_o.m[CONVT ? half * 2 + sub : 0], smem_u32(stg), co0, bw * BW + tile * 8, bh * CH_BH + qd * 4, n);
                            bulk_commit();
                        }
                    }
                }
            }
        }
        if (lane == 0) bulk_wait_read();          // the staging tile must outlive the last store's read
        if (CONVT && STATS) {
            epi_bar();
            const int et = (int)threadIdx.x - 64;
            if (et < 64) {
                atomicAdd(q.stats + et, (double)sh.sstat[et]);
                atomicAdd(q.stats + q.Cout + et, (double)sh.sstat[64 + et]);
            }
        }
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
static int g_halo = -1, g_halo_g = 3, g_halo_s = 2, g_halo_dbg = 0, g_halo_min = 2;   // g_halo_min: items per SM from which the kernel is taken (never slower than conv_tc from 1; tools/halo_tune.py 96)
static bool g_halo_attr = false;
static long long g_halo_count[5] = {0, 0, 0, 0, 0};   // launches per kind (development aid, mstk_debug_halo_count)

static void halo_env() {
    if (g_halo >= 0) return;
    g_halo = 1;
    if (const char* s = getenv("MSTK_CONV_HALO")) g_halo = atoi(s);
    if (const char* s = getenv("MSTK_HALO_DBG")) g_halo_dbg = atoi(s);
    if (const char* s = getenv("MSTK_HALO_MIN")) g_halo_min = atoi(s) > 0 ? atoi(s) : 2;
    if (const char* s = getenv("MSTK_HALO_GS")) sscanf(s, "%d,%d", &g_halo_g, &g_halo_s);
    if (g_halo_g < 1 || g_halo_s < 2 || g_halo_g * g_halo_s > CH_MAX_SUB) { g_halo_g = 3; g_halo_s = 2; }
}

// 0 = not eligible, 1 = 3x3 stride-1 convolution, 2 = ConvTranspose2d(k4, s2, p1) with 64 output channels
static int halo_kind(const ConvP& p) {
    halo_env();
    if (!g_halo) return 0;
    if (p.dtype != MSTK_BF16 || !p.y || p.add || p.add_nc || p.has_ddim) return 0;
    if (p.Cin % 64 != 0) return 0;
    if (p.Hg % CH_BH != 0 || p.Wg % 16 != 0) return 0;
    if (p.s_in == 2) {
        // 3x3 stride-2 convolution = a 2x2 stride-1 convolution over the four parity planes of the input (kind 4)
        if (p.nphases != 1 || p.s_out != 1 || p.ph[0].ntaps != 9 || p.Cout % CH_BN != 0 || p.stats || (p.H & 1) || (p.W & 1)) return 0;
        // OPT-IN (MSTK_CONV_HALO=9 / mstk_debug_halo(9)): measured neutral against conv_tc (config-2 encoder layer 40.6 us both ways,
        // training step 6.86 vs 6.89 ms) -- with one plane per channel block an item loads four 41 KB boxes for the 72 MMAs that
        // the stride-1 layers get from one, and the A loads, not the weights, bound it
        if (p.Hg * 2 != p.H || p.Wg * 2 != p.W || !(g_halo & 8)) return 0;
        for (int t = 0; t < 9; ++t) {
            const Tap& tp = p.ph[0].taps[t];
            if (tp.dy != t / 3 - 1 || tp.dx != t % 3 - 1 || tp.slice != t) return 0;
        }
        const long long items = (long long)p.N * (p.Hg / CH_BH) * (p.Wg / 16) * (p.Cout / CH_BN);
        return (items >= (long long)g_halo_min * tc_num_sms() && items < 0x7fffffffll) ? 4 : 0;
    }
    if (p.s_in != 1 || p.Hg != p.H || p.Wg != p.W) return 0;
    if ((reinterpret_cast<uintptr_t>(p.x) & 15) || (reinterpret_cast<uintptr_t>(p.w) & 15) || (reinterpret_cast<uintptr_t>(p.y) & 15)) return 0;
    if (p.bias && (reinterpret_cast<uintptr_t>(p.bias) & 15)) return 0;
    const long long blocks = (long long)p.N * (p.Hg / CH_BH) * (p.Wg / 16);    // 16 x 16 pixel blocks
    if (p.nphases == 1 && p.s_out == 1 && p.ph[0].ntaps == 9 && p.Cout % CH_BN == 0 && !p.stats) {
        for (int t = 0; t < 9; ++t) {
            const Tap& tp = p.ph[0].taps[t];
            if (tp.dy != t / 3 - 1 || tp.dx != t % 3 - 1 || tp.slice != t) return 0;
        }
        const long long items = blocks * (p.Cout / CH_BN);
        // throughput-bound grids only (at least g_halo_min items per SM)
        return (items >= (long long)g_halo_min * tc_num_sms() && items < 0x7fffffffll) ? 1 : 0;
    }
    if (p.nphases == 4 && p.s_out == 2 && p.Cout % 64 == 0 && p.out_dtype == MSTK_BF16 && !(g_halo & 2)) {
        // conv-transpose: k4 s2 p1 (four taps per output parity, offsets -1..1) or k3 s2 p1 op1 (1 / 2 / 2 / 4 taps, offsets 0..1:
        // the data gradient of the stride-2 convolutions)
        int ntaps = 0;
        bool k3 = true, k4 = true;
        for (int ph = 0; ph < 4; ++ph) {
            if (p.ph[ph].oy != (ph >> 1) || p.ph[ph].ox != (ph & 1)) return 0;
            const int py = ph >> 1, px = ph & 1;
            if (p.ph[ph].ntaps != 4) k4 = false;
            if (p.ph[ph].ntaps != (1 + py) * (1 + px)) k3 = false;
            ntaps += p.ph[ph].ntaps;
            for (int t = 0; t < p.ph[ph].ntaps; ++t) {
                const Tap& tp = p.ph[ph].taps[t];
                if (tp.dy < -1 || tp.dy > 1 || tp.dx < -1 || tp.dx > 1 || tp.slice < 0 || tp.slice >= 16) return 0;
                if (tp.dy < 0 || tp.dx < 0 || tp.dy > py || tp.dx > px) k3 = false;
            }
        }
        const long long items = 2 * blocks * (p.Cout / 64);                             // 16 x 8 pixel tiles x 64-channel blocks
        if (items < (long long)g_halo_min * tc_num_sms() || items >= 0x7fffffffll) return 0;
        if (k4 && ntaps == 16 && p.Cout == 64) return 2;
        if (k3 && ntaps == 9 && !p.stats) return 3;
        return 0;
    }
    return 0;
}

bool conv_halo_supported(const ConvP& p) { return halo_kind(p) != 0; }

int launch_conv_halo(const ConvP& p, cudaStream_t st) {
    if (!tc_init()) return MSTK_EUNSUPPORTED;
    const int kind = halo_kind(p);
    if (!kind) return MSTK_EUNSUPPORTED;
    const bool convt = kind == 2 || kind == 3, k3 = kind == 3, s2 = kind == 4;
    const int bn = convt ? 64 : CH_BN;
    HaloParams q;
    memset(&q, 0, sizeof(q));
    q.bias = p.bias; q.y = p.y; q.stats = p.stats;
    q.N = p.N; q.H = p.Hg; q.W = p.Wg; q.Cin = p.Cin; q.Cout = p.Cout;      // the pixel grid of the items: output (= input for stride 1 / conv-transpose input)
    const int tiles = convt ? 1 : 2, hw = 8 * tiles + 2;
    q.blocks_w = p.Wg / (8 * tiles); q.blocks_h = p.Hg / CH_BH;
    q.nblocks = convt ? p.Cout / 64 : p.Cout / CH_BN;
    q.nitems = p.N * q.blocks_w * q.blocks_h * q.nblocks;
    q.KBC = p.Cin / 64;
    q.KB = s2 ? 4 * q.KBC : q.KBC;
    if (s2) {
        // parity plane of a tap: input row 2 ho + ky - 1 -> plane py = (ky - 1) & 1, plane row ho + (ky - 1 - py) / 2 (same in x)
        for (int pl = 0; pl < 4; ++pl) q.s2_ntaps[pl] = 0;
        for (int t = 0; t < 9; ++t) {
            const int dy = t / 3 - 1, dx = t % 3 - 1;
            const int py = dy & 1, px = dx & 1, oy = (dy - py) / 2, ox = (dx - px) / 2, pl = py * 2 + px;
            const int i = q.s2_ntaps[pl]++;
            q.s2_slice[pl][i] = (int8_t)t;
            q.s2_aoff[pl][i] = (int16_t)((oy + 1) * (8 * 2 + 2) + (ox + 1));
        }
    }
    // conv-transpose: 8 KB boxes.  k4: 8 per stage (two rounds of the box table = 20 MMAs per issuer round trip), 2 stages.
    // k3: 5 per stage (rounds of 5 and 4 boxes), 3 stages.
    static int ct_g = 0;
    if (!ct_g) { const char* e = getenv("MSTK_HALO_CTG"); ct_g = e ? atoi(e) : 8; if (ct_g != 4 && ct_g != 8 && ct_g != 16) ct_g = 8; }
    q.G = s2 ? 2 : !convt ? g_halo_g : k3 ? 5 : ct_g; q.S = s2 ? 3 : !convt ? g_halo_s : k3 ? 3 : 16 / ct_g; q.dbg = g_halo_dbg;
    q.nboxes = k3 ? 9 : 16;
    q.act = p.act; q.affine = p.affine; q.out_f32 = p.out_dtype == MSTK_F32 ? 1 : 0;
    q.out_scale = p.out_scale; q.out_shift = p.out_shift;
    int nslices = 9;
    if (convt) {
        nslices = k3 ? 9 : 16;
        int slice_of[4][3][3];                           // [parity py * 2 + px][dy + 1][dx + 1] -> weight slice, -1 = not a tap
        for (int ph = 0; ph < 4; ++ph) {
            for (int i = 0; i < 9; ++i) slice_of[ph][i / 3][i % 3] = -1;
            for (int t = 0; t < p.ph[ph].ntaps; ++t) {
                const Tap& tp = p.ph[ph].taps[t];
                slice_of[ph][tp.dy + 1][tp.dx + 1] = tp.slice;
            }
        }
        if (!k3) {
        // boxes of the four rounds: {parity, dy, dx}; accumulator column order P00, P01, P11, P10 = parities 0, 1, 3, 2
        static const int box[16][3] = {{0, 0, 0}, {1, 0, 0}, {3, 0, 0}, {2, 0, 0},          // round 0: centre, all four
                                       {0, -1, 0}, {1, -1, 0}, {1, 0, 1}, {3, 0, 1},        // round 1: row above (py = 0), column right (px = 1)
                                       {3, 1, 0}, {2, 1, 0}, {0, 0, -1}, {2, 0, -1},        // round 2: row below (py = 1), column left (px = 0)
                                       {0, -1, -1}, {1, -1, 1}, {3, 1, 1}, {2, 1, -1}};     // round 3: the four corners
        for (int i = 0; i < 16; ++i) {
            const int sl = slice_of[box[i][0]][box[i][1] + 1][box[i][2] + 1];
            MSTK_REQUIRE(sl >= 0, "conv(halo): conv-transpose plan does not have the k4 s2 p1 tap structure");
            q.op_slice[i] = (int16_t)sl;
        }
        // operations: {round, first box, nsel, first column / 64, dy, dx}
        static const int ops[10][6] = {{0, 0, 2, 0, 0, 0},
                                       {1, 0, 1, 0, -1, 0}, {1, 2, 1, 1, 0, 1},
                                       {2, 0, 1, 2, 1, 0}, {2, 2, 0, 0, 0, -1}, {2, 3, 0, 3, 0, -1},
                                       {3, 0, 0, 0, -1, -1}, {3, 1, 0, 1, -1, 1}, {3, 2, 0, 2, 1, 1}, {3, 3, 0, 3, 1, -1}};
        // a stage holds G / 4 rounds of the box table: the issuer's round = those rounds' operations
        const int rps = q.G / 4;
        for (int r = 0; r <= 4; ++r) q.t_first[r] = 10;
        for (int o = 9; o >= 0; --o) {
            q.t_slot[o] = (int8_t)(ops[o][1] + 4 * (ops[o][0] % rps)); q.t_nsel[o] = (int8_t)ops[o][2]; q.t_dcol[o] = (int8_t)ops[o][3];
            q.t_aoff[o] = (int16_t)((ops[o][4] + 1) * hw + (ops[o][5] + 1));
            q.t_first[ops[o][0] / rps] = (int8_t)o;
        }
        } else {
            // k3 s2 p1 op1: offsets 0..1.  Round 0 (5 boxes): the centre offset for all four parities (one N = 256 operation) and
            // the (+1, +1) corner of P11; round 1 (4 boxes): (0, +1) for P01 | P11 and (+1, 0) for P11 | P10 (N = 128 each).
            static const int box3[9][3] = {{0, 0, 0}, {1, 0, 0}, {3, 0, 0}, {2, 0, 0}, {3, 1, 1}, {1, 0, 1}, {3, 0, 1}, {3, 1, 0}, {2, 1, 0}};
            for (int i = 0; i < 9; ++i) {
                const int sl = slice_of[box3[i][0]][box3[i][1] + 1][box3[i][2] + 1];
                MSTK_REQUIRE(sl >= 0, "conv(halo): conv-transpose plan does not have the k3 s2 p1 op1 tap structure");
                q.op_slice[i] = (int16_t)sl;
            }
            // {round, first box of the round, nsel, first column / 64, dy, dx}
            static const int ops3[4][6] = {{0, 0, 2, 0, 0, 0}, {0, 4, 0, 2, 1, 1}, {1, 0, 1, 1, 0, 1}, {1, 2, 1, 2, 1, 0}};
            for (int r = 0; r <= 4; ++r) q.t_first[r] = 4;
            for (int o = 3; o >= 0; --o) {
                q.t_slot[o] = (int8_t)ops3[o][1]; q.t_nsel[o] = (int8_t)ops3[o][2]; q.t_dcol[o] = (int8_t)ops3[o][3];
                q.t_aoff[o] = (int16_t)((ops3[o][4] + 1) * hw + (ops3[o][5] + 1));
                q.t_first[ops3[o][0]] = (int8_t)o;
            }
        }
    }
    auto inv = [](int d) -> uint32_t { return d <= 1 ? 0u : (uint32_t)((0x100000000ull + (uint64_t)d - 1) / (uint64_t)d); };
    q.inv_nblocks = inv(q.nblocks); q.inv_bw = inv(q.blocks_w); q.inv_bwh = inv(q.blocks_w * q.blocks_h);
    MSTK_REQUIRE((long long)q.nitems * (q.blocks_w * q.blocks_h > q.nblocks ? q.blocks_w * q.blocks_h : q.nblocks) < (1ll << 32),
                 "conv(halo): item grid outside the range of the division-free index math");
    CUtensorMap map_a, map_b;
    HaloInMaps maps_a;
    {
        const uint64_t dims[4] = {(uint64_t)p.Cin, (uint64_t)p.W, (uint64_t)p.H, (uint64_t)p.N};
        const uint64_t strides[3] = {(uint64_t)p.Cin * 2, (uint64_t)p.W * p.Cin * 2, (uint64_t)p.H * p.W * p.Cin * 2};
        const uint32_t box[4] = {64u, (uint32_t)hw, (uint32_t)CH_HH, 1u};
        int rc = tc_encode_map(&map_a, 0, p.x, 4, dims, strides, box, 128);
        if (rc != MSTK_OK) return rc;
        for (int i = 0; i < 4; ++i) maps_a.m[i] = map_a;
    }
    if (s2) {
        // one map per parity plane: the input seen with doubled pixel strides from (py, px)
        for (int pl = 0; pl < 4; ++pl) {
            const int py = pl >> 1, px = pl & 1;
            const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(p.x) + ((size_t)py * p.W + px) * p.Cin;
            const uint64_t dims[4] = {(uint64_t)p.Cin, (uint64_t)(p.W / 2), (uint64_t)(p.H / 2), (uint64_t)p.N};
            const uint64_t strides[3] = {(uint64_t)p.Cin * 2 * 2, (uint64_t)p.W * p.Cin * 2 * 2, (uint64_t)p.H * p.W * p.Cin * 2};
            const uint32_t box[4] = {64u, (uint32_t)hw, (uint32_t)CH_HH, 1u};
            int rc = tc_encode_map(&maps_a.m[pl], 0, base, 4, dims, strides, box, 128);
            if (rc != MSTK_OK) return rc;
        }
    }
    {
        const uint64_t dims[2] = {(uint64_t)p.Cin, (uint64_t)nslices * p.Cout};
        const uint64_t strides[1] = {(uint64_t)p.Cin * 2};
        const uint32_t box[2] = {64u, (uint32_t)bn};
        int rc = tc_encode_map(&map_b, 0, p.w, 2, dims, strides, box, 128);
        if (rc != MSTK_OK) return rc;
    }
    HaloOutMaps maps_o;
    for (int i = 0; i < 4; ++i) maps_o.m[i] = map_a;
    if (convt) {
        // one map per output parity: the output [N, 2H, 2W, 64] seen with doubled pixel strides from (py, px), indexed by input pixel
        for (int ph = 0; ph < 4; ++ph) {
            const int py = ph >> 1, px = ph & 1;
            const uint8_t* base = reinterpret_cast<const uint8_t*>(p.y) + ((size_t)py * p.Wo + px) * p.Cout * 2;
            const uint64_t dims[4] = {(uint64_t)p.Cout, (uint64_t)p.W, (uint64_t)p.H, (uint64_t)p.N};
            const uint64_t strides[3] = {(uint64_t)2 * p.Cout * 2, (uint64_t)2 * p.Wo * p.Cout * 2, (uint64_t)p.Ho * p.Wo * p.Cout * 2};
            const uint32_t box[4] = {64u, 8u, 4u, 1u};
            int rc = tc_encode_map(&maps_o.m[ph], 0, base, 4, dims, strides, box, 128);
            if (rc != MSTK_OK) return rc;
        }
    } else if (!q.out_f32) {
        const uint64_t dims[4] = {(uint64_t)p.Cout, (uint64_t)p.Wg, (uint64_t)p.Hg, (uint64_t)p.N};
        const uint64_t strides[3] = {(uint64_t)p.Cout * 2, (uint64_t)p.Wg * p.Cout * 2, (uint64_t)p.Hg * p.Wg * p.Cout * 2};
        const uint32_t box[4] = {64u, 8u, 4u, 1u};
        int rc = tc_encode_map(&maps_o.m[0], 0, p.y, 4, dims, strides, box, 128);
        if (rc != MSTK_OK) return rc;
    }
    const size_t a_stage = ((size_t)hw * CH_HH * 128 + 1023) & ~(size_t)1023;
    q.off_stg = (uint32_t)(2 * a_stage + (size_t)q.G * q.S * bn * 128);
    const size_t smem = (size_t)q.off_stg + 8 * 4096 + 1024;
    MSTK_REQUIRE(smem <= 224 * 1024 && q.S <= CH_MAX_SUB && q.S >= 1, "conv(halo): shared-memory budget");
    if (!g_halo_attr) {
        cudaError_t ea = cudaFuncSetAttribute(conv_halo_kernel<false, 2, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(conv_halo_kernel<false, 2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(conv_halo_kernel<true, 1, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(conv_halo_kernel<true, 1, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
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
    cudaError_t e = s2        ? cudaLaunchKernelEx(&cfg, conv_halo_kernel<false, 2, false, true>, maps_a, map_b, maps_o, q)
                    : !convt  ? cudaLaunchKernelEx(&cfg, conv_halo_kernel<false, 2, false, false>, maps_a, map_b, maps_o, q)
                    : q.stats ? cudaLaunchKernelEx(&cfg, conv_halo_kernel<true, 1, true, false>, maps_a, map_b, maps_o, q)
                              : cudaLaunchKernelEx(&cfg, conv_halo_kernel<true, 1, false, false>, maps_a, map_b, maps_o, q);
    if (e != cudaSuccess) return cuda_fail(e, "conv_halo launch");
    ++g_halo_count[kind];
    return MSTK_OK;
}

long long halo_count(int kind) { return kind >= 1 && kind <= 4 ? g_halo_count[kind] : 0; }
void halo_set(int on) { halo_env(); g_halo = on; }

}  // namespace mstk

/* development aid (not in mstk.h): how many launches took the halo-resident kernel (1 = 3x3 convolution, 2 = conv-transpose k4) */
extern "C" long long mstk_debug_halo_count(int kind) { return mstk::halo_count(kind); }
/* development aid (not in mstk.h): 0 sends every shape back to the one-tile / persistent kernels of conv_tc.cu (1 = default) */
extern "C" int mstk_debug_halo(int on) { mstk::halo_set(on); return 0; }
