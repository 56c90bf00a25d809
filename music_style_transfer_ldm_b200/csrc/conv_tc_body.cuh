// Body of the tcgen05 tap-gather implicit-GEMM convolution (see conv_tc.cu for the design notes), shared by the
// stand-alone kernel (conv_tc.cu) and the persistent layer-program kernel (program.cu).
#pragma once
#include <cuda.h>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace mstk {

constexpr int TC_THREADS = 352;     // warps: 0 A producer, 1 MMA issuer, 2..5 + 7..10 epilogue (column halves), 6 B producer
constexpr int EPI_THREADS = 256;
constexpr int TC_MIN_CTAS = 2;      // registers capped at 65536 / (2 * 352) = 93 so that two CTAs fit an SM
constexpr int MAX_STAGES = 8;

struct TcMaps {
    CUtensorMap a[4];     // input, one per stride-2 parity view (a[0] only for stride 1)
    CUtensorMap b;        // packed weights [taps*cout][cin]
    CUtensorMap out[4];   // output, one per conv-transpose phase (out[0] only for conv)
    CUtensorMap add[4];   // residual tensor, same geometry as out
};

struct TcParams {
    ConvP p;
    int32_t Wt, Ht, Nt;            // tile box; Wt*Ht*Nt == 128
    int32_t tiles_w, tiles_h, tiles_n;
    int32_t KC;                    // channels per k-block: 64 (SWIZZLE_128B) or 32 (SWIZZLE_64B)
    int32_t BN;                    // output channels per CTA (16..128, multiple of 16)
    int32_t kb_per_tap;            // Cin / KC
    int32_t splitk;
    int32_t stages;
    int32_t tmem_cols;
    int32_t out_esz;               // bytes per output element (2 or 4)
    int32_t out_rowbytes;          // bytes per row of one output staging chunk (<= 128)
    int32_t out_chunks;            // BN*out_esz / out_rowbytes
    int32_t add_rowbytes, add_chunks;
    int32_t use_tma_store;         // y written through the staging tile + TMA
    int32_t use_tma_add;           // residual prefetched by TMA
    uint32_t off_stage0;           // shared-memory carve-up (bytes from the 1024-aligned base)
    uint32_t off_part, off_out, off_add, off_bias;
    int32_t kgroup;                // k-blocks per pipeline stage (one barrier round trip and one commit per group)
    // division-free index math on the prologue's critical path: ceil(2^32 / d) multipliers (0 for d == 1, see fdiv)
    // and shifts for the power-of-two tile box and split-K factor; filled by launch_conv_tc
    uint32_t inv_tiles_w, inv_tiles_h, inv_tiles_wh, inv_kb_per_tap, inv_kgroup;
    int32_t wt_shift, ht_shift, splitk_shift;
    int32_t direct_store;          // epilogue writes y from registers instead of staging + TMA store
    int32_t cluster;               // split-K partners are one thread-block cluster (1,1,splitk): reduce through DSMEM, no scratch
    float* ws;                     // split-K workspace
    unsigned int* counters;        // split-K tile counters
    long long* dbg;                // optional timeline of CTA (0,0,0): 8 clock64 stamps
};
#define DBG_STAMP(i) do { if (q.dbg && (PROG ? blockIdx.x == 0 : (bx == 0 && by == 0 && bz == 0))) q.dbg[i] = clock64(); } while (0)

// ---------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------
// Feature flags: the kernel is instantiated per feature set so that each instance's epilogue is a
// few hundred instructions (the all-in-one version was ~4000 SASS instructions = 64 KB of code and
// the epilogue ran at instruction-fetch speed).  A flag that is set still honours the run-time
// argument; a flag that is clear compiles the feature out.
enum : uint32_t { F_SPLITK = 1, F_STATS = 2, F_ADDNC = 4, F_ADD = 8, F_DDIM = 16, F_STORE = 32, F_ACTGEN = 64, F_ALL = 127 };

// Column sums over the 32 rows held by a warp (one row per lane, 16 columns per lane): halving butterfly,
// 16 shuffles instead of 16 x 5.  Returns the sum of column ((lane >> 1) & 15 bit-reversed as below) in every
// lane; lanes l and l^1 hold the same column.  Column index of a lane: colsum_index(lane).
__device__ __forceinline__ float colsum16(float (&a)[16], int lane) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool up = (lane & 16) != 0;
        const float send = up ? a[i] : a[i + 8], mine = up ? a[i + 8] : a[i];
        a[i] = mine + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const bool up = (lane & 8) != 0;
        const float send = up ? a[i] : a[i + 4], mine = up ? a[i + 4] : a[i];
        a[i] = mine + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const bool up = (lane & 4) != 0;
        const float send = up ? a[i] : a[i + 2], mine = up ? a[i + 2] : a[i];
        a[i] = mine + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    {
        const bool up = (lane & 2) != 0;
        const float send = up ? a[0] : a[1], mine = up ? a[1] : a[0];
        a[0] = mine + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    return a[0] + __shfl_xor_sync(0xffffffffu, a[0], 1);
}
__device__ __forceinline__ int colsum_index(int lane) {
    return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
}

// n / d for small non-negative n with inv = ceil(2^32 / d) (exact while n * d < 2^32; inv == 0 encodes d == 1)
__device__ __forceinline__ int fdiv(int n, uint32_t inv) { return inv ? (int)__umulhi((uint32_t)n, inv) : n; }

// Per-CTA resources that outlive one tile (static shared memory of the calling kernel).
struct TcShared {
    uint64_t full_bar[MAX_STAGES];
    uint64_t empty_bar[MAX_STAGES];
    uint64_t accum_bar, add_bar, part_bar;
    uint32_t tmem_base;
    uint32_t is_last;
};

// One output tile (bx, by, bz) of a launch whose grid would be (gx, gy, *).  PROG = false: the body of the
// stand-alone kernel (allocates and frees its TMEM columns, fresh mbarriers).  PROG = true: called in a loop by the
// persistent program kernel (program.cu): TMEM is allocated once by the caller (sh.tmem_base), the mbarriers are
// invalidated and re-initialised per tile, all TC_THREADS threads must call it.
template <uint32_t F, bool PROG>
__device__ __forceinline__ void conv_tc_body(const TcMaps& maps, const TcParams& q, const int bx, const int by, const int bz,
                                             const int gx, const int gy, TcShared& sh, uint8_t* smem_raw) {
    const bool f_split = (F & F_SPLITK) && q.splitk > 1;
    const bool f_stats = (F & F_STATS) && q.p.stats != nullptr;
    const bool f_addnc = (F & F_ADDNC) && q.p.add_nc != nullptr;
    const bool f_add = (F & F_ADD) && q.use_tma_add;
    const bool f_ddim = (F & F_DDIM) && q.p.has_ddim;
    const bool f_store = (F & F_STORE) && q.use_tma_store;
    const bool f_actgen = (F & F_ACTGEN) != 0;
    uint64_t* const full_bar = sh.full_bar;
    uint64_t* const empty_bar = sh.empty_bar;
    uint64_t& accum_bar = sh.accum_bar;
    uint64_t& add_bar = sh.add_bar;
    uint64_t& part_bar = sh.part_bar;
    uint32_t& tmem_base_smem = sh.tmem_base;
    uint32_t& is_last_smem = sh.is_last;

    const ConvP& p = q.p;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // tile coordinates (no integer divisions: each is a ~100-cycle dependent chain in front of the first TMA)
    const int mt = bx;
    const int mrow = fdiv(mt, q.inv_tiles_w);                 // mt / tiles_w
    const int tw = mt - mrow * q.tiles_w;
    const int tn = fdiv(mt, q.inv_tiles_wh);                  // mt / (tiles_w * tiles_h)
    const int th = mrow - tn * q.tiles_h;
    const int w0 = tw * q.Wt, h0 = th * q.Ht, n0 = tn * q.Nt;
    const int co0 = by * q.BN;
    const int phase_idx = bz >> q.splitk_shift;
    const int ks = bz - (phase_idx << q.splitk_shift);
    const Phase& ph = p.ph[phase_idx];
    const int total_kb = ph.ntaps * q.kb_per_tap;
    const int kb_begin = (ks * total_kb) >> q.splitk_shift;
    const int kb_end = ((ks + 1) * total_kb) >> q.splitk_shift;
    const int nkb = kb_end - kb_begin;

    // shared-memory carve-up (1024-byte aligned for the 128B swizzle atom)
    const uint32_t a_bytes = TILE_M * q.KC * 2;
    const uint32_t b_bytes = q.BN * q.KC * 2;
    const uint32_t b_bytes_al = (b_bytes + 1023u) & ~1023u;
    const uint32_t sub_bytes = a_bytes + b_bytes_al;        // one k-block: A tile + B tile
    // pointer arithmetic (not an integer round trip) so that the compiler keeps the shared address space: LDS/STS/ATOMS
    // instead of generic LD/ST/ATOM in the epilogue
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    float* sbias = reinterpret_cast<float*>(smem + q.off_bias);

    if (threadIdx.x == 0) {
        DBG_STAMP(0);
        prefetch_tmap(&maps.a[0]);
        prefetch_tmap(&maps.b);
        if (f_store) prefetch_tmap(&maps.out[phase_idx]);
        if (PROG) {
            // the previous tile's barriers are quiescent (the body ends in __syncthreads): invalidate, then re-arm
            for (int s = 0; s < MAX_STAGES; ++s) { mbar_inval(&full_bar[s]); mbar_inval(&empty_bar[s]); }
            mbar_inval(&accum_bar); mbar_inval(&add_bar); mbar_inval(&part_bar);
        }
        for (int s = 0; s < (PROG ? MAX_STAGES : q.stages); ++s) {
            mbar_init(&full_bar[s], 2);          // one arrive.expect_tx from each of the two producers
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&accum_bar, 1);
        mbar_init(&add_bar, 1);
        mbar_init(&part_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (!PROG && warp == 1) {
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

    // K loop.  Two measurements on B200 shape it (tools/tma_probe.cu and the clock64 stamps, tools/tc_timeline.py):
    //  * TMA loads issued from one instruction stream (a single thread, or a warp running the loop in lock-step)
    //    complete a full latency (~600 cycles) apart however many stages are open; loads issued by divergent lanes
    //    that each run their own loop overlap.  So every sub-slot of the pipeline belongs to one producer lane.
    //  * One round of the MMA issuer (barrier wait, tcgen05.mma x n, commit) costs ~600 cycles + ~15 per MMA even
    //    with every operand already in shared memory, so the round must carry many MMAs: a pipeline stage holds
    //    `kgroup` consecutive k-blocks (each its own A box + B box) behind ONE full/empty barrier pair.
    // Sub-slot j = stage * kgroup + g lives at off_stage0 + j * sub_bytes (A tile, then the B tile).
    //  * Issuing one TMA load occupies its warp for ~200 cycles and the loads of one warp's lanes are issued one after
    //    the other, so the producer jobs are spread over all six non-MMA warps: weights (B, no grid dependency) on
    //    warps 6, 2, 3 and activations (A, after the grid dependency) on warps 0, 4, 5; the epilogue warps go on to
    //    their own work when their producer lanes have issued their last load.
    const int G = q.kgroup;
    const int ngroups = fdiv(nkb + G - 1, q.inv_kgroup);
    const int nsub = q.stages * G;
    const int b_idx = warp == 6 ? 0 : warp == 2 ? 1 : warp == 3 ? 2 : -1;
    const int a_idx = warp == 0 ? 0 : warp == 4 ? 1 : warp == 5 ? 2 : -1;
    if (b_idx >= 0) {
        // ===================== B producers (weights) =====================
        const int j = lane * 3 + b_idx;
        if (j < nsub) {
            const int s = fdiv(j, q.inv_kgroup), g = j - s * G;
            const uint32_t sb = smem_base + q.off_stage0 + (uint32_t)j * sub_bytes + a_bytes;
            int use = 0;
            for (int grp = s; grp < ngroups; grp += q.stages, ++use) {
                const int cnt = nkb - grp * G < G ? nkb - grp * G : G;     // k-blocks in this group (the last may be short)
                if (g >= cnt) break;
                const int kb = kb_begin + grp * G + g;
                const int tap_i = fdiv(kb, q.inv_kb_per_tap);
                const int cb = kb - tap_i * q.kb_per_tap;
                if (use > 0) mbar_wait(&empty_bar[s], (uint32_t)((use - 1) & 1));
                // lane g == 0 posts the group's byte count; its siblings' complete_tx may land first (the transaction
                // count goes transiently negative, the phase cannot complete before both operands' counts are posted)
                if (g == 0) mbar_expect_tx(&full_bar[s], b_bytes * (uint32_t)cnt);
                tma_load_2d(&maps.b, &full_bar[s], sb, cb * q.KC, ph.taps[tap_i].slice * p.Cout + co0);
            }
        }
    }
    if (a_idx >= 0) {
        // ===================== A producers (activations) =====================
        pdl_wait();
        if (warp == 0 && lane == 0) pdl_launch_dependents();
        const int j = lane * 3 + a_idx;
        if (j < nsub) {
            const int s = fdiv(j, q.inv_kgroup), g = j - s * G;
            const uint32_t sa = smem_base + q.off_stage0 + (uint32_t)j * sub_bytes;
            int use = 0;
            for (int grp = s; grp < ngroups; grp += q.stages, ++use) {
                const int cnt = nkb - grp * G < G ? nkb - grp * G : G;
                if (g >= cnt) break;
                const int kb = kb_begin + grp * G + g;
                const int tap_i = fdiv(kb, q.inv_kb_per_tap);
                const int cb = kb - tap_i * q.kb_per_tap;
                const Tap t = ph.taps[tap_i];
                int dy = t.dy, dx = t.dx;
                const CUtensorMap* mA = &maps.a[0];
                if (p.s_in == 2) {
                    const int py = dy & 1, px = dx & 1;        // parity view (works for negative offsets)
                    dy = (dy - py) / 2;
                    dx = (dx - px) / 2;
                    mA = &maps.a[py * 2 + px];
                }
                if (use > 0) mbar_wait(&empty_bar[s], (uint32_t)((use - 1) & 1));
                if (g == 0) mbar_expect_tx(&full_bar[s], a_bytes * (uint32_t)cnt);
                tma_load_4d(mA, &full_bar[s], sa, cb * q.KC, w0 + dx, h0 + dy, n0);
                if (j == 0 && use == 0) DBG_STAMP(2);
            }
        } else if (warp == 0 && lane == 31 && f_add) {
            // residual tile for the epilogue (with split-K any CTA can be the one that finalises)
            mbar_expect_tx(&add_bar, (uint32_t)(TILE_M * q.BN * 2));
            const int cpr = q.add_rowbytes >> 1;
            for (int ch = 0; ch < q.add_chunks; ++ch)
                tma_load_4d(&maps.add[phase_idx], &add_bar, smem_base + q.off_add + (uint32_t)(ch * TILE_M * q.add_rowbytes),
                            co0 + ch * cpr, w0, h0, n0);
        }
    }
    if (warp == 1) {
        // ===================== MMA issuer =====================
        // The whole warp runs the loop converged (all lanes wait on the stage barrier) and one elected lane issues:
        // under an ordinary `lane == 0` branch every tcgen05 instruction is wrapped in a serialisation loop.  Loop
        // invariants are pinned in registers and the barriers addressed by their shared-space address, so the loop
        // body is the wait, the MMAs and the commit and nothing else.
        {
            const uint32_t idesc = keep_reg(make_idesc(q.BN));
            const int ksteps = keep_reg(q.KC / 16);
            const int stages = keep_reg(q.stages);
            const int ng = keep_reg(ngroups);
            const int gsz = keep_reg(G);
            const uint32_t tm = keep_reg(tmem_base);
            const uint64_t adesc0 = make_smem_desc(smem_base + q.off_stage0, q.KC);
            const uint64_t bdesc0 = make_smem_desc(smem_base + q.off_stage0 + a_bytes, q.KC);
            // both descriptors advance together; only the 14-bit start-address field (>>4) changes and it cannot carry
            const uint32_t alo0 = keep_reg((uint32_t)adesc0), blo0 = keep_reg((uint32_t)bdesc0);
            const uint32_t ahi = keep_reg((uint32_t)(adesc0 >> 32)), bhi = keep_reg((uint32_t)(bdesc0 >> 32));
            const uint32_t dsub = keep_reg(sub_bytes >> 4);
            const uint32_t dstage = keep_reg((sub_bytes >> 4) * (uint32_t)G);
            const uint32_t full0 = keep_reg(smem_u32(&full_bar[0])), empty0 = keep_reg(smem_u32(&empty_bar[0]));
            const uint32_t accum_a = keep_reg(smem_u32(&accum_bar));
            const bool dbg_on = q.dbg && (PROG ? blockIdx.x == 0 : (bx == 0 && by == 0 && bz == 0)) && lane == 0;
            int s = 0, left = keep_reg(nkb);
            uint32_t par = 0, soff = 0, acc = 0;
            for (int grp = 0; grp < ng; ++grp) {
                mbar_wait_addr(full0 + 8u * (uint32_t)s, par);
                tc_fence_after();
                if (dbg_on) {
                    if (grp == 0) q.dbg[3] = clock64();
                    if (grp == ng - 1) q.dbg[4] = clock64();
                    if (grp < 8) q.dbg[16 + grp * 8] = clock64();
                }
                const int cnt = left < gsz ? left : gsz;
                left -= cnt;
                if (elect_one()) {
                    uint32_t alo = alo0 + soff, blo = blo0 + soff;
                    for (int g = 0; g < cnt; ++g) {
                        // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in the (>>4) address field
                        if (ksteps == 4) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                umma_bf16(tm, ((uint64_t)ahi << 32) | (alo + 2u * k), ((uint64_t)bhi << 32) | (blo + 2u * k), idesc, acc);
                                acc = 1u;
                            }
                        } else {
                            for (int k = 0; k < ksteps; ++k) {
                                umma_bf16(tm, ((uint64_t)ahi << 32) | (alo + 2u * k), ((uint64_t)bhi << 32) | (blo + 2u * k), idesc, acc);
                                acc = 1u;
                            }
                        }
                        alo += dsub; blo += dsub;
                    }
                    umma_commit_addr(empty0 + 8u * (uint32_t)s);     // frees the stage when these MMAs retire
                    if (grp == ng - 1) umma_commit_addr(accum_a);     // accumulator complete
                }
                acc = 1u;
                __syncwarp();
                if (++s == stages) { s = 0; par ^= 1u; soff = 0; }
                else soff += dstage;
            }
        }
    } else if (warp >= 2 && warp != 6) {
        // ===================== epilogue (warps 2..5: lower half of the columns, 7..10: upper half) =====================
        // one warp per scheduler leaves the epilogue bound by its own instruction latencies; two warps per TMEM lane
        // quarter, each taking half of the BN columns, halve it
        const int qd = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = qd * 32 + lane;          // GEMM row inside the tile
        const int w_l = row & (q.Wt - 1);                        // the tile box sides are powers of two
        const int h_l = (row >> q.wt_shift) & (q.Ht - 1);
        const int n_l = row >> (q.wt_shift + q.ht_shift);
        const int n_img = n0 + n_l, hg = h0 + h_l, wg = w0 + w_l;
        const bool row_ok = n_img < p.N;
        const int half = warp >= 7 ? 1 : 0;
        const int et = half ? (int)threadIdx.x - 96 : (int)threadIdx.x - 64;         // 0..255
        const int BN = q.BN;
        const int cbeg = BN >= 32 ? half * (BN >> 1) : 0;             // this warp's columns [cbeg, cend)
        const int cend = BN >= 32 ? cbeg + (BN >> 1) : (half ? 0 : BN);

        // ---- work that does not need the accumulator: overlaps the main loop ----
        if (et < BN) sbias[et] = p.bias ? __ldg(p.bias + co0 + et) : 0.f;   // parameters: safe before the grid dependency
        float* sstat = sbias + BN;                // [2][BN] per-CTA column sums of the BatchNorm statistics
        if (f_stats) { for (int i = et; i < 2 * BN; i += EPI_THREADS) sstat[i] = 0.f; }
        pdl_wait();
        const size_t pix_off = row_ok ? ((((size_t)n_img * p.Ho + (size_t)(hg * p.s_out + ph.oy)) * p.Wo + (size_t)(wg * p.s_out + ph.ox)) *
                                         (size_t)p.Cout + (size_t)co0)
                                      : 0;
        float xc[16], anc[16];
        const float* xcur_row = f_ddim ? p.ddim.x_cur + pix_off : nullptr;
        const float* anc_row = f_addnc ? p.add_nc + (size_t)(row_ok ? n_img : 0) * p.Cout + co0 : nullptr;
        if (xcur_row && row_ok && cbeg < cend) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                const float4 t = *reinterpret_cast<const float4*>(xcur_row + cbeg + j);
                xc[j] = t.x; xc[j + 1] = t.y; xc[j + 2] = t.z; xc[j + 3] = t.w;
            }
        }
        if (anc_row && cbeg < cend) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(anc_row + cbeg + j));
                anc[j] = t.x; anc[j + 1] = t.y; anc[j + 2] = t.z; anc[j + 3] = t.w;
            }
        }
        epi_bar();                               // sbias visible to all epilogue threads

        mbar_wait(&accum_bar, 0);
        if (et == 0) DBG_STAMP(5);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(qd * 32) << 16);
        uint8_t* part = smem + q.off_part;
        const int pm = (BN >= 32 ? 8 : BN / 4) - 1;       // XOR mask of the private partial-tile layout
        bool finalize = true;

        if (f_split) {
            // ---- split-K: stage the fp32 partial tile, reduce it into L2 with the copy engine ----
            for (int c = cbeg; c < cend; c += 16) {
                float v[16];
                tmem_ld16(taddr + (uint32_t)c, v);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<float4*>(part + (size_t)row * BN * 4 + (size_t)(((c >> 2) + j) ^ (row & pm)) * 16) =
                        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            }
            if (!PROG && q.cluster) {
                // ---- cluster split-K: the partners' partial tiles meet in distributed shared memory ----
                // partner r owns rows [r * 128/splitk, ...): it sums that slice over all partners and writes the sum
                // into partner 0's tile, which then runs the ordinary finalisation from its own shared memory.  Two
                // cluster barriers instead of an L2 reduction, a counter and a reload (~6.6 k -> ~2.5 k cycles).
                if (et == 0) DBG_STAMP(8);
                cluster_sync_all();
                const int sk = q.splitk;
                const int rows_per = TILE_M >> q.splitk_shift;
                const int pieces = rows_per * (BN >> 2);              // 16-byte pieces of this partner's slice
                const uint32_t part_a = smem_u32(part);
                const uint32_t dst0 = cluster_map(part_a, 0u);
                for (int i = et; i < pieces; i += EPI_THREADS) {
                    const int r = ks * rows_per + i / (BN >> 2), c4 = i % (BN >> 2);
                    const uint32_t off = (uint32_t)r * (uint32_t)(BN * 4) + (uint32_t)((c4 ^ (r & pm)) * 16);
                    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int pr = 0; pr < sk; ++pr) {
                        const float4 v = ld_cluster_f4(cluster_map(part_a, (uint32_t)pr) + off);
                        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                    }
                    st_cluster_f4(dst0 + off, acc);
                }
                if (et == 0) DBG_STAMP(9);
                cluster_sync_all();
                if (et == 0) DBG_STAMP(10);
                finalize = ks == 0;
            } else {
            fence_async_smem();
            epi_bar();
            if (et == 0) DBG_STAMP(8);
            const size_t tile_lin = ((size_t)phase_idx * gx + mt) * gy + by;
            float* ws_tile = q.ws + tile_lin * (size_t)(TILE_M * BN);
            const uint32_t tile_bytes = (uint32_t)(TILE_M * BN * 4);
            if (et == 0) {
                bulk_reduce_add_f32(ws_tile, smem_u32(part), tile_bytes);
                bulk_commit();
                bulk_wait_all();
                __threadfence();
                const unsigned int old = atomicAdd(q.counters + tile_lin, 1u);
                const unsigned int last = (old == (unsigned int)(q.splitk - 1)) ? 1u : 0u;
                if (last) {
                    __threadfence();
                    fence_async_all();
                    mbar_expect_tx(&part_bar, tile_bytes);
                    bulk_load(smem_u32(part), ws_tile, tile_bytes, &part_bar);
                    q.counters[tile_lin] = 0u;
                }
                is_last_smem = last;
                DBG_STAMP(9);
            }
            epi_bar();
            finalize = is_last_smem != 0;
            if (finalize) {
                mbar_wait(&part_bar, 0);
                // scratch back to zero for the next launch (the tile now lives in shared memory)
                float4* z = reinterpret_cast<float4*>(ws_tile);
                for (int i = et; i < TILE_M * BN / 4; i += EPI_THREADS) __stcg(z + i, make_float4(0.f, 0.f, 0.f, 0.f));
                if (et == 0) DBG_STAMP(10);
            }
            }
        }

        if (finalize) {
            if (f_add) mbar_wait(&add_bar, 0);
            const int out_cpr = q.out_rowbytes / q.out_esz;
            const int add_cpr = q.add_rowbytes >> 1;
            // columns per staging chunk are powers of two (row bytes 32/64/128): shifts, not divisions, in the loop
            const int out_sh = 31 - __clz(out_cpr), add_sh = 31 - __clz(add_cpr);
            const int out_esz = q.out_esz, out_rowbytes = q.out_rowbytes, add_rowbytes = q.add_rowbytes;
            uint8_t* const out_stage = smem + q.off_out;
            const uint8_t* const add_stage = smem + q.off_add;
            const int act = p.act;
            const bool direct = q.direct_store != 0;
            float vnext[16];
            if (!f_split && cbeg < cend) tmem_ld16_issue(taddr + (uint32_t)cbeg, vnext);
            for (int c = cbeg; c < cend; c += 16) {
                float v[16];
                if (f_split) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 t = *reinterpret_cast<const float4*>(part + (size_t)row * BN * 4 + (size_t)(((c >> 2) + j) ^ (row & pm)) * 16);
                        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
                    }
                } else {
                    // the next chunk's TMEM load is in flight while this one is processed
                    tmem_ld_wait(vnext);
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = vnext[j];
                    if (c + 16 < cend) tmem_ld16_issue(taddr + (uint32_t)(c + 16), vnext);
                }
                if (c == cbeg && et == 0) DBG_STAMP(13);
                // prefetch the next chunk's per-row operands
                float xn_[16], an_[16];
                const bool more = c + 16 < cend;
                if (more && xcur_row && row_ok) {
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        const float4 t = *reinterpret_cast<const float4*>(xcur_row + c + 16 + j);
                        xn_[j] = t.x; xn_[j + 1] = t.y; xn_[j + 2] = t.z; xn_[j + 3] = t.w;
                    }
                }
                if (more && anc_row) {
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        const float4 t = __ldg(reinterpret_cast<const float4*>(anc_row + c + 16 + j));
                        an_[j] = t.x; an_[j + 1] = t.y; an_[j + 2] = t.z; an_[j + 3] = t.w;
                    }
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] += sbias[c + j];
                if (f_stats) {
                    float a1[16], a2[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float x = row_ok ? v[j] : 0.f;
                        a1[j] = x; a2[j] = x * x;
                    }
                    const float s1 = colsum16(a1, lane), s2 = colsum16(a2, lane);
                    if (!(lane & 1)) {
                        const int ci = c + colsum_index(lane);
                        atomicAdd(&sstat[ci], s1);
                        atomicAdd(&sstat[BN + ci], s2);
                    }
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    float x = v[j];
                    if (f_actgen) {
                        x = apply_act(x, act);
                        if (p.affine) x = x * p.out_scale + p.out_shift;
                    } else if (act == MSTK_ACT_RELU) {
                        x = fmaxf(x, 0.f);
                    }
                    v[j] = x;
                }
                if (c == cbeg && et == 0) DBG_STAMP(14);
                if (anc_row) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += anc[j];
                }
                if (f_add) {
                    const uint8_t* ab = add_stage + (size_t)(c >> add_sh) * TILE_M * add_rowbytes;
                    const int j0 = ((c & (add_cpr - 1)) * 2) >> 4;
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const uint4 t = *reinterpret_cast<const uint4*>(ab + swz_off(row, j0 + h2, add_rowbytes));
                        const __nv_bfloat162* hh = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = __bfloat1622float2(hh[e]);
                            v[8 * h2 + 2 * e] += f.x; v[8 * h2 + 2 * e + 1] += f.y;
                        }
                    }
                }
                if (f_ddim && row_ok) {
                    const DdimP& d = p.ddim;
                    float xnew[16], x0v[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float eps = v[j];
                        const float dt = d.s1m_t * eps;
                        const float x0 = (xc[j] - dt) / d.sab_t;
                        const float dn = d.s1m_n * eps;
                        xnew[j] = d.sab_n * x0 + dn + d.eta * (dn - dt);
                        x0v[j] = x0;
                    }
                    float* xo = d.x_next + pix_off + c;
#pragma unroll
                    for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(xo + j) = make_float4(xnew[j], xnew[j + 1], xnew[j + 2], xnew[j + 3]);
                    if (d.x0_pred) {
                        float* po = d.x0_pred + pix_off + c;
#pragma unroll
                        for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(po + j) = make_float4(x0v[j], x0v[j + 1], x0v[j + 2], x0v[j + 3]);
                    }
                    if (d.x_next_lp) {
                        uint4 pk[2];
                        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                        for (int e = 0; e < 8; ++e) hh[e] = __floats2bfloat162_rn(xnew[2 * e], xnew[2 * e + 1]);
                        uint4* lo = reinterpret_cast<uint4*>(d.x_next_lp + pix_off + c);
                        lo[0] = pk[0]; lo[1] = pk[1];
                    }
                }
                if (f_store && direct) {
                    // straight from registers: 32 (bf16) / 64 (fp32) contiguous bytes per row and chunk, whole sectors;
                    // nothing to stage, no bulk store to wait for at the end of the CTA
                    if (row_ok) {
                        if (out_esz == 4) {
                            float4* o4 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.y) + pix_off + c);
#pragma unroll
                            for (int j = 0; j < 4; ++j) o4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                        } else {
                            uint4 pk[2];
                            __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                            for (int e = 0; e < 8; ++e) hh[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
                            uint4* o4 = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.y) + pix_off + c);
                            o4[0] = pk[0]; o4[1] = pk[1];
                        }
                    }
                } else if (f_store) {
                    uint8_t* ob = out_stage + (size_t)(c >> out_sh) * TILE_M * out_rowbytes;
                    const int j0 = ((c & (out_cpr - 1)) * out_esz) >> 4;
                    if (out_esz == 4) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<float4*>(ob + swz_off(row, j0 + j, out_rowbytes)) =
                                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    } else {
                        uint4 pk[2];
                        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                        for (int e = 0; e < 8; ++e) hh[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
                        *reinterpret_cast<uint4*>(ob + swz_off(row, j0, out_rowbytes)) = pk[0];
                        *reinterpret_cast<uint4*>(ob + swz_off(row, j0 + 1, out_rowbytes)) = pk[1];
                    }
                }
                if (c == cbeg && et == 0) DBG_STAMP(15);
                if (more) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) { xc[j] = xn_[j]; anc[j] = an_[j]; }
                }
            }
            if (et == 0) DBG_STAMP(11);
            if (f_stats) {
                epi_bar();
                if (et < BN) {
                    atomicAdd(p.stats + co0 + et, (double)sstat[et]);
                    atomicAdd(p.stats + p.Cout + co0 + et, (double)sstat[BN + et]);
                }
            }
            if (f_store && !direct) {
                fence_async_smem();
                epi_bar();
                if (et == 0) {
                    DBG_STAMP(12);
                    for (int ch = 0; ch < q.out_chunks; ++ch)
                        tma_store_4d(&maps.out[phase_idx], smem_base + q.off_out + (uint32_t)(ch * TILE_M * q.out_rowbytes),
                                     co0 + ch * out_cpr, w0, h0, n0);
                    bulk_commit();
                    bulk_wait_read();
                }
            }
        }
        if (et == 0) DBG_STAMP(6);
        tc_fence_before();
    }
    if (!PROG && f_split && q.cluster && !(warp >= 2 && warp != 6)) {
        // warps 0, 1, 6 have no part in the reduction but every thread of the cluster counts in its two barriers
        __syncwarp();
        cluster_sync_all();
        cluster_sync_all();
    }

    __syncthreads();
    if (!PROG && warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)q.tmem_cols) : "memory");
        if (lane == 0) DBG_STAMP(7);
    }
}

}  // namespace mstk
