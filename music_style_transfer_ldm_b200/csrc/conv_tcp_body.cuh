// Persistent form of the tcgen05 tap-gather implicit GEMM (conv_tc.cu) for the THROUGHPUT-bound grids: the VGGish
// feature stack, the VAE / style-encoder layers at training and config-2 sizes.  There a layer is thousands of
// 128 x BN tiles, and in the one-tile-per-CTA kernel every tile pays its own prologue (barrier init, TMEM allocation,
// first-load latency ~3 k cycles) and its own epilogue + store drain (~4 k cycles) with only the SM's second resident
// CTA to overlap them (the K = 512 convT of config 2 spends 96 us on 34 GFLOP).
//
// One CTA per SM loops over tiles (static round-robin, output-channel tile fastest so that CTAs running together share
// the activation tile in L2):
//   warp 0   A (activation) TMA producer, warp 6  B (weight) TMA producer: one lane per pipeline sub-slot, running
//            AHEAD across tile boundaries (the smem ring never drains between tiles);
//   warp 1   MMA issuer: accumulator of tile i in TMEM buffer i & 1 (2 x BN columns allocated once);
//   8 epilogue warps: tile i's epilogue (tcgen05.ld, bias, BatchNorm sums, activation, staging, TMA store) runs while
//            the MMAs of tile i + 1 fill the other TMEM buffer.
// Barriers: full/empty per smem stage (as conv_tc), tfull/tempty per TMEM buffer.
// Scope: bias (+ReLU / generic activation + affine) (+BatchNorm statistics), y stored; no split-K, residual, time
// embedding or DDIM epilogue (those belong to the latency-bound UNet layers of the one-tile kernel).
#pragma once
#include "conv_tc_body.cuh"

namespace mstk {

struct TcpExtra {
    int32_t ntiles;     // mtiles * ny * nphases
    int32_t mtiles;     // tiles_w * tiles_h * tiles_n
    int32_t ny;         // Cout / BN
    uint32_t off_out;   // staging tile, separate from the stage ring (the ring stays live across tiles)
    uint32_t off_bias;
};

template <uint32_t F>
__device__ __forceinline__ void conv_tcp_body(const TcMaps& maps, const TcParams& q, const TcpExtra& x, uint8_t* smem_raw) {
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES], empty_bar[MAX_STAGES], tfull_bar[2], tempty_bar[2];
    __shared__ uint32_t tmem_base_smem;
    const bool f_stats = (F & F_STATS) && q.p.stats != nullptr;
    const bool f_actgen = (F & F_ACTGEN) != 0;
    const ConvP& p = q.p;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = q.kgroup, S = q.stages, BN = q.BN;
    const uint32_t a_bytes = TILE_M * q.KC * 2;
    const uint32_t b_bytes = BN * q.KC * 2;
    const uint32_t sub_bytes = a_bytes + ((b_bytes + 1023u) & ~1023u);
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t smem_base = smem_u32(smem);
    float* sbias = reinterpret_cast<float*>(smem + x.off_bias);

    if (threadIdx.x == 0) {
        prefetch_tmap(&maps.a[0]);
        prefetch_tmap(&maps.b);
        prefetch_tmap(&maps.out[0]);
        for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 2); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], EPI_THREADS); }
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

    if (warp == 6 || warp == 0) {
        // ===================== producers: lane j owns sub-slot j = stage * G + g of the ring =====================
        const bool is_a = warp == 0;
        if (is_a) {
            pdl_wait();
            if (lane == 0) pdl_launch_dependents();
        }
        const int j = lane;
        if (j < S * G) {
            const int s = j / G, g = j - s * G;
            const uint32_t dst = smem_base + (uint32_t)j * sub_bytes + (is_a ? 0u : a_bytes);
            int gcount = 0;                                            // groups issued by the whole CTA before this tile
            for (int t = blockIdx.x; t < x.ntiles; t += gridDim.x) {
                const int by = t % x.ny, r = t / x.ny;
                const int mt = r % x.mtiles, phase_idx = r / x.mtiles;
                const Phase& ph = p.ph[phase_idx];
                const int total_kb = ph.ntaps * q.kb_per_tap;
                const int ngroups = (total_kb + G - 1) / G;
                const int mrow = mt / q.tiles_w, tw = mt - mrow * q.tiles_w;
                const int tn = mt / (q.tiles_w * q.tiles_h), th = mrow - tn * q.tiles_h;
                const int w0 = tw * q.Wt, h0 = th * q.Ht, n0 = tn * q.Nt, co0 = by * BN;
                int first = s - gcount % S;
                if (first < 0) first += S;
                for (int grp = first; grp < ngroups; grp += S) {
                    const int cnt = total_kb - grp * G < G ? total_kb - grp * G : G;
                    const int use = (gcount + grp) / S;               // earlier uses of this stage
                    // EVERY lane of the stage observes EVERY release of it, also when a short group leaves it nothing to
                    // load: a lane that skipped a use could not tell release #u from release #u+2 by parity and would
                    // overwrite operands the MMAs have not read yet
                    if (use > 0) mbar_wait(&empty_bar[s], (uint32_t)((use - 1) & 1));
                    if (g >= cnt) continue;
                    const int kb = grp * G + g;
                    const int tap_i = kb / q.kb_per_tap, cb = kb - tap_i * q.kb_per_tap;
                    const Tap tp = ph.taps[tap_i];
                    if (is_a) {
                        int dy = tp.dy, dx = tp.dx;
                        const CUtensorMap* mA = &maps.a[0];
                        if (p.s_in == 2) {
                            const int py = dy & 1, px = dx & 1;
                            dy = (dy - py) / 2;
                            dx = (dx - px) / 2;
                            mA = &maps.a[py * 2 + px];
                        }
                        if (g == 0) mbar_expect_tx(&full_bar[s], a_bytes * (uint32_t)cnt);
                        tma_load_4d(mA, &full_bar[s], dst, cb * q.KC, w0 + dx, h0 + dy, n0);
                    } else {
                        if (g == 0) mbar_expect_tx(&full_bar[s], b_bytes * (uint32_t)cnt);
                        tma_load_2d(&maps.b, &full_bar[s], dst, cb * q.KC, tp.slice * p.Cout + co0);
                    }
                }
                gcount += ngroups;
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (converged warp, one elected lane issues) =====================
        const uint32_t idesc = keep_reg(make_idesc(BN));
        const int ksteps = keep_reg(q.KC / 16);
        const uint64_t adesc0 = make_smem_desc(smem_base, q.KC);
        const uint64_t bdesc0 = make_smem_desc(smem_base + a_bytes, q.KC);
        const uint32_t alo0 = keep_reg((uint32_t)adesc0), blo0 = keep_reg((uint32_t)bdesc0);
        const uint32_t ahi = keep_reg((uint32_t)(adesc0 >> 32)), bhi = keep_reg((uint32_t)(bdesc0 >> 32));
        const uint32_t dsub = keep_reg(sub_bytes >> 4);
        const uint32_t full0 = keep_reg(smem_u32(&full_bar[0])), empty0 = keep_reg(smem_u32(&empty_bar[0]));
        int gcount = 0, it = 0;
        for (int t = blockIdx.x; t < x.ntiles; t += gridDim.x, ++it) {
            const int phase_idx = (t / x.ny) / x.mtiles;
            const int total_kb = p.ph[phase_idx].ntaps * q.kb_per_tap;
            const int ngroups = (total_kb + G - 1) / G;
            const int buf = it & 1;
            if (it >= 2) mbar_wait(&tempty_bar[buf], (uint32_t)(((it >> 1) - 1) & 1));      // the epilogue drained this buffer
            tc_fence_after();
            const uint32_t tm = tmem_base + (uint32_t)(buf * BN);
            uint32_t acc = 0;
            int left = total_kb;
            for (int grp = 0; grp < ngroups; ++grp) {
                const int gi = gcount + grp;
                const int s = gi % S;
                mbar_wait_addr(full0 + 8u * (uint32_t)s, (uint32_t)((gi / S) & 1));
                tc_fence_after();
                const int cnt = left < G ? left : G;
                left -= cnt;
                if (elect_one()) {
                    uint32_t alo = alo0 + (uint32_t)(s * G) * dsub, blo = blo0 + (uint32_t)(s * G) * dsub;
                    for (int g = 0; g < cnt; ++g) {
                        for (int k = 0; k < ksteps; ++k) {
                            umma_bf16(tm, ((uint64_t)ahi << 32) | (alo + 2u * k), ((uint64_t)bhi << 32) | (blo + 2u * k), idesc, acc);
                            acc = 1u;
                        }
                        alo += dsub; blo += dsub;
                    }
                    umma_commit_addr(empty0 + 8u * (uint32_t)s);
                    if (grp == ngroups - 1) umma_commit_addr(smem_u32(&tfull_bar[buf]));
                }
                acc = 1u;
                __syncwarp();
            }
            gcount += ngroups;
        }
    } else {
        // ===================== epilogue warps (2..5 lower half of the columns, 7..10 upper half) =====================
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        const int w_l = row & (q.Wt - 1);
        const int h_l = (row >> q.wt_shift) & (q.Ht - 1);
        const int n_l = row >> (q.wt_shift + q.ht_shift);
        const int half = warp >= 7 ? 1 : 0;
        const int et = half ? (int)threadIdx.x - 96 : (int)threadIdx.x - 64;         // 0..255
        // the tile's columns are processed in blocks of CB <= 128 (one staging tile of 128 rows x CB columns, reused per
        // block; BN = 256 tiles have two); within a block the two warps of a TMEM lane quarter take half the columns each
        const int CB = BN < 128 ? BN : 128, nblk = BN / CB;
        const int cb_lo = CB >= 32 ? half * (CB >> 1) : 0;
        const int cb_hi = CB >= 32 ? cb_lo + (CB >> 1) : (half ? 0 : CB);
        const int blk_chunks = CB * q.out_esz / q.out_rowbytes;
        float* sstat = sbias + BN;
        const int out_cpr = q.out_rowbytes / q.out_esz;
        const int out_sh = 31 - __clz(out_cpr);
        const int out_esz = q.out_esz, out_rowbytes = q.out_rowbytes;
        uint8_t* const out_stage = smem + x.off_out;
        const int act = p.act;
        if (f_stats) { for (int i = et; i < 2 * BN; i += EPI_THREADS) sstat[i] = 0.f; }
        int it = 0;
        for (int t = blockIdx.x; t < x.ntiles; t += gridDim.x, ++it) {
            const int by = t % x.ny, r = t / x.ny;
            const int mt = r % x.mtiles, phase_idx = r / x.mtiles;
            const int mrow = mt / q.tiles_w, tw = mt - mrow * q.tiles_w;
            const int tn = mt / (q.tiles_w * q.tiles_h), th = mrow - tn * q.tiles_h;
            const int w0 = tw * q.Wt, h0 = th * q.Ht, n0 = tn * q.Nt, co0 = by * BN;
            const bool row_ok = n0 + n_l < p.N;
            (void)h_l; (void)w_l;
            if (et < BN) sbias[et] = p.bias ? __ldg(p.bias + co0 + et) : 0.f;
            epi_bar();            // bias visible; also: the previous tile's TMA store has finished reading the staging tile
            const int buf = it & 1;
            mbar_wait(&tfull_bar[buf], (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t)(buf * BN) + ((uint32_t)(qd * 32) << 16);
            for (int blk = 0; blk < nblk; ++blk) {
            const int cbeg = blk * CB + cb_lo, cend = blk * CB + cb_hi;
            if (blk > 0) epi_bar();                 // the previous block's TMA store has finished reading the staging tile
            float vnext[16];
            if (cbeg < cend) tmem_ld16_issue(taddr + (uint32_t)cbeg, vnext);
            for (int c = cbeg; c < cend; c += 16) {
                float v[16];
                tmem_ld_wait(vnext);
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) v[jj] = vnext[jj];
                if (c + 16 < cend) tmem_ld16_issue(taddr + (uint32_t)(c + 16), vnext);
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) v[jj] += sbias[c + jj];
                if (f_stats) {
                    float a1[16], a2[16];
#pragma unroll
                    for (int jj = 0; jj < 16; ++jj) {
                        const float xv = row_ok ? v[jj] : 0.f;
                        a1[jj] = xv; a2[jj] = xv * xv;
                    }
                    const float s1 = colsum16(a1, lane), s2 = colsum16(a2, lane);
                    if (!(lane & 1)) {
                        const int ci = c + colsum_index(lane);
                        atomicAdd(&sstat[ci], s1);
                        atomicAdd(&sstat[BN + ci], s2);
                    }
                }
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) {
                    float xv = v[jj];
                    if (f_actgen) {
                        xv = apply_act(xv, act);
                        if (p.affine) xv = xv * p.out_scale + p.out_shift;
                    } else if (act == MSTK_ACT_RELU) {
                        xv = fmaxf(xv, 0.f);
                    }
                    v[jj] = xv;
                }
                uint8_t* ob = out_stage + (size_t)((c - blk * CB) >> out_sh) * TILE_M * out_rowbytes;
                const int j0 = ((c & (out_cpr - 1)) * out_esz) >> 4;
                if (out_esz == 4) {
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj)
                        *reinterpret_cast<float4*>(ob + swz_off(row, j0 + jj, out_rowbytes)) =
                            make_float4(v[4 * jj], v[4 * jj + 1], v[4 * jj + 2], v[4 * jj + 3]);
                } else {
                    uint4 pk[2];
                    __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                    for (int e = 0; e < 8; ++e) hh[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
                    *reinterpret_cast<uint4*>(ob + swz_off(row, j0, out_rowbytes)) = pk[0];
                    *reinterpret_cast<uint4*>(ob + swz_off(row, j0 + 1, out_rowbytes)) = pk[1];
                }
            }
            if (blk == nblk - 1) {
                // this thread's TMEM reads are complete (tcgen05.wait::ld above): hand the accumulator buffer back
                tc_fence_before();
                mbar_arrive(&tempty_bar[buf]);
            }
            fence_async_smem();
            epi_bar();
            if (et == 0) {
                for (int ch = 0; ch < blk_chunks; ++ch)
                    tma_store_4d(&maps.out[phase_idx], smem_base + x.off_out + (uint32_t)(ch * TILE_M * out_rowbytes),
                                 co0 + blk * CB + ch * out_cpr, w0, h0, n0);
                bulk_commit();
                bulk_wait_read();
            }
            }
            if (f_stats) {      // the per-CTA column sums are complete behind the last block's barrier
                if (et < BN) {
                    atomicAdd(p.stats + co0 + et, (double)sstat[et]);
                    atomicAdd(p.stats + p.Cout + co0 + et, (double)sstat[BN + et]);
                    sstat[et] = 0.f;                 // ready for the next tile (its first atomics are behind a barrier)
                    sstat[BN + et] = 0.f;
                }
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

}  // namespace mstk
