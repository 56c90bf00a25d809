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
//             that the small-M layers of the UNet still spread over all 148 SMs.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer
// (single thread issues tcgen05.mma, accumulator in TMEM), warps 2..5 = epilogue.
//
// Epilogue (v2, after the round-1a profile showed it was 60-80 % of the kernel):
//   * everything the epilogue needs that does not depend on the accumulator is fetched
//     WHILE the MMAs run: bias -> smem, residual tile (`add`) -> smem by TMA, x_cur of the
//     fused DDIM update -> registers;
//   * tcgen05.ld 32 columns at a time; the finished tile is staged in shared memory in the
//     TMA swizzle layout (bank-conflict free 16-byte stores) and written with ONE TMA tensor
//     store per 128-byte column chunk (coalescing is done by the copy engine; out-of-range
//     rows are clipped by the tensor map);
//   * split-K partials are reduced in L2 by the copy engine (cp.reduce.async.bulk .add.f32 of
//     the whole contiguous fp32 tile) instead of per-lane REDs; the last CTA to arrive pulls
//     the tile back with one bulk copy, runs the epilogue and re-zeroes the scratch.
// Programmatic dependent launch: the prologue (barrier init, TMEM allocation, descriptor
// prefetch) AND the weight (B operand) loads of the first pipeline stages are issued before
// griddepcontrol.wait, i.e. they overlap the tail of the previous layer's kernel.
#include <cuda.h>
#include <cudaTypedefs.h>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace mstk {

constexpr int TC_THREADS = 192;
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
    float* ws;                     // split-K workspace
    unsigned int* counters;        // split-K tile counters
    long long* dbg;                // optional timeline of CTA (0,0,0): 8 clock64 stamps
};
#define DBG_STAMP(i) do { if (q.dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) q.dbg[i] = clock64(); } while (0)

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

template <uint32_t F>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcParams q) {
    const bool f_split = (F & F_SPLITK) && q.splitk > 1;
    const bool f_stats = (F & F_STATS) && q.p.stats != nullptr;
    const bool f_addnc = (F & F_ADDNC) && q.p.add_nc != nullptr;
    const bool f_add = (F & F_ADD) && q.use_tma_add;
    const bool f_ddim = (F & F_DDIM) && q.p.has_ddim;
    const bool f_store = (F & F_STORE) && q.use_tma_store;
    const bool f_actgen = (F & F_ACTGEN) != 0;
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ __align__(8) uint64_t add_bar;
    __shared__ __align__(8) uint64_t part_bar;
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
    const uint32_t smem_base = smem_u32(smem);
    float* sbias = reinterpret_cast<float*>(smem + q.off_bias);

    if (threadIdx.x == 0) {
        DBG_STAMP(0);
        prefetch_tmap(&maps.a[0]);
        prefetch_tmap(&maps.b);
        if (f_store) prefetch_tmap(&maps.out[phase_idx]);
        for (int s = 0; s < q.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&accum_bar, 1);
        mbar_init(&add_bar, 1);
        mbar_init(&part_bar, 1);
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
            const int npre = nkb < q.stages ? nkb : q.stages;
            // weights do not depend on the previous kernel: fetch them before the grid dependency resolves
            for (int it = 0; it < npre; ++it) {
                const int kb = kb_begin + it;
                const int tap_i = kb / q.kb_per_tap;
                const int cb = kb - tap_i * q.kb_per_tap;
                mbar_expect_tx(&full_bar[it], a_bytes + b_bytes);
                tma_load_2d(&maps.b, &full_bar[it], smem_base + q.off_stage0 + (uint32_t)it * stage_bytes + a_bytes, cb * q.KC,
                            ph.taps[tap_i].slice * p.Cout + co0);
            }
            pdl_wait();
            pdl_launch_dependents();
            for (int it = 0; it < nkb; ++it) {
                const int s = it % q.stages;
                const uint32_t par = (uint32_t)((it / q.stages) & 1);
                const int kb = kb_begin + it;
                const int tap_i = kb / q.kb_per_tap;
                const int cb = kb - tap_i * q.kb_per_tap;
                const Tap t = ph.taps[tap_i];
                const uint32_t sa = smem_base + q.off_stage0 + (uint32_t)s * stage_bytes;
                if (it >= npre) {
                    mbar_wait(&empty_bar[s], par ^ 1u);
                    mbar_expect_tx(&full_bar[s], a_bytes + b_bytes);
                    tma_load_2d(&maps.b, &full_bar[s], sa + a_bytes, cb * q.KC, t.slice * p.Cout + co0);
                }
                int dy = t.dy, dx = t.dx;
                const CUtensorMap* mA = &maps.a[0];
                if (p.s_in == 2) {
                    const int py = dy & 1, px = dx & 1;        // parity view (works for negative offsets)
                    dy = (dy - py) / 2;
                    dx = (dx - px) / 2;
                    mA = &maps.a[py * 2 + px];
                }
                tma_load_4d(mA, &full_bar[s], sa, cb * q.KC, w0 + dx, h0 + dy, n0);
                if (it == 0) DBG_STAMP(2);
                if (it == 0 && f_add) {
                    // residual tile for the epilogue (with split-K any CTA can be the one that finalises)
                    mbar_expect_tx(&add_bar, (uint32_t)(TILE_M * q.BN * 2));
                    const int cpr = q.add_rowbytes >> 1;
                    for (int ch = 0; ch < q.add_chunks; ++ch)
                        tma_load_4d(&maps.add[phase_idx], &add_bar, smem_base + q.off_add + (uint32_t)(ch * TILE_M * q.add_rowbytes),
                                    co0 + ch * cpr, w0, h0, n0);
                }
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
                const uint32_t sa = smem_base + q.off_stage0 + (uint32_t)s * stage_bytes;
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
        const int BN = q.BN;

        // ---- work that does not need the accumulator: overlaps the main loop ----
        if (et < BN) sbias[et] = p.bias ? __ldg(p.bias + co0 + et) : 0.f;   // parameters: safe before the grid dependency
        float* sstat = sbias + BN;                // [2][BN] per-CTA column sums of the BatchNorm statistics
        if (f_stats) { for (int i = et; i < 2 * BN; i += 128) sstat[i] = 0.f; }
        pdl_wait();
        const size_t pix_off = row_ok ? ((((size_t)n_img * p.Ho + (size_t)(hg * p.s_out + ph.oy)) * p.Wo + (size_t)(wg * p.s_out + ph.ox)) *
                                         (size_t)p.Cout + (size_t)co0)
                                      : 0;
        float xc[16], anc[16];
        const float* xcur_row = f_ddim ? p.ddim.x_cur + pix_off : nullptr;
        const float* anc_row = f_addnc ? p.add_nc + (size_t)(row_ok ? n_img : 0) * p.Cout + co0 : nullptr;
        if (xcur_row && row_ok) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                const float4 t = *reinterpret_cast<const float4*>(xcur_row + j);
                xc[j] = t.x; xc[j + 1] = t.y; xc[j + 2] = t.z; xc[j + 3] = t.w;
            }
        }
        if (anc_row) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(anc_row + j));
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
            for (int c = 0; c < BN; c += 16) {
                float v[16];
                tmem_ld16(taddr + (uint32_t)c, v);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<float4*>(part + (size_t)row * BN * 4 + (size_t)(((c >> 2) + j) ^ (row & pm)) * 16) =
                        make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            }
            fence_async_smem();
            epi_bar();
            if (et == 0) DBG_STAMP(8);
            const size_t tile_lin = ((size_t)phase_idx * gridDim.x + mt) * gridDim.y + blockIdx.y;
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
                for (int i = et; i < TILE_M * BN / 4; i += 128) __stcg(z + i, make_float4(0.f, 0.f, 0.f, 0.f));
                if (et == 0) DBG_STAMP(10);
            }
        }

        if (finalize) {
            if (f_add) mbar_wait(&add_bar, 0);
            const int out_cpr = q.out_rowbytes / q.out_esz;
            const int add_cpr = q.add_rowbytes >> 1;
            for (int c = 0; c < BN; c += 16) {
                float v[16];
                if (f_split) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 t = *reinterpret_cast<const float4*>(part + (size_t)row * BN * 4 + (size_t)(((c >> 2) + j) ^ (row & pm)) * 16);
                        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
                    }
                } else {
                    tmem_ld16(taddr + (uint32_t)c, v);
                }
                if (c == 0 && et == 0) DBG_STAMP(13);
                // prefetch the next chunk's per-row operands
                float xn_[16], an_[16];
                const bool more = c + 16 < BN;
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
                        x = apply_act(x, p.act);
                        if (p.affine) x = x * p.out_scale + p.out_shift;
                    } else if (p.act == MSTK_ACT_RELU) {
                        x = fmaxf(x, 0.f);
                    }
                    v[j] = x;
                }
                if (c == 0 && et == 0) DBG_STAMP(14);
                if (anc_row) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += anc[j];
                }
                if (f_add) {
                    const uint8_t* ab = smem + q.off_add + (size_t)(c / add_cpr) * TILE_M * q.add_rowbytes;
                    const int j0 = ((c % add_cpr) * 2) >> 4;
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const uint4 t = *reinterpret_cast<const uint4*>(ab + swz_off(row, j0 + h2, q.add_rowbytes));
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
                if (f_store) {
                    uint8_t* ob = smem + q.off_out + (size_t)(c / out_cpr) * TILE_M * q.out_rowbytes;
                    const int j0 = ((c % out_cpr) * q.out_esz) >> 4;
                    if (q.out_esz == 4) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<float4*>(ob + swz_off(row, j0 + j, q.out_rowbytes)) =
                                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    } else {
                        uint4 pk[2];
                        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
                        for (int e = 0; e < 8; ++e) hh[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
                        *reinterpret_cast<uint4*>(ob + swz_off(row, j0, q.out_rowbytes)) = pk[0];
                        *reinterpret_cast<uint4*>(ob + swz_off(row, j0 + 1, q.out_rowbytes)) = pk[1];
                    }
                }
                if (c == 0 && et == 0) DBG_STAMP(15);
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
            if (f_store) {
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
typedef void (*TcKernel)(const TcMaps, const TcParams);
struct TcVariant { uint32_t flags; TcKernel fn; };
#define TCV(f) {(f), conv_tc_kernel<(f)>}
static const TcVariant kVariants[] = {
    TCV(F_STORE), TCV(F_STORE | F_SPLITK),                                   // bias (+ReLU): most layers
    TCV(F_STORE | F_ADDNC), TCV(F_STORE | F_ADDNC | F_SPLITK),               // + time embedding (UNet enc2)
    TCV(F_STORE | F_ADD), TCV(F_STORE | F_ADD | F_SPLITK),                   // + skip connection (UNet dec2..4)
    TCV(F_DDIM), TCV(F_DDIM | F_SPLITK), TCV(F_DDIM | F_STORE),              // fused DDIM update (UNet dec1)
    TCV(F_STORE | F_STATS), TCV(F_STORE | F_STATS | F_SPLITK),               // BatchNorm statistics (VAE, train mode)
    TCV(F_STORE | F_ACTGEN),                                                 // tanh / affine epilogues
    TCV(F_ALL),                                                              // anything else
};
static const int kNumVariants = (int)(sizeof(kVariants) / sizeof(kVariants[0]));

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static int g_tc_state = 0;   // 0 unknown, 1 ok, -1 failed
static int g_num_sms = 148;
static int g_pdl = 1;
static int g_force_bn = 0, g_force_sk = 0, g_force_stages = 0;
static long long* g_dbg = nullptr;
void tc_set_debug(long long* p) { g_dbg = p; }
void tc_set_force(int bn, int sk, int stages, int pdl) {
    g_force_bn = bn; g_force_sk = sk; g_force_stages = stages;
    if (pdl >= 0) g_pdl = pdl;
}
static int g_last_cfg[8] = {0, 0, 0, 0, 0, 0, 0, 0};
const int* tc_last_config() { return g_last_cfg; }

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
    // development knobs (tuning sweeps): MSTK_PDL=0 disables programmatic dependent launch,
    // MSTK_TC_FORCE="BN,splitk,stages" overrides the tile heuristic (0 = keep heuristic value)
    if (const char* s = getenv("MSTK_PDL")) g_pdl = atoi(s) != 0;
    if (const char* s = getenv("MSTK_TC_FORCE")) sscanf(s, "%d,%d,%d", &g_force_bn, &g_force_sk, &g_force_stages);
    for (int i = 0; i < kNumVariants; ++i) {
        cudaError_t ea = cudaFuncSetAttribute(kVariants[i].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (ea != cudaSuccess) {
            cuda_fail(ea, "cudaFuncSetAttribute(conv_tc_kernel)");
            g_tc_state = -1;
            return false;
        }
    }
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
    if ((reinterpret_cast<uintptr_t>(p.y) & 15) || (reinterpret_cast<uintptr_t>(p.add) & 15)) return false;
    if (p.has_ddim && ((reinterpret_cast<uintptr_t>(p.ddim.x_cur) & 15) || (reinterpret_cast<uintptr_t>(p.ddim.x_next) & 15) ||
                       (reinterpret_cast<uintptr_t>(p.ddim.x_next_lp) & 15) || (reinterpret_cast<uintptr_t>(p.ddim.x0_pred) & 15)))
        return false;
    if (p.add_nc && (reinterpret_cast<uintptr_t>(p.add_nc) & 15)) return false;
    return true;
}

struct TcConfig {
    int Wt, Ht, Nt, tiles_w, tiles_h, tiles_n, KC, BN, kb_per_tap, splitk, stages, tmem_cols;
    int out_esz, out_rowbytes, out_chunks, add_rowbytes, add_chunks;
    uint32_t off_part, off_out, off_add, off_bias;
    size_t smem_bytes;
};

// shared-memory carve-up for a (BN, splitk, stages) choice; fills the offsets and smem_bytes
static void layout_smem(const ConvP& p, TcConfig& c) {
    const size_t a_bytes = (size_t)TILE_M * c.KC * 2;
    const size_t b_bytes = (((size_t)c.BN * c.KC * 2) + 1023) & ~(size_t)1023;
    const size_t pipe = (a_bytes + b_bytes) * c.stages;
    c.out_esz = p.out_dtype == MSTK_F32 ? 4 : 2;
    c.out_rowbytes = c.BN * c.out_esz >= 128 ? 128 : c.BN * c.out_esz;
    c.out_chunks = c.BN * c.out_esz / c.out_rowbytes;
    c.add_rowbytes = c.BN * 2 >= 128 ? 128 : c.BN * 2;
    c.add_chunks = c.BN * 2 / c.add_rowbytes;
    const size_t part = c.splitk > 1 ? (size_t)TILE_M * c.BN * 4 : 0;
    const size_t outb = p.y ? (size_t)TILE_M * c.BN * c.out_esz : 0;
    const size_t addb = p.add ? (size_t)TILE_M * c.BN * 2 : 0;
    c.off_part = 0;
    c.off_out = (uint32_t)part;                       // part and out tiles re-use the pipeline stages (main loop is over)
    size_t main_bytes = pipe > part + outb ? pipe : part + outb;
    main_bytes = (main_bytes + 1023) & ~(size_t)1023;
    c.off_add = (uint32_t)main_bytes;
    c.off_bias = (uint32_t)(main_bytes + addb);
    c.smem_bytes = main_bytes + addb + (size_t)c.BN * 4 * 3 + 1024;   // bias + 2 x BN stat sums + alignment slack
}

static TcConfig choose_config(const ConvP& p) {
    TcConfig c;
    memset(&c, 0, sizeof(c));
    c.Wt = pow2_divisor(p.Wg, TILE_M);
    c.Ht = pow2_divisor(p.Hg, TILE_M / c.Wt);
    c.Nt = TILE_M / (c.Wt * c.Ht);
    c.tiles_w = p.Wg / c.Wt;
    c.tiles_h = p.Hg / c.Ht;
    c.tiles_n = (p.N + c.Nt - 1) / c.Nt;
    c.KC = (p.Cin % 64 == 0) ? 64 : 32;
    c.kb_per_tap = p.Cin / c.KC;
    const long long mtiles = (long long)c.tiles_w * c.tiles_h * c.tiles_n * p.nphases;
    int min_kb = 1 << 30, max_kb = 0;
    for (int i = 0; i < p.nphases; ++i) {
        const int kb = p.ph[i].ntaps * c.kb_per_tap;
        if (kb < min_kb) min_kb = kb;
        if (kb > max_kb) max_kb = kb;
    }
    // cost model (cycles) over (BN, splitk): L2-fed K loop vs MMA floor, wave quantisation, epilogue
    double best = 1e30;
    int best_bn = 16, best_sk = 1, best_st = 2;
    for (int bn = 128; bn >= 16; bn >>= 1) {
        if (p.Cout % bn != 0) continue;
        if (g_force_bn && bn != g_force_bn && p.Cout % g_force_bn == 0) continue;
        for (int sk = 1; sk <= 16; sk <<= 1) {
            if (sk > min_kb) break;
            if (g_force_sk && sk != g_force_sk && g_force_sk <= min_kb) continue;
            if (sk > 1 && bn > 64) continue;          // keep the fp32 partial tile <= 32 KB
            TcConfig t = c;
            t.BN = bn; t.splitk = sk;
            const int kb_cta = (max_kb + sk - 1) / sk;
            int st = kb_cta < MAX_STAGES ? kb_cta : MAX_STAGES;
            if (st < 2) st = 2;
            t.stages = st;
            layout_smem(p, t);
            while (t.stages > 3 && t.smem_bytes > 100 * 1024) { --t.stages; layout_smem(p, t); }
            if (t.smem_bytes > 220 * 1024) continue;
            int occ = (int)((228 * 1024) / (t.smem_bytes + 1024));
            if (occ > 4) occ = 4;
            if (occ < 1) occ = 1;
            const long long ctas = mtiles * (p.Cout / bn) * sk;
            const long long slots = (long long)g_num_sms * occ;
            const long long waves = (ctas + slots - 1) / slots;
            const double conc = (double)(ctas < slots ? ctas : slots);
            const double per_sm = conc / g_num_sms < 1.0 ? 1.0 : conc / g_num_sms;
            double bw = 90.0 / per_sm;                               // bytes/cycle one CTA can pull through its SM
            if (bw > 6000.0 / conc) bw = 6000.0 / conc;              // L2 cap shared by everyone streaming at once
            const double stage_b = (double)TILE_M * c.KC * 2 + (double)bn * c.KC * 2;
            const double t_load = kb_cta * stage_b / bw;
            const double t_mma = (double)kb_cta * (c.KC / 16) * (bn / 2.0) * per_sm;
            const double t_k = (t_load > t_mma ? t_load : t_mma) + 900.0;
            const double t_epi = 500.0 + bn * 14.0 + (sk > 1 ? 1800.0 + bn * 10.0 : 0.0);
            const double cost = waves * (t_k + t_epi) + 1200.0;
            if (cost < best) { best = cost; best_bn = bn; best_sk = sk; best_st = t.stages; }
        }
    }
    c.BN = best_bn; c.splitk = best_sk; c.stages = g_force_stages ? g_force_stages : best_st;
    if (c.stages > MAX_STAGES) c.stages = MAX_STAGES;
    if (c.stages < 2) c.stages = 2;
    layout_smem(p, c);
    int cols = 32;
    while (cols < c.BN) cols <<= 1;
    c.tmem_cols = cols;
    return c;
}

static void ws_sizes(const ConvP& p, const TcConfig& c, size_t* ws, size_t* ctr) {
    const size_t tiles = (size_t)c.tiles_w * c.tiles_h * c.tiles_n * p.nphases * (size_t)(p.Cout / c.BN);
    *ws = tiles * TILE_M * (size_t)c.BN * sizeof(float);
    *ctr = (tiles * sizeof(unsigned int) + 255) & ~(size_t)255;
}

size_t conv_tc_workspace_bytes(const ConvP& p) {
    if (!tc_init()) return 0;
    const TcConfig c = choose_config(p);
    if (c.splitk == 1) return 0;
    size_t ws, ctr;
    ws_sizes(p, c, &ws, &ctr);
    return ws + ctr;
}

static int encode_map(CUtensorMap* m, CUtensorMapDataType dt, const void* base, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes, const cuuint32_t* box, CUtensorMapSwizzle sw) {
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = g_encode(m, dt, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]", (int)r, rank,
                  (unsigned long long)dims[0], (unsigned long long)dims[1], rank > 2 ? (unsigned long long)dims[2] : 0ull,
                  rank > 3 ? (unsigned long long)dims[3] : 0ull, box[0], box[1], rank > 2 ? box[2] : 0u, rank > 3 ? box[3] : 0u);
        return MSTK_ECUDA;
    }
    return MSTK_OK;
}

// exported to wgrad_tc.cu: encode a bf16/f32 tiled map with the swizzle that matches `rowbytes`
int tc_encode_map(CUtensorMap* m, int is_f32, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box, int rowbytes) {
    cuuint64_t d[5], s[5];
    cuuint32_t b[5];
    for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; }
    for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
    const CUtensorMapSwizzle sw = rowbytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : rowbytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
    return encode_map(m, is_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, d, s, b, sw);
}
int tc_num_sms() { return g_num_sms; }

static CUtensorMapSwizzle swizzle_for(int rowbytes) {
    return rowbytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : rowbytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
}

// tensor maps of an output-shaped NHWC tensor (y or the residual): one per phase
static int encode_out_maps(const ConvP& p, const TcConfig& c, const void* base, int esz, int rowbytes, CUtensorMap* maps) {
    const CUtensorMapDataType dt = esz == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    for (int i = 0; i < p.nphases; ++i) {
        const Phase& ph = p.ph[i];
        const uint8_t* b = reinterpret_cast<const uint8_t*>(base) + ((size_t)ph.oy * p.Wo + ph.ox) * p.Cout * esz;
        const cuuint64_t s = (cuuint64_t)p.s_out;
        cuuint64_t dims[4] = {(cuuint64_t)p.Cout, (cuuint64_t)p.Wg, (cuuint64_t)p.Hg, (cuuint64_t)p.N};
        cuuint64_t strides[3] = {s * p.Cout * esz, s * p.Wo * p.Cout * esz, (cuuint64_t)p.Ho * p.Wo * p.Cout * esz};
        cuuint32_t box[4] = {(cuuint32_t)(rowbytes / esz), (cuuint32_t)c.Wt, (cuuint32_t)c.Ht, (cuuint32_t)c.Nt};
        int rc = encode_map(&maps[i], dt, b, 4, dims, strides, box, swizzle_for(rowbytes));
        if (rc != MSTK_OK) return rc;
    }
    for (int i = p.nphases; i < 4; ++i) maps[i] = maps[0];
    return MSTK_OK;
}

int launch_conv_tc(const ConvP& p, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (!tc_init()) return MSTK_EUNSUPPORTED;
    TcConfig c = choose_config(p);
    TcParams q;
    memset(&q, 0, sizeof(q));
    q.p = p;
    if (c.splitk > 1) {
        size_t ws, ctr;
        ws_sizes(p, c, &ws, &ctr);
        MSTK_REQUIRE(workspace != nullptr && workspace_bytes >= ws + ctr,
                     "conv(tcgen05): split-K needs %zu bytes of zero-initialised workspace (mstk_conv2d_workspace_bytes), got %zu",
                     ws + ctr, workspace_bytes);
        MSTK_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "conv(tcgen05): workspace must be 16-byte aligned");
        q.ws = reinterpret_cast<float*>(workspace);
        q.counters = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(workspace) + ws);
    }
    q.Wt = c.Wt; q.Ht = c.Ht; q.Nt = c.Nt;
    q.tiles_w = c.tiles_w; q.tiles_h = c.tiles_h; q.tiles_n = c.tiles_n;
    q.dbg = g_dbg;
    q.KC = c.KC; q.BN = c.BN; q.kb_per_tap = c.kb_per_tap; q.splitk = c.splitk; q.stages = c.stages; q.tmem_cols = c.tmem_cols;
    q.out_esz = c.out_esz; q.out_rowbytes = c.out_rowbytes; q.out_chunks = c.out_chunks;
    q.add_rowbytes = c.add_rowbytes; q.add_chunks = c.add_chunks;
    q.use_tma_store = p.y ? 1 : 0;
    q.use_tma_add = p.add ? 1 : 0;
    q.off_stage0 = 0; q.off_part = c.off_part; q.off_out = c.off_out; q.off_add = c.off_add; q.off_bias = c.off_bias;

    const CUtensorMapSwizzle sw = (c.KC == 64) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    TcMaps maps;
    memset(&maps, 0, sizeof(maps));
    const int s = p.s_in;
    const int nviews = (s == 2) ? 4 : 1;
    for (int v = 0; v < nviews; ++v) {
        const int py = v >> 1, px = v & 1;
        const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(p.x) + ((size_t)py * p.W + px) * p.Cin;
        cuuint64_t dims[4] = {(cuuint64_t)p.Cin, (cuuint64_t)(p.W / s), (cuuint64_t)(p.H / s), (cuuint64_t)p.N};
        cuuint64_t strides[3] = {(cuuint64_t)p.Cin * 2 * s, (cuuint64_t)p.W * p.Cin * 2 * s, (cuuint64_t)p.H * p.W * p.Cin * 2};
        cuuint32_t box[4] = {(cuuint32_t)c.KC, (cuuint32_t)c.Wt, (cuuint32_t)c.Ht, (cuuint32_t)c.Nt};
        int rc = encode_map(&maps.a[v], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, 4, dims, strides, box, sw);
        if (rc != MSTK_OK) return rc;
    }
    for (int v = nviews; v < 4; ++v) maps.a[v] = maps.a[0];
    {
        int nslices = 0;
        for (int i = 0; i < p.nphases; ++i)
            for (int t = 0; t < p.ph[i].ntaps; ++t)
                if (p.ph[i].taps[t].slice + 1 > nslices) nslices = p.ph[i].taps[t].slice + 1;
        cuuint64_t dims[2] = {(cuuint64_t)p.Cin, (cuuint64_t)nslices * p.Cout};
        cuuint64_t strides[1] = {(cuuint64_t)p.Cin * 2};
        cuuint32_t box[2] = {(cuuint32_t)c.KC, (cuuint32_t)c.BN};
        int rc = encode_map(&maps.b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, p.w, 2, dims, strides, box, sw);
        if (rc != MSTK_OK) return rc;
    }
    if (p.y) {
        int rc = encode_out_maps(p, c, p.y, c.out_esz, c.out_rowbytes, maps.out);
        if (rc != MSTK_OK) return rc;
    } else {
        for (int i = 0; i < 4; ++i) maps.out[i] = maps.a[0];
    }
    if (p.add) {
        int rc = encode_out_maps(p, c, p.add, 2, c.add_rowbytes, maps.add);
        if (rc != MSTK_OK) return rc;
    } else {
        for (int i = 0; i < 4; ++i) maps.add[i] = maps.a[0];
    }

    g_last_cfg[0] = c.BN; g_last_cfg[1] = c.splitk; g_last_cfg[2] = c.stages; g_last_cfg[3] = (int)c.smem_bytes;
    g_last_cfg[4] = c.tiles_w * c.tiles_h * c.tiles_n; g_last_cfg[5] = p.Cout / c.BN; g_last_cfg[6] = p.nphases * c.splitk; g_last_cfg[7] = c.KC;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(c.tiles_w * c.tiles_h * c.tiles_n), (unsigned)(p.Cout / c.BN), (unsigned)(p.nphases * c.splitk));
    cfg.blockDim = dim3(TC_THREADS, 1, 1);
    cfg.dynamicSmemBytes = c.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_pdl ? 1 : 0;
    uint32_t need = 0;
    if (c.splitk > 1) need |= F_SPLITK;
    if (p.stats) need |= F_STATS;
    if (p.add_nc) need |= F_ADDNC;
    if (p.add) need |= F_ADD;
    if (p.has_ddim) need |= F_DDIM;
    if (p.y) need |= F_STORE;
    if (p.affine || (p.act != MSTK_ACT_NONE && p.act != MSTK_ACT_RELU)) need |= F_ACTGEN;
    TcKernel fn = kVariants[kNumVariants - 1].fn;
    for (int i = 0; i < kNumVariants; ++i)
        if (kVariants[i].flags == need) { fn = kVariants[i].fn; break; }
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, maps, q);
    if (e != cudaSuccess) return cuda_fail(e, "conv_tc launch");
    return MSTK_OK;
}

}  // namespace mstk
