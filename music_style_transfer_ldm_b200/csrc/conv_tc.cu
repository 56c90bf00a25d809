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
// Warp roles (224 threads): warp 0 = A (activation) TMA producer, warp 6 = B (weight) TMA producer -- one issuing LANE per
// pipeline stage, because loads issued by a single thread do not overlap (tools/tma_probe.cu) --, warp 1 = TMEM allocator + MMA issuer
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
#include "conv_tc_body.cuh"
#include "conv_tcp_body.cuh"

namespace mstk {

// persistent form (conv_tcp_body.cuh): one CTA per SM looping over tiles, double-buffered TMEM accumulators
template <uint32_t F>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tcp_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcParams q, const __grid_constant__ TcpExtra x) {
    extern __shared__ uint8_t smem_raw[];
    conv_tcp_body<F>(maps, q, x, smem_raw);
}

template <uint32_t F>
__global__ void __launch_bounds__(TC_THREADS, (F & F_DDIM) ? 1 : TC_MIN_CTAS)
conv_tc_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcParams q) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) TcShared sh;
    conv_tc_body<F, false>(maps, q, (int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z, (int)gridDim.x, (int)gridDim.y, sh, smem_raw);
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

typedef void (*TcpKernel)(const TcMaps, const TcParams, const TcpExtra);
struct TcpVariant { uint32_t flags; TcpKernel fn; };
#define TCPV(f) {(f), conv_tcp_kernel<(f)>}
static const TcpVariant kTcpVariants[] = {TCPV(F_STORE), TCPV(F_STORE | F_STATS), TCPV(F_STORE | F_ACTGEN), TCPV(F_STORE | F_STATS | F_ACTGEN)};
static bool tcp_set_attrs() {
    for (const TcpVariant& v : kTcpVariants) {
        cudaError_t ea = cudaFuncSetAttribute(v.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (ea != cudaSuccess) {
            cuda_fail(ea, "cudaFuncSetAttribute(conv_tcp_kernel)");
            return false;
        }
    }
    return true;
}

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static int g_tc_state = 0;   // 0 unknown, 1 ok, -1 failed
static int g_num_sms = 148;
static int g_pdl = 1;
static int g_force_group = 0, g_trace = 0;
// Split-K partners as a thread-block cluster reducing through DSMEM (MSTK_TC_CLUSTER=1).  Measured on B200: the
// reduction itself is faster (6.6 k -> 4.4 k cycles per tile) but the sampling loop is SLOWER (10 607 -> 10 323
// steps/s): a cluster needs 4-8 free SMs of one GPC at once, which stalls the programmatic-dependent-launch overlap
// of consecutive layers.  Off by default.
static int g_cluster = 0;
static int g_tcp_grp = 0, g_tcp_stages = 0, g_tcp_bn = 0;   // MSTK_TCP_FORCE="kgroup,stages,BN" (tuning sweeps)
// Persistent tile loop (conv_tcp_body.cuh), OFF by default.  Measured on B200 (tools/tcp_tune.py, round 2): with one
// CTA per SM the smem ring (<= 160 KB in flight) and the single MMA issuer bound it at 640-1080 TFLOP/s on the VGGish
// stack against 690-1184 for two resident one-tile CTAs (two issuers, ~200 KB in flight), and the config-2 VAE layers do
// not move (92 us for the K = 512 convT either way: they are bound by the 16x re-read of the activation tile through
// L2, not by per-tile prologues).  MSTK_TC_PERSIST=1 / mstk_debug_tc_persist(1) select it; the tests keep it correct.
static int g_persist = 0;
static int g_direct = 0;      // MSTK_TC_DIRECT=1: register -> global output stores instead of staging + TMA store
static double g_split_cost = 1800.0;   // cost-model cycles charged to a split-K tile (MSTK_TC_SPLIT_COST)
static int g_force_bn = 0, g_force_sk = 0, g_force_stages = 0;
static long long* g_dbg = nullptr;
void tc_set_debug(long long* p) { g_dbg = p; }
void tc_set_force(int bn, int sk, int stages, int pdl) {
    g_force_bn = bn; g_force_sk = sk; g_force_stages = stages;
    if (pdl >= 0) g_pdl = pdl;
}
void tc_set_persist(int on) { g_persist = on; }
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
    if (const char* s = getenv("MSTK_TC_TRACE")) g_trace = atoi(s);
    if (const char* s = getenv("MSTK_TC_CLUSTER")) g_cluster = atoi(s);
    if (const char* s = getenv("MSTK_TC_DIRECT")) g_direct = atoi(s);
    if (const char* s = getenv("MSTK_TC_PERSIST")) g_persist = atoi(s);
    if (const char* s = getenv("MSTK_TCP_FORCE")) sscanf(s, "%d,%d,%d", &g_tcp_grp, &g_tcp_stages, &g_tcp_bn);
    if (const char* s = getenv("MSTK_TC_SPLIT_COST")) g_split_cost = atof(s);
    if (const char* s = getenv("MSTK_PDL")) g_pdl = atoi(s) != 0;
    if (const char* s = getenv("MSTK_TC_FORCE")) sscanf(s, "%d,%d,%d,%d", &g_force_bn, &g_force_sk, &g_force_stages, &g_force_group);
    for (int i = 0; i < kNumVariants; ++i) {
        cudaError_t ea = cudaFuncSetAttribute(kVariants[i].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (ea != cudaSuccess) {
            cuda_fail(ea, "cudaFuncSetAttribute(conv_tc_kernel)");
            g_tc_state = -1;
            return false;
        }
    }
    if (!tcp_set_attrs()) {
        g_tc_state = -1;
        return false;
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
    int Wt, Ht, Nt, tiles_w, tiles_h, tiles_n, KC, BN, kb_per_tap, splitk, stages, kgroup, tmem_cols;
    int out_esz, out_rowbytes, out_chunks, add_rowbytes, add_chunks;
    uint32_t off_part, off_out, off_add, off_bias;
    size_t smem_bytes;
};

// shared-memory carve-up for a (BN, splitk, stages) choice; fills the offsets and smem_bytes
static void layout_smem(const ConvP& p, TcConfig& c) {
    const size_t a_bytes = (size_t)TILE_M * c.KC * 2;
    const size_t b_bytes = (((size_t)c.BN * c.KC * 2) + 1023) & ~(size_t)1023;
    const size_t pipe = (a_bytes + b_bytes) * c.stages * c.kgroup;
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
    int best_bn = 16, best_sk = 1, best_st = 2, best_grp = 1;
    for (int bn = 128; bn >= 16; bn >>= 1) {
        if (p.Cout % bn != 0) continue;
        if (g_force_bn && bn != g_force_bn && p.Cout % g_force_bn == 0) continue;
        for (int sk = 1; sk <= 16; sk <<= 1) {
            if (sk > min_kb) break;
            if (g_force_sk && sk != g_force_sk && g_force_sk <= min_kb) continue;
            if (sk > 1 && bn > 64) continue;          // keep the fp32 partial tile <= 32 KB
            const int kb_cta = (max_kb + sk - 1) / sk;
            // k-blocks per stage: one issuer round (wait, MMAs, commit) costs ~600 cycles whatever it carries, so
            // short K loops want few, fat rounds; at most 16 MMAs per round and 30 producer lanes
            for (int grp = 1; grp <= 4; ++grp) {
                if (grp > 1 && (grp * (c.KC / 16) > 16 || grp > kb_cta)) break;
                if (g_force_group && grp != g_force_group && g_force_group <= kb_cta) continue;
                TcConfig t = c;
                t.BN = bn; t.splitk = sk; t.kgroup = grp;
                const int ngroups = (kb_cta + grp - 1) / grp;
                int st = ngroups < MAX_STAGES ? ngroups : MAX_STAGES;
                if (st < 2) st = 2;
                while (st * grp > 30) --st;
                t.stages = st;
                layout_smem(p, t);
                const int min_st = grp > 1 ? 2 : 3;
                const long long ctas = mtiles * (p.Cout / bn) * sk;
                // a grid of at most one CTA per SM may as well hold its whole K loop in flight
                const size_t budget = ctas <= g_num_sms ? 200 * 1024 : 100 * 1024;
                while (t.stages > min_st && t.smem_bytes > budget) { --t.stages; layout_smem(p, t); }
                if (t.smem_bytes > 220 * 1024) continue;
                int occ = (int)((228 * 1024) / (t.smem_bytes + 1024));
                if (occ > 4) occ = 4;
                if (occ < 1) occ = 1;
                const long long slots = (long long)g_num_sms * occ;
                const long long waves = (ctas + slots - 1) / slots;
                const double conc = (double)(ctas < slots ? ctas : slots);
                const double per_sm = conc / g_num_sms < 1.0 ? 1.0 : conc / g_num_sms;
                double bw = 90.0 / per_sm;                               // bytes/cycle one CTA can pull through its SM
                if (bw > 6000.0 / conc) bw = 6000.0 / conc;              // L2 cap shared by everyone streaming at once
                const double stage_b = (double)TILE_M * c.KC * 2 + (double)bn * c.KC * 2;
                const double t_load = kb_cta * stage_b / bw;
                const double t_mma = (double)kb_cta * (c.KC / 16) * (bn / 2.0) * per_sm;
                const double t_issue = ngroups * 600.0 + (double)kb_cta * (c.KC / 16) * 15.0;
                double t_k = t_load > t_mma ? t_load : t_mma;
                if (t_issue > t_k) t_k = t_issue;
                t_k += 600.0 + (t.stages * grp < kb_cta ? 0.0 : 0.0);
                const double t_epi = 500.0 + bn * 14.0 + (sk > 1 ? g_split_cost + bn * 10.0 : 0.0);
                const double cost = waves * (t_k + t_epi) + 1200.0;
                if (cost < best) { best = cost; best_bn = bn; best_sk = sk; best_st = t.stages; best_grp = grp; }
            }
        }
    }
    c.BN = best_bn; c.splitk = best_sk; c.kgroup = best_grp; c.stages = g_force_stages ? g_force_stages : best_st;
    if (c.stages > MAX_STAGES) c.stages = MAX_STAGES;
    if (c.stages < 2) c.stages = 2;
    while (c.stages * c.kgroup > 30) --c.stages;
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

// Persistent tile loop: throughput-bound grids only (several tiles per SM), plain bias / activation / BatchNorm-sum
// epilogues.  BN = the widest tile that divides Cout (<= 128: two accumulators fit 256 TMEM columns), the smem ring as
// deep as 200 KB allow next to the separate output staging tile.
static bool choose_config_persistent(const ConvP& p, TcConfig& c, TcpExtra& x) {
    if (!g_persist || !p.y || p.add || p.add_nc || p.has_ddim) return false;
    memset(&c, 0, sizeof(c));
    c.Wt = pow2_divisor(p.Wg, TILE_M);
    c.Ht = pow2_divisor(p.Hg, TILE_M / c.Wt);
    c.Nt = TILE_M / (c.Wt * c.Ht);
    c.tiles_w = p.Wg / c.Wt;
    c.tiles_h = p.Hg / c.Ht;
    c.tiles_n = (p.N + c.Nt - 1) / c.Nt;
    c.KC = (p.Cin % 64 == 0) ? 64 : 32;
    c.kb_per_tap = p.Cin / c.KC;
    c.BN = 0;
    for (int bn = 256; bn >= 16; bn >>= 1)
        if (p.Cout % bn == 0 && (!g_tcp_bn || bn <= g_tcp_bn)) { c.BN = bn; break; }
    if (!c.BN) return false;
    c.splitk = 1;
    const long long mtiles = (long long)c.tiles_w * c.tiles_h * c.tiles_n;
    const long long ntiles = mtiles * (p.Cout / c.BN) * p.nphases;
    if (ntiles < 3ll * g_num_sms || ntiles > 0x7fffffffll) return false;
    int min_kb = 1 << 30;
    for (int i = 0; i < p.nphases; ++i) {
        const int kb = p.ph[i].ntaps * c.kb_per_tap;
        if (kb < min_kb) min_kb = kb;
    }
    c.out_esz = p.out_dtype == MSTK_F32 ? 4 : 2;
    c.out_rowbytes = c.BN * c.out_esz >= 128 ? 128 : c.BN * c.out_esz;
    c.out_chunks = c.BN * c.out_esz / c.out_rowbytes;
    c.add_rowbytes = c.out_rowbytes; c.add_chunks = 0;
    const size_t a_bytes = (size_t)TILE_M * c.KC * 2;
    const size_t b_bytes = (((size_t)c.BN * c.KC * 2) + 1023) & ~(size_t)1023;
    const size_t sub = a_bytes + b_bytes;
    const size_t outb = ((size_t)TILE_M * (c.BN < 128 ? c.BN : 128) * c.out_esz + 1023) & ~(size_t)1023;   // one column block of <= 128
    const size_t budget = 200 * 1024 - outb - (size_t)c.BN * 12 - 1024;
    // k-blocks per ring stage: up to 8 MMAs behind one barrier round trip, but at least 3 stages in flight
    int grp = 8 / (c.KC / 16);
    if (grp > min_kb) grp = min_kb;
    while (grp > 1 && budget / (sub * grp) < 3) --grp;
    if (g_tcp_grp > 0 && g_tcp_grp <= min_kb) grp = g_tcp_grp;
    int stages = (int)(budget / (sub * grp));
    if (g_tcp_stages > 0 && g_tcp_stages <= stages) stages = g_tcp_stages;
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    while (stages * grp > 30) --stages;
    if (stages < 2) return false;
    c.kgroup = grp; c.stages = stages;
    int cols = 32;
    while (cols < 2 * c.BN) cols <<= 1;
    c.tmem_cols = cols;
    x.ntiles = (int)ntiles; x.mtiles = (int)mtiles; x.ny = p.Cout / c.BN;
    x.off_out = (uint32_t)(sub * grp * stages);
    x.off_bias = (uint32_t)(x.off_out + outb);
    c.smem_bytes = x.off_bias + (size_t)c.BN * 12 + 1024;
    return true;
}

int launch_conv_tc(const ConvP& p, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (!tc_init()) return MSTK_EUNSUPPORTED;
    TcConfig c;
    TcpExtra px;
    const bool persistent = choose_config_persistent(p, c, px);
    if (!persistent) c = choose_config(p);
    TcParams q;
    memset(&q, 0, sizeof(q));
    q.p = p;
    // opt-in: split-K partners as one thread-block cluster (portable size <= 8)
    q.cluster = (c.splitk > 1 && c.splitk <= 8 && g_cluster) ? 1 : 0;
    if (c.splitk > 1 && !q.cluster) {
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
    q.KC = c.KC; q.BN = c.BN; q.kb_per_tap = c.kb_per_tap; q.splitk = c.splitk; q.stages = c.stages; q.kgroup = c.kgroup; q.tmem_cols = c.tmem_cols;
    q.out_esz = c.out_esz; q.out_rowbytes = c.out_rowbytes; q.out_chunks = c.out_chunks;
    q.add_rowbytes = c.add_rowbytes; q.add_chunks = c.add_chunks;
    q.use_tma_store = p.y ? 1 : 0;
    q.direct_store = g_direct;
    q.use_tma_add = p.add ? 1 : 0;
    if (g_trace)
        fprintf(stderr, "[mstk conv_tc] N%d %dx%d Cin%d Cout%d k%d: BN %d KC %d splitk %d stages %d kgroup %d smem %zu\n", p.N, p.H, p.W, p.Cin,
                p.Cout, p.nphases, c.BN, c.KC, c.splitk, c.stages, c.kgroup, c.smem_bytes);
    {
        auto inv = [](int d) -> uint32_t { return d <= 1 ? 0u : (uint32_t)((0x100000000ull + (uint64_t)d - 1) / (uint64_t)d); };
        auto lg2 = [](int v) { int s = 0; while ((1 << s) < v) ++s; return s; };
        q.inv_tiles_w = inv(c.tiles_w); q.inv_tiles_h = inv(c.tiles_h); q.inv_tiles_wh = inv(c.tiles_w * c.tiles_h);
        q.inv_kb_per_tap = inv(c.kb_per_tap); q.inv_kgroup = inv(c.kgroup);
        q.wt_shift = lg2(c.Wt); q.ht_shift = lg2(c.Ht); q.splitk_shift = lg2(c.splitk);
        const long long mt_max = (long long)c.tiles_w * c.tiles_h * c.tiles_n;
        MSTK_REQUIRE(mt_max * c.tiles_w * c.tiles_h < (1ll << 32) && (1 << q.splitk_shift) == c.splitk && (1 << q.wt_shift) == c.Wt &&
                         (1 << q.ht_shift) == c.Ht,
                     "conv(tcgen05): tile grid outside the range of the division-free index math");
    }
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

    if (persistent) {
        uint32_t need = F_STORE;
        if (p.stats) need |= F_STATS;
        if (p.affine || (p.act != MSTK_ACT_NONE && p.act != MSTK_ACT_RELU)) need |= F_ACTGEN;
        TcpKernel pfn = nullptr;
        for (const TcpVariant& v : kTcpVariants)
            if (v.flags == need) pfn = v.fn;
        cudaLaunchConfig_t pc;
        memset(&pc, 0, sizeof(pc));
        pc.gridDim = dim3((unsigned)(px.ntiles < g_num_sms ? px.ntiles : g_num_sms), 1, 1);
        pc.blockDim = dim3(TC_THREADS, 1, 1);
        pc.dynamicSmemBytes = c.smem_bytes;
        pc.stream = st;
        cudaLaunchAttribute pattr[1];
        pattr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        pattr[0].val.programmaticStreamSerializationAllowed = 1;
        pc.attrs = pattr;
        pc.numAttrs = g_pdl ? 1 : 0;
        g_last_cfg[0] = c.BN; g_last_cfg[1] = -1; g_last_cfg[2] = c.stages; g_last_cfg[3] = (int)c.smem_bytes;
        g_last_cfg[4] = px.mtiles; g_last_cfg[5] = px.ny; g_last_cfg[6] = p.nphases; g_last_cfg[7] = c.KC;
        if (g_trace)
            fprintf(stderr, "[mstk conv_tcp] persistent: %d tiles on %u CTAs, BN %d KC %d stages %d kgroup %d smem %zu\n", px.ntiles, pc.gridDim.x,
                    c.BN, c.KC, c.stages, c.kgroup, c.smem_bytes);
        cudaError_t pe = cudaLaunchKernelEx(&pc, pfn, maps, q, px);
        if (pe != cudaSuccess) return cuda_fail(pe, "conv_tcp launch");
        return MSTK_OK;
    }
    g_last_cfg[0] = c.BN; g_last_cfg[1] = c.splitk; g_last_cfg[2] = c.stages; g_last_cfg[3] = (int)c.smem_bytes;
    g_last_cfg[4] = c.tiles_w * c.tiles_h * c.tiles_n; g_last_cfg[5] = p.Cout / c.BN; g_last_cfg[6] = p.nphases * c.splitk; g_last_cfg[7] = c.KC;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(c.tiles_w * c.tiles_h * c.tiles_n), (unsigned)(p.Cout / c.BN), (unsigned)(p.nphases * c.splitk));
    cfg.blockDim = dim3(TC_THREADS, 1, 1);
    cfg.dynamicSmemBytes = c.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int nattr = 0;
    if (g_pdl) {
        attr[nattr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[nattr].val.programmaticStreamSerializationAllowed = 1;
        ++nattr;
    }
    if (q.cluster) {
        attr[nattr].id = cudaLaunchAttributeClusterDimension;
        attr[nattr].val.clusterDim.x = 1; attr[nattr].val.clusterDim.y = 1; attr[nattr].val.clusterDim.z = (unsigned)c.splitk;
        ++nattr;
    }
    cfg.attrs = attr;
    cfg.numAttrs = nattr;
    uint32_t need = 0;
    if (c.splitk > 1) need |= F_SPLITK;
    if (p.stats) need |= F_STATS;
    if (p.add_nc) need |= F_ADDNC;
    if (p.add) need |= F_ADD;
    if (p.has_ddim) need |= F_DDIM;
    if (p.y) need |= F_STORE;
    if (p.affine || (p.act != MSTK_ACT_NONE && p.act != MSTK_ACT_RELU)) need |= F_ACTGEN;
    TcKernel fn = kVariants[kNumVariants - 1].fn;
    uint32_t picked = F_ALL;
    for (int i = 0; i < kNumVariants; ++i)
        if (kVariants[i].flags == need) { fn = kVariants[i].fn; picked = need; break; }
    (void)picked;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, maps, q);
    if (e != cudaSuccess) return cuda_fail(e, "conv_tc launch");
    return MSTK_OK;
}

}  // namespace mstk
