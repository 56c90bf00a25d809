// Host side of the convolution family: ABI entry, plan construction, dispatch.
#include <stdarg.h>
#include "common.cuh"

namespace mstk {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return MSTK_ECUDA;
}
const char* last_error() { return g_err; }

int build_conv_plan(const mstk_conv_args* a, ConvP* p) {
    memset(p, 0, sizeof(*p));
    MSTK_REQUIRE(a != nullptr, "conv: null args");
    MSTK_REQUIRE(a->x && a->wpacked, "conv: x / wpacked must be non-null");
    MSTK_REQUIRE(a->y || a->ddim, "conv: y is null and no ddim epilogue");
    MSTK_REQUIRE(a->n > 0 && a->h > 0 && a->w > 0 && a->cin > 0 && a->cout > 0, "conv: non-positive dims");
    MSTK_REQUIRE(a->dtype == MSTK_F32 || a->dtype == MSTK_BF16, "conv: bad dtype %d", a->dtype);
    MSTK_REQUIRE(a->out_dtype == MSTK_F32 || a->out_dtype == MSTK_BF16, "conv: bad out_dtype %d", a->out_dtype);
    MSTK_REQUIRE(a->act >= MSTK_ACT_NONE && a->act <= MSTK_ACT_TANH, "conv: bad act %d", a->act);
    p->x = a->x; p->w = a->wpacked; p->bias = a->bias; p->y = a->y; p->add = a->add; p->add_nc = a->add_nc;
    p->stats = a->stats;
    p->N = a->n; p->H = a->h; p->W = a->w; p->Cin = a->cin; p->Cout = a->cout;
    p->act = a->act;
    p->out_scale = a->out_scale; p->out_shift = a->out_shift;
    p->affine = (a->out_scale != 1.0f || a->out_shift != 0.0f) ? 1 : 0;
    p->dtype = a->dtype; p->out_dtype = a->out_dtype;
    if (a->ddim) {
        MSTK_REQUIRE(a->ddim->x_cur && a->ddim->x_next, "conv: ddim needs x_cur and x_next");
        p->has_ddim = 1;
        p->ddim.x_cur = a->ddim->x_cur; p->ddim.x_next = a->ddim->x_next;
        p->ddim.x_next_lp = reinterpret_cast<__nv_bfloat16*>(a->ddim->x_next_lp);
        p->ddim.x0_pred = a->ddim->x0_pred;
        p->ddim.sab_t = a->ddim->sab_t; p->ddim.s1m_t = a->ddim->s1m_t;
        p->ddim.sab_n = a->ddim->sab_n; p->ddim.s1m_n = a->ddim->s1m_n; p->ddim.eta = a->ddim->eta;
    }
    const int k = a->ksize, s = a->stride;
    if (!a->transposed) {
        MSTK_REQUIRE(k == 1 || k == 3 || k == 4, "conv: ksize %d unsupported (1, 3 or 4)", k);
        MSTK_REQUIRE(s == 1 || s == 2, "conv: stride %d unsupported", s);
        MSTK_REQUIRE(k != 1 || s == 1, "conv: k1 needs stride 1");
        MSTK_REQUIRE(k != 4 || s == 2, "conv: k4 needs stride 2 (data gradient of ConvTranspose2d k4 s2 p1)");
        const int pad = k == 4 ? 1 : k / 2;
        p->Ho = (a->h + 2 * pad - k) / s + 1;
        p->Wo = (a->w + 2 * pad - k) / s + 1;
        p->Hg = p->Ho; p->Wg = p->Wo;
        p->s_in = s; p->s_out = 1;
        p->nphases = 1;
        Phase& ph = p->ph[0];
        ph.oy = ph.ox = 0;
        ph.ntaps = 0;
        for (int ky = 0; ky < k; ++ky)
            for (int kx = 0; kx < k; ++kx) {
                Tap& t = ph.taps[ph.ntaps++];
                t.dy = (int16_t)(ky - pad); t.dx = (int16_t)(kx - pad); t.slice = ky * k + kx;
            }
    } else {
        MSTK_REQUIRE((k == 3 || k == 4) && s == 2, "convT: only k3 s2 p1 op1 and k4 s2 p1 are supported");
        // out = 2*in for both (k3: (h-1)*2 - 2 + 3 + 1 ; k4: (h-1)*2 - 2 + 4)
        p->Ho = 2 * a->h; p->Wo = 2 * a->w;
        p->Hg = a->h; p->Wg = a->w;
        p->s_in = 1; p->s_out = 2;
        p->nphases = 4;
        for (int py = 0; py < 2; ++py)
            for (int px = 0; px < 2; ++px) {
                Phase& ph = p->ph[py * 2 + px];
                ph.oy = py; ph.ox = px; ph.ntaps = 0;
                // oy = 2*iy - 1 + ky  =>  iy = hg + (py + 1 - ky)/2 when (py + 1 - ky) is even
                for (int ky = 0; ky < k; ++ky) {
                    if (((py + 1 - ky) & 1) != 0) continue;
                    for (int kx = 0; kx < k; ++kx) {
                        if (((px + 1 - kx) & 1) != 0) continue;
                        Tap& t = ph.taps[ph.ntaps++];
                        t.dy = (int16_t)((py + 1 - ky) / 2); t.dx = (int16_t)((px + 1 - kx) / 2);
                        t.slice = ky * k + kx;
                    }
                }
            }
    }
    return MSTK_OK;
}

}  // namespace mstk

extern "C" {

int mstk_version(void) { return MSTK_ABI_VERSION; }
const char* mstk_last_error(void) { return mstk::last_error(); }
int mstk_tcgen05_available(void) { return mstk::tc_init() ? 1 : 0; }
/* development aid (not in mstk.h): CTA (0,0,0) of every tcgen05 conv launch writes 8 clock64 stamps here */
int mstk_debug_tc_timeline(long long* dev_buf) { mstk::tc_set_debug(dev_buf); return MSTK_OK; }
/* development aids (not in mstk.h): override the tile heuristic (0 = heuristic; pdl < 0 = keep) and read back
 * {BN, splitk, stages, smem, grid.x, grid.y, grid.z, KC} of the most recent tcgen05 launch */
int mstk_debug_tc_force(int bn, int sk, int stages, int pdl) { mstk::tc_set_force(bn, sk, stages, pdl); return MSTK_OK; }
const int* mstk_debug_tc_last_config(void) { return mstk::tc_last_config(); }
/* 0: never take the persistent tile loop (conv_tcp_body.cuh); last_config()[1] == -1 marks a persistent launch */
int mstk_debug_tc_persist(int on) { mstk::tc_set_persist(on); return MSTK_OK; }

int mstk_conv2d_fwd(const mstk_conv_args* a, void* stream) {
    mstk::ConvP p;
    int rc = mstk::build_conv_plan(a, &p);
    if (rc != MSTK_OK) return rc;
    // every kernel behind this entry point moves its tensors with 128-bit accesses or TMA boxes
    MSTK_REQUIRE(mstk::aligned16({a->x, a->wpacked, a->y, a->add, a->add_nc, a->bias, a->stats}), "conv: tensor pointers must be 16-byte aligned");
    if (a->ddim)
        MSTK_REQUIRE(mstk::aligned16({a->ddim->x_cur, a->ddim->x_next, a->ddim->x_next_lp, a->ddim->x0_pred}),
                     "conv: the tensors of the ddim epilogue must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int backend = a->backend;
    if (backend == MSTK_BACKEND_AUTO) {
        if (mstk::conv_edge_supported(p)) return mstk::launch_conv_edge(p, st);
        backend = mstk::conv_tc_supported(p) ? MSTK_BACKEND_TCGEN05 : MSTK_BACKEND_SIMT;
    }
    if (backend == MSTK_BACKEND_TCGEN05) {
        if (!mstk::conv_tc_supported(p)) {
            mstk::set_error("conv: tcgen05 backend requested but shape unsupported (dtype=%d cin=%d cout=%d hg=%d wg=%d)",
                            p.dtype, p.Cin, p.Cout, p.Hg, p.Wg);
            return MSTK_EUNSUPPORTED;
        }
        if (mstk::tc_init() && mstk::conv_halo_supported(p)) return mstk::launch_conv_halo(p, st);
        return mstk::launch_conv_tc(p, a->workspace, a->workspace_bytes, st);
    }
    if (backend == MSTK_BACKEND_SIMT) return mstk::launch_conv_simt(p, st);
    mstk::set_error("conv: bad backend %d", a->backend);
    return MSTK_EINVAL;
}

size_t mstk_conv2d_workspace_bytes(const mstk_conv_args* a) {
    mstk::ConvP p;
    if (mstk::build_conv_plan(a, &p) != MSTK_OK) return 0;
    if (a->backend == MSTK_BACKEND_SIMT || !mstk::conv_tc_supported(p)) return 0;
    return mstk::conv_tc_workspace_bytes(p);
}

}  // extern "C"
