// =====================================================================================
// cvx_api.cu -- the C-ABI (include/curvix_b200.h): context, device memory, per-pass constants, launch order.
//
// Host work here is limited to what the reference does once per pass on a handful of 3x3 matrices
// (Set_Hypothesis_Views_Camera, EdgeSketch_Core.cpp:115-143, and the per-view F / epipoles that the
// reference recomputes inside every getOre* call).  Everything per edge / per hypothesis runs on the GPU.
// =====================================================================================
#include "../../include/curvix_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "cvx_internal.h"
#include "cvx_kernels.cuh"

using namespace cvx;

namespace {

// ----------------------------------------------------------------------------- host 3x3 helpers (same op order as the device code)
struct M3 { double m[9]; };
struct V3 { double v[3]; };
inline double hdot3(double a0, double a1, double a2, double b0, double b1, double b2) { return (a0 * b0 + a1 * b1) + a2 * b2; }
M3 h_mul(const M3& a, const M3& b) {
    M3 c;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) c.m[i * 3 + j] = hdot3(a.m[i * 3], a.m[i * 3 + 1], a.m[i * 3 + 2], b.m[j], b.m[3 + j], b.m[6 + j]);
    return c;
}
V3 h_mv(const M3& a, const V3& x) {
    V3 r;
    for (int i = 0; i < 3; i++) r.v[i] = hdot3(a.m[i * 3], a.m[i * 3 + 1], a.m[i * 3 + 2], x.v[0], x.v[1], x.v[2]);
    return r;
}
M3 h_tr(const M3& a) {
    M3 t;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) t.m[i * 3 + j] = a.m[j * 3 + i];
    return t;
}
// Eigen Matrix3d::inverse(): cofactors times 1/det
double h_cof(const M3& a, int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return a.m[i1 * 3 + j1] * a.m[i2 * 3 + j2] - a.m[i1 * 3 + j2] * a.m[i2 * 3 + j1];
}
M3 h_inv(const M3& a) {
    const double det = (h_cof(a, 0, 0) * a.m[0] + h_cof(a, 1, 0) * a.m[3]) + h_cof(a, 2, 0) * a.m[6];
    const double invdet = 1.0 / det;
    M3 r;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.m[j * 3 + i] = h_cof(a, i, j) * invdet;
    return r;
}
struct PairGeomH { M3 R21; V3 T21; M3 F; V3 e1pix, e2pix; };
// util.cpp:44-66 (relative pose, fundamental matrix) + getOrientationList.cpp:47-50 (epipoles)
PairGeomH h_pair_geometry(const M3& R1, const V3& T1, const M3& R2, const V3& T2, const M3& K1, const M3& K2) {
    PairGeomH g;
    g.R21 = h_mul(R2, h_tr(R1));
    M3 nR1t = h_tr(R1), nR2t = h_tr(R2);
    for (int i = 0; i < 9; i++) { nR1t.m[i] = -1 * nR1t.m[i]; nR2t.m[i] = -1 * nR2t.m[i]; }
    const V3 C1 = h_mv(nR1t, T1), C2 = h_mv(nR2t, T2);
    const V3 d = {{C1.v[0] - C2.v[0], C1.v[1] - C2.v[1], C1.v[2] - C2.v[2]}};
    g.T21 = h_mv(R2, d);
    const M3 Tx = {{0., -g.T21.v[2], g.T21.v[1], g.T21.v[2], 0., -g.T21.v[0], -g.T21.v[1], g.T21.v[0], 0.}};
    g.F = h_mul(h_mul(h_tr(h_inv(K2)), h_mul(Tx, g.R21)), h_inv(K1));
    const V3 rt = h_mv(h_tr(g.R21), g.T21);
    const V3 met1 = {{rt.v[0] / rt.v[2], rt.v[1] / rt.v[2], 1.0}};
    const V3 met2 = {{g.T21.v[0] / g.T21.v[2], g.T21.v[1] / g.T21.v[2], 1.0}};
    g.e1pix = h_mv(K1, met1);
    g.e2pix = h_mv(K2, met2);
    return g;
}

template <typename T>
struct DevBuf {   // grow-only device buffer
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 8 + 64;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace

struct cvx_ctx {
    cvx_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;             // side stream: the clustering launches of the larger size classes run beside the small one
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    std::string err;
    // cameras (host)
    int n_views = 0;
    std::vector<M3> R, K;
    std::vector<V3> T;
    bool multi_K = false, have_cams = false;
    // edge arena
    size_t arena_cap = 0, arena_used = 0;
    EdgeArena ea = {};
    std::vector<int64_t> edge_off;   // per view
    std::vector<int32_t> edge_n, edge_cap;
    DevBuf<int64_t> d_edge_off;
    DevBuf<int32_t> d_edge_n;
    bool tables_dirty = true;
    // pass
    int h1 = -1, h2 = -1, n_vali = 0;
    PassConst pc;
    std::vector<int32_t> vali_views;
    DevBuf<ViewConst> d_vc;
    DevBuf<int32_t> d_vali, d_cell_end, d_cell_end_b, d_scan_tmp;
    DevBuf<ViewGrid> d_vg;
    DevBuf<unsigned long long> d_minmax;
    DevBuf<double> d_ang2, d_h2_sorted_ang;
    DevBuf<int32_t> d_h2_sorted_idx, d_h2_bin_end, d_h2_bin_of;
    AngleIndex h2ix = {};
    DevBuf<double2> d_wedge_h1;
    DevBuf<int32_t> d_wflag_h1;
    // run
    bool have_run = false, have_final = false;
    int run_begin = 0, run_end = 0;
    DevBuf<int32_t> d_m, d_n, d_ncl, d_nvalid, d_nsel;
    DevBuf<int64_t> d_row_off, d_u_off, d_valid_off, d_pair_off;
    int64_t H = 0, U = 0, n_valid = 0, P = 0;
    // row table storage
    DevBuf<int32_t> r_h1, r_h2orig, r_local, r_rawat, r_label, r_h2rep, r_count, r_flags, r_urank, r_row_u;
    DevBuf<double> r_xc, r_yc, r_th, r_xbar, r_ybar, r_thbar, r_thn, r_orient;
    // unique-hypothesis table (one entry per cluster)
    DevBuf<int32_t> u_h1, u_row, u_mult, u_count;
    DevBuf<double2> u_xy;
    DevBuf<double4> u_geom;
    DevBuf<double2> d_w31, d_w32, d_tgt;
    bool keep_inter = false, have_inter = false;   // wedge32 / tangent of every (unique hypothesis, view): parity debugging only
    DevBuf<int32_t> d_support, d_support_t;
    DevBuf<DevCounters> d_ctr;
    DevBuf<int32_t> d_big_list;
    // run-based clustering of the > 128-row edges: work lists [long runs | short runs] and the scratch table of the long runs
    DevBuf<int4> d_run_desc;
    DevBuf<int32_t> d_run_ctr, vr_h2orig, vr_label, vr_h2rep, vr_urank, vr_oi, vr_n, vr_ncl, vr_order;
    DevBuf<double> vr_xc, vr_yc, vr_th, vr_xbar, vr_ybar, vr_thbar;
    DevBuf<int64_t> vr_off;
    DevBuf<double> d_ingest;            // cvx_set_all_view_edges: packed columns before the scatter into the arena slots
    DevBuf<int64_t> d_ingest_tab;
    std::vector<int64_t> h_ingest_tab;
    DevBuf<int32_t> v_h1, v_row, v_count, v_sel, p_h1, p_h2rep, p_row, p_count, p_sup, f_nan;
    DevBuf<double> f_gc, f_gw, f_tw, f_h2c;
    std::vector<int64_t> h_row_off;             // fetched lazily by cvx_get_h1_table
    bool h_row_off_valid = false;
    struct HostPin { int64_t H, U, tot[2]; int32_t cnts[8]; DevCounters hc; };
    HostPin* pin = nullptr;                     // pinned: the scalars a run reads back arrive by truly asynchronous copies
    bool timings_pending = false;
    // claimed-edge bitmap over the arena (one bit per edge slot) + per-view popcounts
    uint32_t* d_claim = nullptr;
    size_t claim_words = 0;
    DevBuf<unsigned long long> d_claim_cnt;
    DevBuf<uint8_t> d_claim_mask, d_keep;
    CvxCommState* comm = nullptr;               // multi-GPU gather state (cvx_comm.cu), NULL until cvx_comm_init / cvx_comm_attach
    cvx_counts counts;
    cvx_timings tm;
    cudaEvent_t ev[10];                         // run: 0,1,3,4,5; finalize: 6,7; set_pair: 8,9
    bool final_timing_pending = false;
    bool set_pair_timing_pending = false;
};

struct cvx_support_problem {
    cvx_ctx* ctx;
    int G, V;
    int64_t H, n_edges;
    DevBuf<int64_t> view_off;
    DevBuf<int32_t> cell_end, cell_end_b, scan_tmp, support, cell, idx, idx_b, row_group;
    DevBuf<double4> rec, rec_b;
    DevBuf<ViewGrid> vg;
    DevBuf<unsigned long long> minmax;
    GridIndex gi;
    DevBuf<double> a31, a32, cs, sn;
    DevBuf<double2> w31, w32, tgt;
    double thr, par;
    cudaEvent_t e0, e1;
};

namespace {

int fail(cvx_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    return code;
}
#define CU(call)                                                                                                  \
    do {                                                                                                          \
        cudaError_t _e = (call);                                                                                  \
        if (_e != cudaSuccess)                                                                                    \
            return fail(ctx, CVX_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e));                   \
    } while (0)

RowTable row_table(cvx_ctx* c) {
    RowTable rt;
    rt.h1 = c->r_h1.p; rt.h2orig = c->r_h2orig.p; rt.local = c->r_local.p; rt.rawat = c->r_rawat.p; rt.label = c->r_label.p;
    rt.h2rep = c->r_h2rep.p; rt.count = c->r_count.p; rt.flags = c->r_flags.p; rt.urank = c->r_urank.p; rt.row_u = c->r_row_u.p;
    rt.xc = c->r_xc.p; rt.yc = c->r_yc.p; rt.th = c->r_th.p; rt.xbar = c->r_xbar.p; rt.ybar = c->r_ybar.p; rt.thbar = c->r_thbar.p;
    rt.thn = c->r_thn.p; rt.orient = c->r_orient.p;
    return rt;
}

template <typename T>
cudaError_t regrow(T*& p, size_t keep, size_t cap, cudaStream_t st) {   // new allocation of cap elements, first `keep` copied over
    T* q = nullptr;
    cudaError_t e = cudaMalloc(&q, cap * sizeof(T));
    if (e != cudaSuccess) return e;
    if (p && keep) e = cudaMemcpyAsync(q, p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (p) cudaFree(p);
    p = q;
    return e;
}
void free_arena(EdgeArena& ea) {
    void* ps[] = {ea.x, ea.y, ea.th, ea.cs, ea.sn, ea.ang, ea.cell, ea.gi.A.rec, ea.gi.A.idx, ea.gi.B.rec, ea.gi.B.idx};
    for (void* p : ps) if (p) cudaFree(p);
    const GridIndex keep = ea.gi;   // cell_end / vg / minmax are owned by DevBufs
    ea = EdgeArena{};
    ea.gi.A.cell_end = keep.A.cell_end; ea.gi.B.cell_end = keep.B.cell_end; ea.gi.vg = keep.vg; ea.gi.minmax = keep.minmax;
    ea.gi.CS = keep.CS;
}
int grow_arena(cvx_ctx* ctx, size_t need) {
    if (need <= ctx->arena_cap) return CVX_OK;
    size_t cap = std::max(need, ctx->arena_cap * 2);
    cap = (cap + 1023) & ~size_t(1023);
    EdgeArena& ea = ctx->ea;
    const size_t used = ctx->arena_used;
    cudaStream_t st = ctx->stream;
    CU(regrow(ea.x, used, cap, st)); CU(regrow(ea.y, used, cap, st)); CU(regrow(ea.th, used, cap, st));
    CU(regrow(ea.cs, used, cap, st)); CU(regrow(ea.sn, used, cap, st));
    CU(regrow(ea.ang, 0, cap, st)); CU(regrow(ea.cell, 0, cap, st));
    CU(regrow(ea.gi.A.rec, 0, cap, st)); CU(regrow(ea.gi.A.idx, 0, cap, st)); CU(regrow(ea.gi.B.rec, 0, cap, st)); CU(regrow(ea.gi.B.idx, 0, cap, st));
    if (ctx->d_claim) {   // the claim bitmap covers the arena: grow it with the arena, new words cleared
        const size_t words = cap / 32;
        uint32_t* q = nullptr;
        CU(cudaMalloc(&q, words * 4));
        CU(cudaMemsetAsync(q, 0, words * 4, st));
        CU(cudaMemcpyAsync(q, ctx->d_claim, ctx->claim_words * 4, cudaMemcpyDeviceToDevice, st));
        CU(cudaStreamSynchronize(st));
        cudaFree(ctx->d_claim);
        ctx->d_claim = q; ctx->claim_words = words;
    }
    ctx->arena_cap = cap;
    return CVX_OK;
}

int sync_tables(cvx_ctx* ctx) {
    if (!ctx->tables_dirty) return CVX_OK;
    CU(ctx->d_edge_off.reserve(ctx->n_views + 1));
    CU(ctx->d_edge_n.reserve(ctx->n_views));
    CU(cudaMemcpyAsync(ctx->d_edge_off.p, ctx->edge_off.data(), ctx->n_views * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_edge_n.p, ctx->edge_n.data(), ctx->n_views * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->tables_dirty = false;
    return CVX_OK;
}

const M3& K_of(const cvx_ctx* c, int v) { return c->multi_K ? c->K[v] : c->K[0]; }


}  // namespace

extern "C" {

void cvx_default_config(cvx_config* c) {
    std::memset(c, 0, sizeof(*c));
    c->delta = 0.4; c->delta_theta_deg = 15; c->parallel_angle_deg = 15;
    c->lowes_ratio = 0.6; c->epip_tangency_displ_thresh = 3; c->location_perturbation = 0.4; c->orient_perturbation = 0.174533;
    c->cluster_dist_thresh = 1; c->cluster_orient_thresh_deg = 20.0; c->cluster_gauss_sigma = 2.0; c->pi_truncated = 3.1415926;
    c->n_views = 50; c->img_rows = 800; c->img_cols = 800; c->max_support_views = 4; c->max_cluster_size = 10;
}

int cvx_create(const cvx_config* cfg, int device, cvx_ctx** out) {
    if (!cfg || !out || cfg->n_views < 3) return CVX_ERR_INVALID;
    *out = nullptr;
    // With lowes_ratio >= 1 no row passes count > ratio * max and the reference falls back to the first element of its (unstable)
    // std::sort (EdgeSketch_Core.cpp:505-507); that order is not reproduced on the device, so the range is rejected.
    if (!(cfg->lowes_ratio >= 0.0 && cfg->lowes_ratio < 1.0) || cfg->max_support_views < 1) return CVX_ERR_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return CVX_ERR_NO_DEVICE;   // no CPU fallback by design
    if (device < 0 || device >= ndev) return CVX_ERR_INVALID;
    cvx_ctx* ctx = new cvx_ctx();
    ctx->cfg = *cfg; ctx->device = device; ctx->n_views = cfg->n_views;
    ctx->edge_off.assign(cfg->n_views, 0); ctx->edge_n.assign(cfg->n_views, 0); ctx->edge_cap.assign(cfg->n_views, 0);
    std::memset(&ctx->counts, 0, sizeof(ctx->counts));
    std::memset(&ctx->tm, 0, sizeof(ctx->tm));
    if (const char* b = getenv("CVX_KEEP_INTERMEDIATES")) ctx->keep_inter = atoi(b) != 0;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return CVX_ERR_CUDA;
    }
    for (auto& e : ctx->ev) cudaEventCreate(&e);
    if (cudaMallocHost(reinterpret_cast<void**>(&ctx->pin), sizeof(cvx_ctx::HostPin)) != cudaSuccess) { delete ctx; return CVX_ERR_NOMEM; }
    cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming); cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming);
    if (ctx->d_ctr.reserve(1) != cudaSuccess) { delete ctx; return CVX_ERR_NOMEM; }
    *out = ctx;
    return CVX_OK;
}

int cvx_destroy(cvx_ctx* ctx) {
    if (!ctx) return CVX_ERR_INVALID;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->comm) { cvx_comm_release(ctx->comm); ctx->comm = nullptr; }
    free_arena(ctx->ea);
    DevBuf<int32_t>* ib[] = {&ctx->d_edge_n, &ctx->d_vali, &ctx->d_cell_end, &ctx->d_cell_end_b, &ctx->d_scan_tmp, &ctx->d_wflag_h1, &ctx->d_m, &ctx->d_n, &ctx->d_ncl,
                             &ctx->d_nvalid, &ctx->d_nsel, &ctx->r_h1, &ctx->r_h2orig, &ctx->r_local, &ctx->r_rawat, &ctx->r_label, &ctx->r_h2rep,
                             &ctx->r_count, &ctx->r_flags, &ctx->r_urank, &ctx->r_row_u, &ctx->u_h1, &ctx->u_row, &ctx->u_mult, &ctx->u_count,
                             &ctx->d_support, &ctx->d_support_t, &ctx->v_h1, &ctx->v_row, &ctx->v_count, &ctx->v_sel, &ctx->p_h1, &ctx->p_h2rep,
                             &ctx->p_row, &ctx->p_count, &ctx->p_sup, &ctx->f_nan};
    for (auto* b : ib) b->release();
    DevBuf<int64_t>* lb[] = {&ctx->d_edge_off, &ctx->d_row_off, &ctx->d_u_off, &ctx->d_valid_off, &ctx->d_pair_off};
    for (auto* b : lb) b->release();
    ctx->d_h2_sorted_idx.release(); ctx->d_h2_bin_end.release(); ctx->d_h2_bin_of.release();
    DevBuf<double>* db[] = {&ctx->d_ang2, &ctx->d_h2_sorted_ang, &ctx->r_xc, &ctx->r_yc, &ctx->r_th, &ctx->r_xbar, &ctx->r_ybar, &ctx->r_thbar, &ctx->r_thn, &ctx->r_orient,
                            &ctx->f_gc, &ctx->f_gw, &ctx->f_tw, &ctx->f_h2c};
    for (auto* b : db) b->release();
    DevBuf<double2>* d2[] = {&ctx->d_wedge_h1, &ctx->u_xy, &ctx->d_w31, &ctx->d_w32, &ctx->d_tgt};
    for (auto* b : d2) b->release();
    ctx->u_geom.release(); ctx->d_vc.release(); ctx->d_ctr.release(); ctx->d_big_list.release(); ctx->d_ingest.release(); ctx->d_ingest_tab.release(); ctx->d_vg.release(); ctx->d_minmax.release();
    if (ctx->d_claim) cudaFree(ctx->d_claim);
    ctx->d_claim_cnt.release(); ctx->d_claim_mask.release(); ctx->d_keep.release(); ctx->d_run_desc.release(); ctx->d_run_ctr.release();
    { DevBuf<int32_t>* vb[] = {&ctx->vr_h2orig, &ctx->vr_label, &ctx->vr_h2rep, &ctx->vr_urank, &ctx->vr_oi, &ctx->vr_n, &ctx->vr_ncl, &ctx->vr_order};
      for (auto* b : vb) b->release();
      DevBuf<double>* vd[] = {&ctx->vr_xc, &ctx->vr_yc, &ctx->vr_th, &ctx->vr_xbar, &ctx->vr_ybar, &ctx->vr_thbar};
      for (auto* b : vd) b->release();
      ctx->vr_off.release(); }
    for (auto& e : ctx->ev) cudaEventDestroy(e);
    cudaStreamDestroy(ctx->stream);
    if (ctx->pin) cudaFreeHost(ctx->pin);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    delete ctx;
    return CVX_OK;
}

const char* cvx_last_error(const cvx_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int cvx_set_cameras(cvx_ctx* ctx, int n, const double* R, const double* T, const double* K, int k_per_view) {
    if (!ctx || !R || !T || !K || n != ctx->n_views) return fail(ctx, CVX_ERR_INVALID, "cvx_set_cameras: bad arguments");
    ctx->R.resize(n); ctx->T.resize(n);
    for (int i = 0; i < n; i++) { std::memcpy(ctx->R[i].m, R + 9 * i, 72); std::memcpy(ctx->T[i].v, T + 3 * i, 24); }
    ctx->multi_K = k_per_view != 0;
    ctx->K.resize(ctx->multi_K ? n : 1);
    for (size_t i = 0; i < ctx->K.size(); i++) std::memcpy(ctx->K[i].m, K + 9 * i, 72);
    ctx->have_cams = true;
    return CVX_OK;
}

int cvx_set_view_edges(cvx_ctx* ctx, int view, int n, const double* x, const double* y, const double* theta, const double* strength) {
    (void)strength;   // carried through by the host shim only; never used in any computation on the path
    if (!ctx || view < 0 || view >= ctx->n_views || n < 0 || (n > 0 && (!x || !y || !theta)))
        return fail(ctx, CVX_ERR_INVALID, "cvx_set_view_edges: bad arguments");
    CU(cudaSetDevice(ctx->device));
    if (n > ctx->edge_cap[view]) {   // new slot at the end of the arena (256-byte aligned)
        const size_t slot = (size_t(n) + 31) & ~size_t(31);
        const int rc = grow_arena(ctx, ctx->arena_used + slot);
        if (rc) return rc;
        ctx->edge_off[view] = int64_t(ctx->arena_used);
        ctx->edge_cap[view] = int32_t(slot);
        ctx->arena_used += slot;
    }
    ctx->edge_n[view] = n;
    ctx->tables_dirty = true;
    const int64_t off = ctx->edge_off[view];
    if (ctx->d_claim && ctx->edge_cap[view] > 0)   // new edges for this view: its claims no longer mean anything
        CU(cudaMemsetAsync(ctx->d_claim + off / 32, 0, size_t(ctx->edge_cap[view]) / 32 * 4, ctx->stream));
    if (n > 0) {
        CU(cudaMemcpyAsync(ctx->ea.x + off, x, size_t(n) * 8, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->ea.y + off, y, size_t(n) * 8, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->ea.th + off, theta, size_t(n) * 8, cudaMemcpyHostToDevice, ctx->stream));
        launch_edge_trig(ctx->ea.th + off, ctx->ea.cs + off, ctx->ea.sn + off, n, ctx->stream);
        ctx->tm.kernel_launches++;
    }
    ctx->have_run = false;
    return CVX_OK;
}

int cvx_set_all_view_edges(cvx_ctx* ctx, int n_views, const int32_t* n, const double* x, const double* y, const double* theta) {
    if (!ctx || n_views != ctx->n_views || !n) return fail(ctx, CVX_ERR_INVALID, "cvx_set_all_view_edges: bad arguments (n_views must equal cfg.n_views)");
    CU(cudaSetDevice(ctx->device));
    int64_t total = 0, max_n = 0;
    for (int v = 0; v < n_views; v++) {
        if (n[v] < 0) return fail(ctx, CVX_ERR_INVALID, "cvx_set_all_view_edges: negative edge count");
        total += n[v]; max_n = std::max<int64_t>(max_n, n[v]);
    }
    if (total > 0 && (!x || !y || !theta)) return fail(ctx, CVX_ERR_INVALID, "cvx_set_all_view_edges: null column");
    // arena slots first (growing moves the arena: do it before any pointer is taken)
    for (int v = 0; v < n_views; v++) {
        if (n[v] > ctx->edge_cap[v]) {
            const size_t slot = (size_t(n[v]) + 31) & ~size_t(31);
            const int rc = grow_arena(ctx, ctx->arena_used + slot);
            if (rc) return rc;
            ctx->edge_off[v] = int64_t(ctx->arena_used);
            ctx->edge_cap[v] = int32_t(slot);
            ctx->arena_used += slot;
        }
        ctx->edge_n[v] = n[v];
    }
    ctx->tables_dirty = true; ctx->have_run = false;
    cudaStream_t st = ctx->stream;
    if (ctx->d_claim) CU(cudaMemsetAsync(ctx->d_claim, 0, ctx->claim_words * 4, st));   // every view is new: no claim survives
    if (total == 0) return CVX_OK;
    CU(ctx->d_ingest.reserve(size_t(3) * size_t(total)));
    CU(ctx->d_ingest_tab.reserve(size_t(3) * n_views));
    std::vector<int64_t>& tab = ctx->h_ingest_tab;
    CU(cudaStreamSynchronize(st));            // tab may still be the source of the previous call's copy
    tab.assign(size_t(3) * n_views, 0);
    int64_t acc = 0;
    for (int v = 0; v < n_views; v++) { tab[v] = acc; tab[n_views + v] = ctx->edge_off[v]; tab[2 * n_views + v] = n[v]; acc += n[v]; }
    double* sx = ctx->d_ingest.p; double* sy = sx + total; double* sth = sy + total;
    CU(cudaMemcpyAsync(ctx->d_ingest_tab.p, tab.data(), tab.size() * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(sx, x, size_t(total) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(sy, y, size_t(total) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(sth, theta, size_t(total) * 8, cudaMemcpyHostToDevice, st));
    launch_ingest_views(sx, sy, sth, ctx->d_ingest_tab.p, n_views, max_n, ctx->ea, st);
    ctx->tm.kernel_launches++;
    CU(cudaGetLastError());
    return CVX_OK;
}

int cvx_set_hypothesis_pair(cvx_ctx* ctx, int h1, int h2) {
    if (!ctx || !ctx->have_cams || h1 < 0 || h2 < 0 || h1 >= ctx->n_views || h2 >= ctx->n_views || h1 == h2)
        return fail(ctx, CVX_ERR_INVALID, "cvx_set_hypothesis_pair: bad view indices or cameras not set");
    CU(cudaSetDevice(ctx->device));
    const cvx_config& cfg = ctx->cfg;
    ctx->h1 = h1; ctx->h2 = h2;
    // ---- Set_Hypothesis_Views_Camera, EdgeSketch_Core.cpp:115-143
    const M3 K1 = K_of(ctx, h1), K2 = K_of(ctx, h2);
    const PairGeomH g21 = h_pair_geometry(ctx->R[h1], ctx->T[h1], ctx->R[h2], ctx->T[h2], K1, K2);
    PassConst& pc = ctx->pc;
    std::memset(&pc, 0, sizeof(pc));
    std::memcpy(pc.F21, g21.F.m, 72); std::memcpy(pc.R21, g21.R21.m, 72); std::memcpy(pc.T21, g21.T21.v, 24);
    pc.e1_21[0] = g21.e1pix.v[0]; pc.e1_21[1] = g21.e1pix.v[1]; pc.e2_21[0] = g21.e2pix.v[0]; pc.e2_21[1] = g21.e2pix.v[1];
    const M3 K1i = h_inv(K1), K2i = h_inv(K2);
    std::memcpy(pc.K1inv, K1i.m, 72); std::memcpy(pc.K2inv, K2i.m, 72);
    std::memcpy(pc.R1, ctx->R[h1].m, 72); std::memcpy(pc.T1, ctx->T[h1].v, 24);
    pc.delta = cfg.delta;
    pc.cos_thresh = std::cos(double(cfg.delta_theta_deg) / 180 * cfg.pi_truncated);   // getSupportedEdgels.cpp:35 (host libm, as the reference)
    pc.parallel_deg = cfg.parallel_angle_deg;
    pc.loc_pert = cfg.location_perturbation; pc.tang_thresh = cfg.epip_tangency_displ_thresh; pc.orient_pert = cfg.orient_perturbation;
    pc.cl_dist = cfg.cluster_dist_thresh;
    {   // exact critical value of the squared distance (IEEE sqrt is correctly rounded and monotone; a few ulp steps from cl_dist^2)
        double sq = pc.cl_dist * pc.cl_dist;
        if (pc.cl_dist > 0) {
            while (std::sqrt(sq) >= pc.cl_dist) sq = std::nextafter(sq, 0.0);
            while (std::sqrt(sq) < pc.cl_dist) sq = std::nextafter(sq, INFINITY);
        } else sq = 0.0;   // sqrt(s) < non-positive threshold never holds for s >= 0 (s < 0 cannot occur)
        pc.cl_dist_sq = sq;
    }
    pc.cl_orient_rad = cfg.cluster_orient_thresh_deg * (M_PI / 180.0);                // util.hpp deg_to_rad
    pc.cl_wrap_deg = -90.0 + (cfg.cluster_orient_thresh_deg / 2.0);                   // EdgeClusterer.cpp:41
    pc.cl_sigma = cfg.cluster_gauss_sigma; pc.lowes_ratio = cfg.lowes_ratio;
    { int ex = 0; pc.cl_sigma_inv = (pc.cl_sigma > 0 && std::frexp(pc.cl_sigma, &ex) == 0.5) ? 1.0 / pc.cl_sigma : 0.0; }
    pc.rows = cfg.img_rows; pc.cols = cfg.img_cols; pc.max_support = cfg.max_support_views; pc.max_cluster = cfg.max_cluster_size;
    ctx->vali_views.clear();
    for (int v = 0; v < ctx->n_views; v++) if (v != h1 && v != h2) ctx->vali_views.push_back(v);
    const int V = int(ctx->vali_views.size());
    ctx->n_vali = V; pc.n_vali = V;
    std::vector<ViewConst> vc(V);
    int max_edges = 1;
    for (int i = 0; i < V; i++) {
        const int v = ctx->vali_views[i];
        const M3 K3 = K_of(ctx, v);
        const PairGeomH g31 = h_pair_geometry(ctx->R[h1], ctx->T[h1], ctx->R[v], ctx->T[v], K1, K3);
        const PairGeomH g32 = h_pair_geometry(ctx->R[h2], ctx->T[h2], ctx->R[v], ctx->T[v], K2, K3);
        ViewConst& c = vc[i];
        std::memset(&c, 0, sizeof(c));
        std::memcpy(c.F31, g31.F.m, 72); c.e1_31[0] = g31.e1pix.v[0]; c.e1_31[1] = g31.e1pix.v[1]; c.e2_31[0] = g31.e2pix.v[0]; c.e2_31[1] = g31.e2pix.v[1];
        std::memcpy(c.F32, g32.F.m, 72); c.e1_32[0] = g32.e1pix.v[0]; c.e1_32[1] = g32.e1pix.v[1]; c.e2_32[0] = g32.e2pix.v[0]; c.e2_32[1] = g32.e2pix.v[1];
        std::memcpy(c.R31, g31.R21.m, 72); std::memcpy(c.T31, g31.T21.v, 24);
        max_edges = std::max(max_edges, ctx->edge_n[v]);
    }
    int rc = sync_tables(ctx);
    if (rc) return rc;
    CU(ctx->d_vc.reserve(V)); CU(ctx->d_vali.reserve(V));
    cudaEventRecord(ctx->ev[8], ctx->stream);
    CU(cudaMemcpyAsync(ctx->d_vc.p, vc.data(), V * sizeof(ViewConst), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->d_vali.p, ctx->vali_views.data(), V * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    // ---- stage every validation view into its bin grid (hoisted getOreListVali + cos/sin, EdgeSketch_Core.cpp:315-342)
    GridIndex& gi = ctx->ea.gi;
    gi.CS = pick_cell_stride(max_edges);
    CU(ctx->d_scan_tmp.reserve(size_t(2 * std::max(V, 1)) * size_t(scan_tiles(gi.CS))));
    CU(ctx->d_cell_end.reserve(size_t(std::max(V, 1)) * size_t(gi.CS))); CU(ctx->d_cell_end_b.reserve(size_t(std::max(V, 1)) * size_t(gi.CS)));
    CU(ctx->d_vg.reserve(std::max(V, 1)));
    CU(ctx->d_minmax.reserve(size_t(std::max(V, 1)) * 4));
    gi.A.cell_end = ctx->d_cell_end.p; gi.B.cell_end = ctx->d_cell_end_b.p; gi.vg = ctx->d_vg.p; gi.minmax = ctx->d_minmax.p;
    GridSample gs;
    std::memset(&gs, 0, sizeof(gs));
    gs.x1 = ctx->ea.x + ctx->edge_off[h1]; gs.y1 = ctx->ea.y + ctx->edge_off[h1]; gs.E1 = ctx->edge_n[h1];
    gs.x2 = ctx->ea.x + ctx->edge_off[h2]; gs.y2 = ctx->ea.y + ctx->edge_off[h2]; gs.E2 = ctx->edge_n[h2];
    gs.delta = pc.delta; gs.rows = pc.rows; gs.cols = pc.cols;
    launch_stage_views(ctx->ea, ctx->d_edge_off.p, ctx->d_edge_n.p, ctx->d_vali.p, V, max_edges, ctx->d_vc.p, gs, ctx->d_scan_tmp.p, ctx->stream,
                       &ctx->tm.kernel_launches);
    // ---- Set_Hypothesis_Views_Edgels, EdgeSketch_Core.cpp:145-166
    const int E1 = ctx->edge_n[h1], E2 = ctx->edge_n[h2];
    CU(ctx->d_ang2.reserve(std::max(E2, 1))); CU(ctx->d_wedge_h1.reserve(std::max(E1, 1))); CU(ctx->d_wflag_h1.reserve(std::max(E1, 1)));
    launch_h2_angles(ctx->ea.x + ctx->edge_off[h2], ctx->ea.y + ctx->edge_off[h2], E2, pc.e2_21[0], pc.e2_21[1], ctx->d_ang2.p, ctx->stream);
    {   // angle-sorted index of view 2 for the candidate search
        AngleIndex& ix = ctx->h2ix;
        ix.B = pick_angle_bins(E2); ix.scale = ix.B / 180.0;
        CU(ctx->d_h2_sorted_ang.reserve(std::max(E2, 1))); CU(ctx->d_h2_sorted_idx.reserve(std::max(E2, 1)));
        CU(ctx->d_h2_bin_end.reserve(size_t(ix.B) + 1)); CU(ctx->d_h2_bin_of.reserve(std::max(E2, 1)));
        ix.ang = ctx->d_h2_sorted_ang.p; ix.idx = ctx->d_h2_sorted_idx.p; ix.bin_end = ctx->d_h2_bin_end.p;
        launch_h2_index(ctx->d_ang2.p, E2, ctx->d_h2_bin_of.p, ix, ctx->stream, &ctx->tm.kernel_launches);
    }
    launch_h1_wedges(pc, ctx->ea.x + ctx->edge_off[h1], ctx->ea.y + ctx->edge_off[h1], E1, ctx->d_wedge_h1.p, ctx->d_wflag_h1.p, ctx->stream);
    ctx->tm.kernel_launches += 2;
    cudaEventRecord(ctx->ev[9], ctx->stream);
    // No synchronisation: the copies above come from pageable host memory, which the runtime stages before cudaMemcpyAsync
    // returns, and the pass that follows is ordered on the same stream.  The phase time is read in cvx_get_timings.
    ctx->set_pair_timing_pending = true;
    CU(cudaGetLastError());
    ctx->have_run = false; ctx->have_final = false;
    return CVX_OK;
}

static UniqTable uniq_table(cvx_ctx* c) {
    UniqTable ut;
    ut.h1 = c->u_h1.p; ut.row = c->u_row.p; ut.mult = c->u_mult.p; ut.count = c->u_count.p; ut.xy = c->u_xy.p; ut.geom = c->u_geom.p;
    return ut;
}

int cvx_run_edge_sketch(cvx_ctx* ctx, int h1_begin, int h1_end) {
    if (!ctx || ctx->h1 < 0) return fail(ctx, CVX_ERR_INVALID, "cvx_run_edge_sketch: no hypothesis pair set");
    CU(cudaSetDevice(ctx->device));
    const int E1 = ctx->edge_n[ctx->h1], E2 = ctx->edge_n[ctx->h2], V = ctx->n_vali;
    if (h1_begin < 0) h1_begin = 0;
    if (h1_end < 0 || h1_end > E1) h1_end = E1;
    if (h1_begin > h1_end) h1_begin = h1_end;
    ctx->run_begin = h1_begin; ctx->run_end = h1_end;
    cudaStream_t st = ctx->stream;
    const PassConst& pc = ctx->pc;
    const double *x1 = ctx->ea.x + ctx->edge_off[ctx->h1], *y1 = ctx->ea.y + ctx->edge_off[ctx->h1], *th1 = ctx->ea.th + ctx->edge_off[ctx->h1];
    const double *x2 = ctx->ea.x + ctx->edge_off[ctx->h2], *y2 = ctx->ea.y + ctx->edge_off[ctx->h2], *th2 = ctx->ea.th + ctx->edge_off[ctx->h2];
    const int E1a = std::max(E1, 1), Va = std::max(V, 1);
    CU(ctx->d_m.reserve(E1a)); CU(ctx->d_n.reserve(E1a)); CU(ctx->d_ncl.reserve(E1a)); CU(ctx->d_nvalid.reserve(E1a)); CU(ctx->d_nsel.reserve(E1a));
    CU(ctx->d_row_off.reserve(E1a + 1)); CU(ctx->d_u_off.reserve(E1a + 1)); CU(ctx->d_valid_off.reserve(E1a + 1)); CU(ctx->d_pair_off.reserve(E1a + 1));
    CU(cudaMemsetAsync(ctx->d_m.p, 0, E1a * 4, st)); CU(cudaMemsetAsync(ctx->d_n.p, 0, E1a * 4, st)); CU(cudaMemsetAsync(ctx->d_ncl.p, 0, E1a * 4, st));
    CU(cudaMemsetAsync(ctx->d_nvalid.p, 0, E1a * 4, st)); CU(cudaMemsetAsync(ctx->d_nsel.p, 0, E1a * 4, st));
    CU(cudaMemsetAsync(ctx->d_ctr.p, 0, sizeof(DevCounters), st));
    CU(ctx->d_big_list.reserve(size_t(E1a) + 8));                     // clustering order of the view-1 edges + the class counts
    CU(cudaMemsetAsync(ctx->d_big_list.p + E1a, 0, 8 * 4, st));
    cudaEventRecord(ctx->ev[0], st);
    // ---- hypothesis formation
    launch_form_count(pc, x1, y1, h1_begin, h1_end, ctx->d_wedge_h1.p, x2, y2, th2, ctx->d_ang2.p, E2, ctx->h2ix, ctx->d_m.p, ctx->d_n.p, st);
    launch_scan_i32_i64(ctx->d_n.p, ctx->d_row_off.p, E1, st);
    ctx->tm.kernel_launches += 2;
    ctx->h_row_off_valid = false;
    CU(cudaMemcpyAsync(&ctx->pin->H, ctx->d_row_off.p + E1, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int64_t H = ctx->pin->H;
    ctx->H = H;
    const size_t Ha = size_t(std::max<int64_t>(H, 1));
    DevBuf<int32_t>* ib[] = {&ctx->r_h1, &ctx->r_h2orig, &ctx->r_local, &ctx->r_rawat, &ctx->r_label, &ctx->r_h2rep, &ctx->r_count, &ctx->r_flags,
                             &ctx->r_urank, &ctx->r_row_u};
    for (auto* b : ib) CU(b->reserve(Ha));
    DevBuf<double>* db[] = {&ctx->r_xc, &ctx->r_yc, &ctx->r_th, &ctx->r_xbar, &ctx->r_ybar, &ctx->r_thbar, &ctx->r_thn, &ctx->r_orient};
    for (auto* b : db) CU(b->reserve(Ha));
    RowTable rt = row_table(ctx);
    // scratch of the run-based clustering, sized for the worst case: H / 2 short runs, H / 33 long ones holding at most H rows
    RunScratch rs;
    std::memset(&rs, 0, sizeof(rs));
    const bool any_big = H > 0;
    if (any_big) {
        const int64_t cap_short = H / 2 + 1, cap_long = H / 33 + 1;
        CU(ctx->d_run_desc.reserve(size_t(cap_short + cap_long))); CU(ctx->d_run_ctr.reserve(4));
        DevBuf<double>* vd[] = {&ctx->vr_xc, &ctx->vr_yc, &ctx->vr_th, &ctx->vr_xbar, &ctx->vr_ybar, &ctx->vr_thbar};
        for (auto* b : vd) CU(b->reserve(Ha));
        DevBuf<int32_t>* vi[] = {&ctx->vr_h2orig, &ctx->vr_label, &ctx->vr_h2rep, &ctx->vr_urank, &ctx->vr_oi};
        for (auto* b : vi) CU(b->reserve(Ha));
        CU(ctx->vr_n.reserve(size_t(cap_long))); CU(ctx->vr_ncl.reserve(size_t(cap_long))); CU(ctx->vr_order.reserve(size_t(cap_long) + 8));
        CU(ctx->vr_off.reserve(size_t(cap_long)));
        rs.desc_short = ctx->d_run_desc.p; rs.desc_long = ctx->d_run_desc.p + cap_short; rs.ctr = ctx->d_run_ctr.p;
        rs.cap_long = int32_t(cap_long); rs.cap_rows = H;
        rs.vrt.xc = ctx->vr_xc.p; rs.vrt.yc = ctx->vr_yc.p; rs.vrt.th = ctx->vr_th.p; rs.vrt.xbar = ctx->vr_xbar.p; rs.vrt.ybar = ctx->vr_ybar.p;
        rs.vrt.thbar = ctx->vr_thbar.p; rs.vrt.h2orig = ctx->vr_h2orig.p; rs.vrt.label = ctx->vr_label.p; rs.vrt.h2rep = ctx->vr_h2rep.p;
        rs.vrt.urank = ctx->vr_urank.p;
        rs.v_n = ctx->vr_n.p; rs.v_ncl = ctx->vr_ncl.p; rs.v_order = ctx->vr_order.p; rs.v_off = ctx->vr_off.p; rs.v_oi = ctx->vr_oi.p;
    }
    launch_form_fill(pc, x1, y1, h1_begin, h1_end, ctx->d_wedge_h1.p, x2, y2, th2, ctx->d_ang2.p, E2, ctx->h2ix, ctx->d_m.p, ctx->d_n.p, ctx->d_row_off.p, rt,
                     ctx->d_ncl.p, ctx->d_big_list.p, ctx->d_big_list.p + E1a, any_big ? &rs : nullptr, st, ctx->stream2, ctx->ev_fork,
                     ctx->ev_join);   // order[E1], cnts[5]
    launch_scan_i32_i64(ctx->d_ncl.p, ctx->d_u_off.p, E1, st);
    ctx->tm.kernel_launches += 14;
    cudaEventRecord(ctx->ev[1], st);
    CU(cudaMemcpyAsync(&ctx->pin->U, ctx->d_u_off.p + E1, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int64_t U = ctx->pin->U;
    ctx->U = U;
    // ---- unique hypotheses (one per cluster) + wedge of the view-1 edge in every validation view
    const size_t Ua = size_t(std::max<int64_t>(U, 1));
    CU(ctx->u_h1.reserve(Ua)); CU(ctx->u_row.reserve(Ua)); CU(ctx->u_mult.reserve(Ua)); CU(ctx->u_count.reserve(Ua));
    CU(ctx->u_xy.reserve(Ua)); CU(ctx->u_geom.reserve(Ua));
    CU(ctx->d_w31.reserve(size_t(Va) * E1a));
    CU(ctx->d_support.reserve(size_t(Va) * Ua));
    ctx->have_inter = false;
    if (ctx->keep_inter) { CU(ctx->d_w32.reserve(size_t(Va) * Ua)); CU(ctx->d_tgt.reserve(size_t(Va) * Ua)); ctx->have_inter = true; }
    CU(cudaMemsetAsync(ctx->u_mult.p, 0, Ua * 4, st));
    UniqTable ut = uniq_table(ctx);
    launch_uniq_fill(pc, x1, y1, th1, rt, H, ctx->d_row_off.p, ctx->d_u_off.p, ut, st);
    launch_wedge31(pc, ctx->d_vc.p, x1, y1, ctx->d_n.p, h1_begin, h1_end, E1, ctx->d_w31.p, ctx->d_ctr.p, st);
    ctx->tm.kernel_launches += 2;
    cudaEventRecord(ctx->ev[4], st);
    // ---- validation: reprojection + wedge + parallel test + support test, one thread per (unique hypothesis, view)
    ValidateArgs a;
    std::memset(&a, 0, sizeof(a));
    a.ut = ut; a.U = U; a.x1 = x1; a.y1 = y1; a.w31 = ctx->d_w31.p; a.w31_stride = E1; a.vc = ctx->d_vc.p;
    a.vali_view = ctx->d_vali.p; a.edge_off = ctx->d_edge_off.p; a.gi = ctx->ea.gi;
    a.cos_thresh = pc.cos_thresh; a.parallel_deg = pc.parallel_deg; a.delta = pc.delta; a.rows = pc.rows; a.cols = pc.cols; a.n_vali = V; a.strict = ctx->cfg.strict_libm_order;
    a.support = ctx->d_support.p; a.sup_stride = int64_t(Ua); a.ctr = ctx->d_ctr.p;
    if (ctx->keep_inter) { a.w32_out = ctx->d_w32.p; a.tgt_out = ctx->d_tgt.p; }
    launch_validate(pc, a, false, st);
    ctx->tm.kernel_launches++;
    cudaEventRecord(ctx->ev[5], st);
    // ---- vote + ratio test
    launch_count_u(ctx->d_support.p, int64_t(Ua), U, V, ctx->u_count.p, st);
    launch_select(pc, ctx->d_row_off.p, h1_begin, h1_end, ctx->u_count.p, rt, ctx->d_nvalid.p, ctx->d_nsel.p, ctx->d_big_list.p + E1a + 6, st);   // cnts[6] = max selected rows of an edge
    launch_scan_i32_i64(ctx->d_nvalid.p, ctx->d_valid_off.p, E1, st);
    launch_scan_i32_i64(ctx->d_nsel.p, ctx->d_pair_off.p, E1, st);
    ctx->tm.kernel_launches += 4;
    int64_t* tot = ctx->pin->tot;
    CU(cudaMemcpyAsync(&tot[0], ctx->d_valid_off.p + E1, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&tot[1], ctx->d_pair_off.p + E1, 8, cudaMemcpyDeviceToHost, st));
    static_assert(sizeof(DevCounters) == 6 * CVX_CTR_SLOTS * 8, "DevCounters layout");
    const DevCounters& hc = ctx->pin->hc;
    CU(cudaMemcpyAsync(&ctx->pin->hc, ctx->d_ctr.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(ctx->pin->cnts, ctx->d_big_list.p + E1a, 8 * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    ctx->n_valid = tot[0]; ctx->P = tot[1];
    const size_t NVa = size_t(std::max<int64_t>(ctx->n_valid, 1)), Pa = size_t(std::max<int64_t>(ctx->P, 1));
    CU(ctx->v_h1.reserve(NVa)); CU(ctx->v_row.reserve(NVa)); CU(ctx->v_count.reserve(NVa)); CU(ctx->v_sel.reserve(NVa));
    CU(ctx->p_h1.reserve(Pa)); CU(ctx->p_h2rep.reserve(Pa)); CU(ctx->p_row.reserve(Pa)); CU(ctx->p_count.reserve(Pa));
    CU(ctx->p_sup.reserve(Pa * Va));
    launch_select_fill(ctx->d_row_off.p, h1_begin, h1_end, rt, ctx->d_valid_off.p, ctx->d_pair_off.p, ctx->v_h1.p, ctx->v_row.p, ctx->v_count.p,
                       ctx->v_sel.p, ctx->p_h1.p, ctx->p_h2rep.p, ctx->p_row.p, ctx->p_count.p, st);
    launch_gather_supports(ctx->p_row.p, ctx->r_row_u.p, ctx->P, V, ctx->d_support.p, int64_t(Ua), ctx->p_sup.p, st);
    ctx->tm.kernel_launches += 2;
    cudaEventRecord(ctx->ev[3], st);
    CU(cudaGetLastError());
    ctx->timings_pending = true;      // the last two launches are still in flight: phase times are read in cvx_get_timings
    // ---- counts
    cvx_counts& c = ctx->counts;
    std::memset(&c, 0, sizeof(c));
    c.n_h1_total = E1; c.n_hypotheses = H; c.n_validations = H * V; c.n_valid_rows = ctx->n_valid; c.n_pairs = ctx->P;
    c.n_unique_hypotheses = U;
    for (int k = 0; k < CVX_CTR_SLOTS; k++) {
        c.n_degenerate_wedges += int64_t(hc.n_degenerate[k]); c.n_supported += int64_t(hc.n_supported[k]);
        c.n_parallel += int64_t(hc.n_parallel[k]); c.n_inliers += int64_t(hc.n_inliers[k]);
        c.n_entries_visited += int64_t(hc.n_visited[k]); c.n_heavy_queries += int64_t(hc.n_heavy[k]);
    }
    c.n_h1_processed = ctx->pin->cnts[5];                             // view-1 edges with at least one hypothesis (k_order_by_n)
    c.n_candidates = -1;   // filled lazily by cvx_get_counts (needs m[])
    ctx->have_run = true; ctx->have_final = false;
    return CVX_OK;
}

int cvx_get_counts(cvx_ctx* ctx, cvx_counts* out) {
    if (!ctx || !out || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "cvx_get_counts: no run");
    if (ctx->counts.n_candidates < 0) {
        CU(cudaSetDevice(ctx->device));
        const int E1 = ctx->edge_n[ctx->h1];
        std::vector<int32_t> m(std::max(E1, 1));
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaMemcpy(m.data(), ctx->d_m.p, size_t(E1) * 4, cudaMemcpyDeviceToHost));
        int64_t s = 0;
        for (int i = 0; i < E1; i++) s += m[i];
        ctx->counts.n_candidates = s;
    }
    *out = ctx->counts;
    return CVX_OK;
}

#define D2HA(dst, src, n, T)                                                                                      \
    do { if ((dst) && (n) > 0) CU(cudaMemcpyAsync((dst), (src), size_t(n) * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream)); } while (0)
// stream-ordered and complete on return: the context stream is non-blocking, so a plain cudaMemcpy would not wait for the
// launches cvx_run_edge_sketch / run_finalize leave in flight
#define D2H(dst, src, n, T)                                                                                       \
    do { if ((dst) && (n) > 0) { CU(cudaMemcpyAsync((dst), (src), size_t(n) * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream)); \
                                 CU(cudaStreamSynchronize(ctx->stream)); } } while (0)

int cvx_get_h2_angles(cvx_ctx* ctx, double* out) {
    if (!ctx || ctx->h2 < 0) return fail(ctx, CVX_ERR_INVALID, "no pair");
    CU(cudaSetDevice(ctx->device));
    D2H(out, ctx->d_ang2.p, ctx->edge_n[ctx->h2], double);
    return CVX_OK;
}
int cvx_get_h1_wedges(cvx_ctx* ctx, double* out, int32_t* flag) {
    if (!ctx || ctx->h1 < 0) return fail(ctx, CVX_ERR_INVALID, "no pair");
    CU(cudaSetDevice(ctx->device));
    D2H(out, ctx->d_wedge_h1.p, ctx->edge_n[ctx->h1], double2);
    D2H(flag, ctx->d_wflag_h1.p, ctx->edge_n[ctx->h1], int32_t);
    return CVX_OK;
}
int cvx_get_h1_table(cvx_ctx* ctx, int32_t* m, int32_t* n, int64_t* row_off) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run");
    CU(cudaSetDevice(ctx->device));
    const int E1 = ctx->edge_n[ctx->h1];
    D2H(m, ctx->d_m.p, E1, int32_t);
    D2H(n, ctx->d_n.p, E1, int32_t);
    if (row_off && !ctx->h_row_off_valid) {
        ctx->h_row_off.resize(size_t(E1) + 1);
        D2H(ctx->h_row_off.data(), ctx->d_row_off.p, E1 + 1, int64_t);
        ctx->h_row_off_valid = true;
    }
    if (row_off)
        for (int i = 0; i < E1; i++) row_off[i] = (ctx->h_row_off[i + 1] > ctx->h_row_off[i]) ? ctx->h_row_off[i] : -1;
    return CVX_OK;
}
int cvx_get_rows(cvx_ctx* ctx, int32_t* h1, int32_t* h2_orig, int32_t* local_idx, int32_t* raw_at_local, double* xc, double* yc,
                 double* theta_orig, double* xbar, double* ybar, double* thbar, int32_t* label, int32_t* h2_rep, int32_t* count) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run");
    CU(cudaSetDevice(ctx->device));
    const int64_t H = ctx->H;
    D2H(h1, ctx->r_h1.p, H, int32_t); D2H(h2_orig, ctx->r_h2orig.p, H, int32_t); D2H(local_idx, ctx->r_local.p, H, int32_t);
    D2H(raw_at_local, ctx->r_rawat.p, H, int32_t); D2H(xc, ctx->r_xc.p, H, double); D2H(yc, ctx->r_yc.p, H, double);
    D2H(theta_orig, ctx->r_th.p, H, double); D2H(xbar, ctx->r_xbar.p, H, double); D2H(ybar, ctx->r_ybar.p, H, double);
    D2H(thbar, ctx->r_thbar.p, H, double); D2H(label, ctx->r_label.p, H, int32_t); D2H(h2_rep, ctx->r_h2rep.p, H, int32_t);
    D2H(count, ctx->r_count.p, H, int32_t);
    return CVX_OK;
}
int cvx_get_support_matrix(cvx_ctx* ctx, int32_t* out) {
    if (!ctx || !ctx->have_run || !out) return fail(ctx, CVX_ERR_INVALID, "no run");
    CU(cudaSetDevice(ctx->device));
    const int64_t H = ctx->H;
    const int V = ctx->n_vali;
    if (H == 0 || V == 0) return CVX_OK;
    CU(ctx->d_support_t.reserve(size_t(H) * V));
    launch_expand_support(ctx->d_support.p, std::max<int64_t>(ctx->U, 1), ctx->r_row_u.p, H, V, ctx->d_support_t.p, ctx->stream);
    ctx->tm.kernel_launches++;
    CU(cudaStreamSynchronize(ctx->stream));
    D2H(out, ctx->d_support_t.p, H * V, int32_t);
    return CVX_OK;
}
int cvx_set_keep_intermediates(cvx_ctx* ctx, int on) {
    if (!ctx) return CVX_ERR_INVALID;
    ctx->keep_inter = on != 0;
    return CVX_OK;
}
int cvx_get_intermediates(cvx_ctx* ctx, double* wedge31, double* wedge32, double* tgt) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run");
    if ((wedge32 || tgt) && !ctx->have_inter && ctx->H > 0)
        return fail(ctx, CVX_ERR_INVALID, "intermediates were not kept: call cvx_set_keep_intermediates(ctx, 1) before the run");
    CU(cudaSetDevice(ctx->device));
    const int E1 = ctx->edge_n[ctx->h1], V = ctx->n_vali;
    const int64_t H = ctx->H, U = ctx->U;
    // device layout is view-major over unique hypotheses; the ABI returns [E1][V][2] and [H][V][2]
    CU(cudaStreamSynchronize(ctx->stream));
    if (wedge31) {
        std::vector<double> t(size_t(V) * E1 * 2);
        CU(cudaMemcpy(t.data(), ctx->d_w31.p, t.size() * 8, cudaMemcpyDeviceToHost));
        for (int v = 0; v < V; v++) for (int i = 0; i < E1; i++) for (int k = 0; k < 2; k++)
            wedge31[(size_t(i) * V + v) * 2 + k] = t[(size_t(v) * E1 + i) * 2 + k];
    }
    if ((wedge32 || tgt) && H > 0) {
        std::vector<int32_t> row_u(H);
        CU(cudaMemcpy(row_u.data(), ctx->r_row_u.p, size_t(H) * 4, cudaMemcpyDeviceToHost));
        for (int which = 0; which < 2; which++) {
            double* dst = which ? tgt : wedge32;
            if (!dst) continue;
            std::vector<double> t(size_t(V) * U * 2);
            CU(cudaMemcpy(t.data(), which ? (void*)ctx->d_tgt.p : (void*)ctx->d_w32.p, t.size() * 8, cudaMemcpyDeviceToHost));
            for (int v = 0; v < V; v++) for (int64_t r = 0; r < H; r++) for (int k = 0; k < 2; k++)
                dst[(size_t(r) * V + v) * 2 + k] = t[(size_t(v) * U + row_u[r]) * 2 + k];
        }
    }
    return CVX_OK;
}
int cvx_get_valid_rows(cvx_ctx* ctx, int32_t* h1, int32_t* row, int32_t* count, int32_t* selected) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run");
    CU(cudaSetDevice(ctx->device));
    const int64_t n = ctx->n_valid;
    D2H(h1, ctx->v_h1.p, n, int32_t); D2H(row, ctx->v_row.p, n, int32_t); D2H(count, ctx->v_count.p, n, int32_t); D2H(selected, ctx->v_sel.p, n, int32_t);
    return CVX_OK;
}
int cvx_get_pairs(cvx_ctx* ctx, int32_t* h1, int32_t* h2rep, int32_t* row, int32_t* count, int32_t* supports) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run");
    CU(cudaSetDevice(ctx->device));
    const int64_t P = ctx->P;
    // asynchronous copies + one synchronize: pinned destinations stream at link speed, pageable ones behave as before
    D2HA(h1, ctx->p_h1.p, P, int32_t); D2HA(h2rep, ctx->p_h2rep.p, P, int32_t); D2HA(row, ctx->p_row.p, P, int32_t);
    D2HA(count, ctx->p_count.p, P, int32_t); D2HA(supports, ctx->p_sup.p, P * ctx->n_vali, int32_t);
    CU(cudaStreamSynchronize(ctx->stream));
    return CVX_OK;
}

// EdgeSketch_Core.cpp:480-526 on the host: per view-1 edge std::sort of the valid rows by count, ratio-test filter, n_views-row capacity
static int reference_order(cvx_ctx* ctx, std::vector<int64_t>& idx) {
    const int64_t nv = ctx->n_valid;
    std::vector<int32_t> h1(nv), cnt(nv), sel(nv);
    D2H(h1.data(), ctx->v_h1.p, nv, int32_t); D2H(cnt.data(), ctx->v_count.p, nv, int32_t); D2H(sel.data(), ctx->v_sel.p, nv, int32_t);
    idx.clear();
    idx.reserve(size_t(ctx->P));
    int64_t pair_base = 0;
    for (int64_t i = 0; i < nv;) {
        int64_t j = i;
        while (j < nv && h1[j] == h1[i]) j++;
        // valid_pairs_with_counts in ascending row order, EdgeSketch_Core.cpp:480-485
        std::vector<std::pair<int, int>> vp;
        std::vector<int64_t> pair_of(j - i, -1);
        int64_t q = pair_base;
        for (int64_t k = i; k < j; k++) { vp.push_back({int(k - i), cnt[k]}); if (sel[k]) pair_of[k - i] = q++; }
        std::sort(vp.begin(), vp.end(), [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.second > b.second; });   // :488-491
        int taken = 0;
        for (const auto& pr : vp)
            if (pair_of[pr.first] >= 0 && taken < ctx->cfg.n_views) { idx.push_back(pair_of[pr.first]); taken++; }   // :500-526
        pair_base = q;
        i = j;
    }
    return CVX_OK;
}
int cvx_get_reference_order(cvx_ctx* ctx, int64_t* idx, int64_t* n_out) {
    if (!ctx || !ctx->have_run || !idx || !n_out) return fail(ctx, CVX_ERR_INVALID, "cvx_get_reference_order: bad arguments");
    CU(cudaSetDevice(ctx->device));
    std::vector<int64_t> order;
    const int rc = reference_order(ctx, order);
    if (rc) return rc;
    std::copy(order.begin(), order.end(), idx);
    *n_out = int64_t(order.size());
    return CVX_OK;
}

static int run_finalize(cvx_ctx* ctx) {
    if (ctx->have_final) return CVX_OK;
    const int64_t P = ctx->P;
    const size_t Pa = size_t(std::max<int64_t>(P, 1));
    CU(ctx->f_gc.reserve(Pa * 3)); CU(ctx->f_gw.reserve(Pa * 3)); CU(ctx->f_tw.reserve(Pa * 3)); CU(ctx->f_h2c.reserve(Pa * 2)); CU(ctx->f_nan.reserve(Pa));
    const double *x1 = ctx->ea.x + ctx->edge_off[ctx->h1], *y1 = ctx->ea.y + ctx->edge_off[ctx->h1], *th1 = ctx->ea.th + ctx->edge_off[ctx->h1];
    const double *x2 = ctx->ea.x + ctx->edge_off[ctx->h2], *y2 = ctx->ea.y + ctx->edge_off[ctx->h2], *th2 = ctx->ea.th + ctx->edge_off[ctx->h2];
    cudaEventRecord(ctx->ev[6], ctx->stream);
    launch_finalize(ctx->pc, x1, y1, th1, x2, y2, th2, ctx->p_h1.p, ctx->p_h2rep.p, P, ctx->f_gc.p, ctx->f_gw.p, ctx->f_tw.p, ctx->f_h2c.p,
                    ctx->f_nan.p, ctx->stream);
    ctx->tm.kernel_launches++;
    cudaEventRecord(ctx->ev[7], ctx->stream);
    CU(cudaGetLastError());
    ctx->final_timing_pending = true;          // asynchronous: results are consumed on the same stream (or by synchronising getters)
    ctx->have_final = true;
    return CVX_OK;
}

int cvx_finalize(cvx_ctx* ctx, double* gamma_cam, double* gamma_world, double* tangent_world, double* h2_corrected_xy, int32_t* nan_flag) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "cvx_finalize: no run");
    CU(cudaSetDevice(ctx->device));
    const int rc = run_finalize(ctx);
    if (rc) return rc;
    const int64_t P = ctx->P;
    D2HA(gamma_cam, ctx->f_gc.p, P * 3, double); D2HA(gamma_world, ctx->f_gw.p, P * 3, double); D2HA(tangent_world, ctx->f_tw.p, P * 3, double);
    D2HA(h2_corrected_xy, ctx->f_h2c.p, P * 2, double); D2HA(nan_flag, ctx->f_nan.p, P, int32_t);
    CU(cudaStreamSynchronize(ctx->stream));
    return CVX_OK;
}

int cvx_get_device_results(cvx_ctx* ctx, cvx_device_results* out) {
    if (!ctx || !out || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run");
    CU(cudaSetDevice(ctx->device));
    const int rc = run_finalize(ctx);
    if (rc) return rc;
    out->n_pairs = ctx->P; out->n_vali_views = ctx->n_vali; out->pad = 0;
    out->pair_h1 = ctx->p_h1.p; out->pair_h2rep = ctx->p_h2rep.p; out->pair_count = ctx->p_count.p; out->pair_supports = ctx->p_sup.p;
    out->gamma_cam = ctx->f_gc.p; out->gamma_world = ctx->f_gw.p; out->tangent_world = ctx->f_tw.p;
    return CVX_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// claimed edges (EdgeSketch_Core.cpp:776-813)
// ------------------------------------------------------------------------------------------------------------------
int cvx_claims_reset(cvx_ctx* ctx) {
    if (!ctx) return CVX_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    const size_t words = std::max<size_t>(ctx->arena_cap / 32, 1);
    if (!ctx->d_claim || ctx->claim_words != words) {
        if (ctx->d_claim) cudaFree(ctx->d_claim);
        ctx->d_claim = nullptr;
        CU(cudaMalloc(&ctx->d_claim, words * 4));
        ctx->claim_words = words;
    }
    CU(cudaMemsetAsync(ctx->d_claim, 0, words * 4, ctx->stream));
    return CVX_OK;
}
int cvx_claims_accumulate(cvx_ctx* ctx) {
    if (!ctx || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "cvx_claims_accumulate: no run");
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_claim) { const int rc = cvx_claims_reset(ctx); if (rc) return rc; }
    const int rc = run_finalize(ctx);   // the NaN / dropped-representative flags decide which pairs reach the insertions
    if (rc) return rc;
    ClaimArgs a;
    a.p_h1 = ctx->p_h1.p; a.p_h2rep = ctx->p_h2rep.p; a.p_sup = ctx->p_sup.p; a.nan_flag = ctx->f_nan.p; a.P = ctx->P;
    a.row_off = ctx->d_row_off.p; a.rawat = ctx->r_rawat.p; a.label = ctx->r_label.p;
    a.vali_view = ctx->d_vali.p; a.edge_off = ctx->d_edge_off.p; a.n_vali = ctx->n_vali; a.view1 = ctx->h1; a.view2 = ctx->h2;
    a.E2 = ctx->edge_n[ctx->h2]; a.bits = ctx->d_claim; a.keep = nullptr;
    if (ctx->pin->cnts[6] > ctx->cfg.n_views && ctx->P > 0) {
        // Some view-1 edge selected more rows than paired_edge holds for it (n_views, EdgeSketch_Core.cpp:513-526): only the rows the
        // reference keeps -- in its std::sort order -- reach the EdgeMapping insertions (:682-726).  Rare (never on ABC-NEF).
        std::vector<int64_t> order;
        const int rc2 = reference_order(ctx, order);
        if (rc2) return rc2;
        std::vector<uint8_t> keep(size_t(ctx->P), 0);
        for (int64_t i : order) keep[size_t(i)] = 1;
        CU(ctx->d_keep.reserve(size_t(ctx->P)));
        CU(cudaMemcpyAsync(ctx->d_keep.p, keep.data(), size_t(ctx->P), cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));   // keep[] is a local
        a.keep = ctx->d_keep.p;
    }
    launch_claim_pairs(a, ctx->stream);
    ctx->tm.kernel_launches++;
    CU(cudaGetLastError());
    return CVX_OK;
}
int cvx_claims_get_counts(cvx_ctx* ctx, int64_t* claimed, int64_t* total) {
    if (!ctx) return CVX_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    const int N = ctx->n_views;
    if (total) for (int v = 0; v < N; v++) total[v] = ctx->edge_n[v];
    if (!claimed) return CVX_OK;
    if (!ctx->d_claim) { for (int v = 0; v < N; v++) claimed[v] = 0; return CVX_OK; }
    const int rc = sync_tables(ctx);
    if (rc) return rc;
    CU(ctx->d_claim_cnt.reserve(N));
    launch_claim_count(ctx->d_claim, ctx->d_edge_off.p, ctx->d_edge_n.p, N, ctx->d_claim_cnt.p, ctx->stream);
    ctx->tm.kernel_launches++;
    std::vector<unsigned long long> h(N);
    CU(cudaMemcpyAsync(h.data(), ctx->d_claim_cnt.p, size_t(N) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaGetLastError());
    for (int v = 0; v < N; v++) claimed[v] = int64_t(h[v]);
    return CVX_OK;
}
int cvx_claims_get_mask(cvx_ctx* ctx, int view, uint8_t* out) {
    if (!ctx || !out || view < 0 || view >= ctx->n_views) return fail(ctx, CVX_ERR_INVALID, "cvx_claims_get_mask: bad arguments");
    CU(cudaSetDevice(ctx->device));
    const int n = ctx->edge_n[view];
    if (n == 0) return CVX_OK;
    if (!ctx->d_claim) { std::memset(out, 0, size_t(n)); return CVX_OK; }
    CU(ctx->d_claim_mask.reserve(size_t(n)));
    launch_claim_expand(ctx->d_claim, ctx->edge_off[view], n, ctx->d_claim_mask.p, ctx->stream);
    ctx->tm.kernel_launches++;
    CU(cudaMemcpyAsync(out, ctx->d_claim_mask.p, size_t(n), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaGetLastError());
    return CVX_OK;
}

// Row-balanced split of view 1 (include/curvix_b200.h): the count pass over all view-1 edges, an exclusive scan, and the edges at
// which the running row count crosses k / parts of the total.
int cvx_split_h1_by_rows(cvx_ctx* ctx, int parts, int32_t* bounds) {
    if (!ctx || ctx->h1 < 0 || parts < 1 || !bounds) return fail(ctx, CVX_ERR_INVALID, "cvx_split_h1_by_rows: bad arguments / no pair");
    CU(cudaSetDevice(ctx->device));
    const int E1 = ctx->edge_n[ctx->h1], E2 = ctx->edge_n[ctx->h2];
    const int E1a = std::max(E1, 1);
    cudaStream_t st = ctx->stream;
    const double *x1 = ctx->ea.x + ctx->edge_off[ctx->h1], *y1 = ctx->ea.y + ctx->edge_off[ctx->h1];
    const double *x2 = ctx->ea.x + ctx->edge_off[ctx->h2], *y2 = ctx->ea.y + ctx->edge_off[ctx->h2], *th2 = ctx->ea.th + ctx->edge_off[ctx->h2];
    CU(ctx->d_m.reserve(E1a)); CU(ctx->d_n.reserve(E1a)); CU(ctx->d_row_off.reserve(E1a + 1));
    CU(cudaMemsetAsync(ctx->d_m.p, 0, E1a * 4, st)); CU(cudaMemsetAsync(ctx->d_n.p, 0, E1a * 4, st));
    launch_form_count(ctx->pc, x1, y1, 0, E1, ctx->d_wedge_h1.p, x2, y2, th2, ctx->d_ang2.p, E2, ctx->h2ix, ctx->d_m.p, ctx->d_n.p, st);
    launch_scan_i32_i64(ctx->d_n.p, ctx->d_row_off.p, E1, st);
    ctx->tm.kernel_launches += 2;
    std::vector<int64_t> off(size_t(E1) + 1);
    D2H(off.data(), ctx->d_row_off.p, E1 + 1, int64_t);
    ctx->have_run = false; ctx->have_final = false; ctx->h_row_off_valid = false;   // the tables of an earlier run were overwritten
    const int64_t H = off[size_t(E1)];
    bounds[0] = 0;
    for (int k = 1; k < parts; k++) {
        const int64_t target = (H * k) / parts;
        const auto it = std::lower_bound(off.begin(), off.end(), target);
        bounds[k] = std::max<int32_t>(bounds[k - 1], int32_t(std::min<int64_t>(it - off.begin(), E1)));
    }
    bounds[parts] = E1;
    return CVX_OK;
}

int cvx_get_timings(cvx_ctx* ctx, cvx_timings* out) {
    if (!ctx || !out) return CVX_ERR_INVALID;
    if (ctx->timings_pending) {
        CU(cudaSetDevice(ctx->device));
        CU(cudaEventSynchronize(ctx->ev[3]));
        cudaEventElapsedTime(&ctx->tm.run_ms, ctx->ev[0], ctx->ev[3]);
        cudaEventElapsedTime(&ctx->tm.form_ms, ctx->ev[0], ctx->ev[1]);
        cudaEventElapsedTime(&ctx->tm.prep_ms, ctx->ev[1], ctx->ev[4]);
        cudaEventElapsedTime(&ctx->tm.validate_ms, ctx->ev[4], ctx->ev[5]);
        cudaEventElapsedTime(&ctx->tm.select_ms, ctx->ev[5], ctx->ev[3]);
        ctx->timings_pending = false;
    }
    if (ctx->set_pair_timing_pending) {
        CU(cudaSetDevice(ctx->device));
        CU(cudaEventSynchronize(ctx->ev[9]));
        cudaEventElapsedTime(&ctx->tm.set_pair_ms, ctx->ev[8], ctx->ev[9]);
        ctx->set_pair_timing_pending = false;
    }
    if (ctx->final_timing_pending) {
        CU(cudaSetDevice(ctx->device));
        CU(cudaEventSynchronize(ctx->ev[7]));
        cudaEventElapsedTime(&ctx->tm.finalize_ms, ctx->ev[6], ctx->ev[7]);
        ctx->final_timing_pending = false;
    }
    *out = ctx->tm;
    return CVX_OK;
}
int cvx_get_stream(cvx_ctx* ctx, void** stream) {
    if (!ctx || !stream) return CVX_ERR_INVALID;
    *stream = (void*)ctx->stream;
    return CVX_OK;
}
int cvx_synchronize(cvx_ctx* ctx) {
    if (!ctx) return CVX_ERR_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    return CVX_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// stand-alone support test (BASELINE.json config 5)
// ------------------------------------------------------------------------------------------------------------------
int cvx_support_problem_create(cvx_ctx* ctx, int G, const int64_t* group_off, int V, const int64_t* view_off, const double* ang31,
                               const double* ang32, const double* cos_t, const double* sin_t, const double* wedge31, const double* wedge32,
                               const double* tgt, double cos_thresh, double parallel_deg, cvx_support_problem** out) {
    if (!ctx || !out || G <= 0 || V <= 0 || !group_off || !view_off) return fail(ctx, CVX_ERR_INVALID, "cvx_support_problem_create: bad arguments");
    CU(cudaSetDevice(ctx->device));
    cvx_support_problem* p = new cvx_support_problem();
    p->ctx = ctx; p->G = G; p->V = V; p->H = group_off[G]; p->n_edges = view_off[V]; p->thr = cos_thresh; p->par = parallel_deg;
    cudaStream_t st = ctx->stream;
    const size_t NE = size_t(std::max<int64_t>(p->n_edges, 1)), H = size_t(std::max<int64_t>(p->H, 1));
    int64_t max_edges = 1;
    for (int v = 0; v < V; v++) max_edges = std::max(max_edges, view_off[v + 1] - view_off[v]);
    p->gi.CS = pick_cell_stride(max_edges);
    CU(p->scan_tmp.reserve(size_t(2 * V) * size_t(scan_tiles(p->gi.CS))));
    CU(p->view_off.reserve(V + 1)); CU(p->cell_end.reserve(size_t(V) * size_t(p->gi.CS))); CU(p->cell_end_b.reserve(size_t(V) * size_t(p->gi.CS))); CU(p->vg.reserve(V)); CU(p->minmax.reserve(size_t(V) * 4));
    CU(p->support.reserve(size_t(V) * H)); CU(p->cell.reserve(NE)); CU(p->idx.reserve(NE)); CU(p->idx_b.reserve(NE));
    CU(p->row_group.reserve(H));
    CU(p->a31.reserve(NE)); CU(p->a32.reserve(NE)); CU(p->cs.reserve(NE)); CU(p->sn.reserve(NE));
    CU(p->rec.reserve(NE)); CU(p->rec_b.reserve(NE));
    p->gi.A = GridLayout{p->rec.p, p->idx.p, p->cell_end.p}; p->gi.B = GridLayout{p->rec_b.p, p->idx_b.p, p->cell_end_b.p};
    p->gi.vg = p->vg.p; p->gi.minmax = p->minmax.p;
    CU(p->w31.reserve(size_t(V) * G)); CU(p->w32.reserve(size_t(V) * H)); CU(p->tgt.reserve(size_t(V) * H));
    std::vector<int32_t> rg(H, 0);
    for (int g = 0; g < G; g++) for (int64_t r = group_off[g]; r < group_off[g + 1]; r++) rg[size_t(r)] = g;
    CU(cudaMemcpyAsync(p->row_group.p, rg.data(), size_t(p->H) * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->view_off.p, view_off, (V + 1) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->a31.p, ang31, p->n_edges * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->a32.p, ang32, p->n_edges * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->cs.p, cos_t, p->n_edges * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->sn.p, sin_t, p->n_edges * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->w31.p, wedge31, size_t(V) * G * 16, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->w32.p, wedge32, size_t(V) * p->H * 16, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(p->tgt.p, tgt, size_t(V) * p->H * 16, cudaMemcpyHostToDevice, st));
    GridSample gs;
    std::memset(&gs, 0, sizeof(gs));
    gs.w31 = p->w31.p; gs.n31 = G; gs.w32 = p->w32.p; gs.n32 = p->H;
    launch_stage_given(p->a31.p, p->a32.p, p->cs.p, p->sn.p, p->view_off.p, V, max_edges, p->cell.p, p->gi, gs, p->scan_tmp.p, st,
                       &ctx->tm.kernel_launches);
    cudaEventCreate(&p->e0); cudaEventCreate(&p->e1);
    CU(cudaStreamSynchronize(st));   // rg (host vector) must outlive the copy
    CU(cudaGetLastError());
    *out = p;
    return CVX_OK;
}

int cvx_support_problem_run(cvx_support_problem* p, float* kernel_ms) {
    if (!p) return CVX_ERR_INVALID;
    cvx_ctx* ctx = p->ctx;
    CU(cudaSetDevice(ctx->device));
    ValidateArgs a;
    std::memset(&a, 0, sizeof(a));
    a.ut.h1 = p->row_group.p; a.U = p->H; a.w31 = p->w31.p; a.w31_stride = p->G; a.w32_in = p->w32.p; a.tgt_in = p->tgt.p;
    a.vali_view = nullptr; a.edge_off = p->view_off.p; a.gi = p->gi;
    a.cos_thresh = p->thr; a.parallel_deg = p->par; a.n_vali = p->V;
    a.support = p->support.p; a.sup_stride = std::max<int64_t>(p->H, 1); a.ctr = nullptr;
    cudaEventRecord(p->e0, ctx->stream);
    launch_validate(ctx->pc, a, true, ctx->stream);
    cudaEventRecord(p->e1, ctx->stream);
    ctx->tm.kernel_launches++;
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaGetLastError());
    float ms = 0;
    cudaEventElapsedTime(&ms, p->e0, p->e1);
    if (kernel_ms) *kernel_ms = ms;
    return CVX_OK;
}
int cvx_support_problem_stats(cvx_support_problem* p, int64_t* n_inliers, int64_t* n_visited, int64_t* n_supported, int64_t* n_parallel) {
    if (!p) return CVX_ERR_INVALID;
    cvx_ctx* ctx = p->ctx;
    CU(cudaSetDevice(ctx->device));
    ValidateArgs a;
    std::memset(&a, 0, sizeof(a));
    a.ut.h1 = p->row_group.p; a.U = p->H; a.w31 = p->w31.p; a.w31_stride = p->G; a.w32_in = p->w32.p; a.tgt_in = p->tgt.p;
    a.vali_view = nullptr; a.edge_off = p->view_off.p; a.gi = p->gi;
    a.cos_thresh = p->thr; a.parallel_deg = p->par; a.n_vali = p->V;
    a.support = p->support.p; a.sup_stride = std::max<int64_t>(p->H, 1); a.ctr = ctx->d_ctr.p;
    CU(cudaMemsetAsync(ctx->d_ctr.p, 0, sizeof(DevCounters), ctx->stream));
    launch_validate(ctx->pc, a, true, ctx->stream);
    ctx->tm.kernel_launches++;
    DevCounters hc;
    CU(cudaMemcpyAsync(&hc, ctx->d_ctr.p, sizeof(hc), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaGetLastError());
    int64_t inl = 0, vis = 0, sup = 0, par = 0;
    for (int k = 0; k < CVX_CTR_SLOTS; k++) {
        inl += int64_t(hc.n_inliers[k]); vis += int64_t(hc.n_visited[k]); sup += int64_t(hc.n_supported[k]); par += int64_t(hc.n_parallel[k]);
    }
    if (n_inliers) *n_inliers = inl;
    if (n_visited) *n_visited = vis;
    if (n_supported) *n_supported = sup;
    if (n_parallel) *n_parallel = par;
    return CVX_OK;
}
int cvx_support_problem_result(cvx_support_problem* p, int32_t* support) {
    if (!p || !support) return CVX_ERR_INVALID;
    cvx_ctx* ctx = p->ctx;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpy(support, p->support.p, size_t(p->V) * p->H * 4, cudaMemcpyDeviceToHost));
    return CVX_OK;
}
int cvx_support_problem_destroy(cvx_support_problem* p) {
    if (!p) return CVX_ERR_INVALID;
    cudaSetDevice(p->ctx->device);
    p->view_off.release(); p->scan_tmp.release(); p->cell_end.release(); p->cell_end_b.release(); p->idx_b.release(); p->rec_b.release(); p->vg.release(); p->minmax.release(); p->support.release(); p->cell.release();
    p->idx.release(); p->row_group.release(); p->a31.release(); p->a32.release(); p->cs.release(); p->sn.release();
    p->rec.release(); p->w31.release(); p->w32.release(); p->tgt.release();
    cudaEventDestroy(p->e0); cudaEventDestroy(p->e1);
    delete p;
    return CVX_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------
// internal interface of cvx_comm.cu (cvx_internal.h)
// ------------------------------------------------------------------------------------------------------------------
int cvx_internal_fail(cvx_ctx* ctx, int code, const char* msg) { return fail(ctx, code, msg ? msg : ""); }
CvxCommState** cvx_internal_comm_slot(cvx_ctx* ctx) { return &ctx->comm; }
int cvx_internal_device(cvx_ctx* ctx, int* device, cudaStream_t* stream) {
    if (device) *device = ctx->device;
    if (stream) *stream = ctx->stream;
    return CVX_OK;
}
void cvx_internal_count_launch(cvx_ctx* ctx, int n) { ctx->tm.kernel_launches += n; }
int cvx_internal_results(cvx_ctx* ctx, CvxResultsView* out) {
    if (!ctx || !out || !ctx->have_run) return fail(ctx, CVX_ERR_INVALID, "no run to gather");
    CU(cudaSetDevice(ctx->device));
    const int rc = run_finalize(ctx);
    if (rc) return rc;
    out->P = ctx->P; out->V = ctx->n_vali; out->device = ctx->device;
    out->view1 = ctx->h1; out->view2 = ctx->h2; out->run_begin = ctx->run_begin; out->run_end = ctx->run_end;
    out->h1 = ctx->p_h1.p; out->h2rep = ctx->p_h2rep.p; out->count = ctx->p_count.p; out->nan = ctx->f_nan.p; out->sup = ctx->p_sup.p;
    out->gc = ctx->f_gc.p; out->gw = ctx->f_gw.p; out->tw = ctx->f_tw.p; out->h2c = ctx->f_h2c.p;
    out->stream = ctx->stream;
    return CVX_OK;
}
int cvx_internal_claim_words(cvx_ctx* ctx, uint32_t** words, size_t* n_words) {
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_claim) { const int rc = cvx_claims_reset(ctx); if (rc) return rc; }
    *words = ctx->d_claim; *n_words = ctx->claim_words;
    return CVX_OK;
}
