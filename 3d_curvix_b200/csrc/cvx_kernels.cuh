// =====================================================================================
// cvx_kernels.cuh -- launch interface between the C-ABI layer (cvx_api.cu) and the sm_100a kernels
// (cvx_kernels.cu).  All pointers are device pointers unless noted.
// =====================================================================================
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cvx_math.cuh"

namespace cvx {

// One angular index of the validation views: every view's edges counting-sorted by the bin of one of their two
// epipolar angles.  Entry e of view slot v lives at [edge_off[view] + e]; bins[v][b] = END (exclusive) of bin b.
struct ViewIndex {
    double2* ang;    // (a31, a32): epipolar angle about the epipole of hypothesis view 1 / view 2 (degrees, [0,180))
    double2* tan;    // (cos theta, sin theta) of the edge orientation
    int32_t* idx;    // original edge index inside its view
    int32_t* bins;   // [V][BS]
};

// Per-edge arrays of all views, concatenated; view v occupies [edge_off[v], edge_off[v] + edge_n[v]).
struct EdgeArena {
    double *x, *y, *th, *cs, *sn;   // static per view (uploaded / derived once per cvx_set_view_edges)
    double2* ang;                   // per pass scratch: (a31, a32)
    int32_t *bin1, *bin2;           // per pass scratch: bin of a31 / a32
    ViewIndex ix1, ix2;             // per pass: index keyed by a31 and index keyed by a32
};

#define CVX_CTR_SLOTS 64
struct DevCounters {   // spread over slots to keep same-address atomics off the critical path; summed on the host
    unsigned long long n_degenerate[CVX_CTR_SLOTS], n_supported[CVX_CTR_SLOTS], n_parallel[CVX_CTR_SLOTS], n_inliers[CVX_CTR_SLOTS];
};

// Hypothesis table (one row per epipolar-corrected candidate), grouped by view-1 edge.
struct RowTable {
    int32_t *h1, *h2orig, *local, *rawat, *label, *h2rep, *count, *flags;
    int32_t *urank;          // rank of the row's cluster among the clusters of its view-1 edge
    int32_t *row_u;          // index of the row's cluster in the unique-hypothesis table
    double *xc, *yc, *th, *xbar, *ybar, *thbar;
    double *thn, *orient;    // clustering scratch of the large-n fallback
};

// Unique hypotheses: one entry per cluster (every member row of a cluster carries the same averaged edge, so
// reprojection, wedge and support test are evaluated once per cluster and shared by its members).
struct UniqTable {
    int32_t *h1, *row, *mult, *count;   // view-1 edge, row of the cluster root, cluster size, supporting views
    double2* xy;                        // clustered view-2 location (xbar, ybar)
    double4* geom;                      // view-independent reprojection terms: (rho1, Tv1x, Tv1y, Tv1z)
};

struct BinGrid {
    int32_t B, BS;     // B angular bins over [0,180], BS = B + 1 (last = NaN bin)
    double scale;      // B / 180
};

void launch_edge_trig(const double* th, double* cs, double* sn, int64_t n, cudaStream_t st);

// stage all validation views of a pass: angles -> two histograms -> scan -> scatter into the two indexes
void launch_stage_views(const EdgeArena& ea, const int64_t* edge_off, const int32_t* edge_n, const int32_t* vali_view, int n_vali,
                        int max_edges, const ViewConst* vc, const BinGrid& bg, cudaStream_t st, int* launches);
// stage externally supplied per-view arrays (stand-alone support test): a31/a32/cs/sn concatenated with view_off[V+1]
void launch_stage_given(const double* a31, const double* a32, const double* cs, const double* sn, const int64_t* view_off, int n_vali,
                        int64_t max_edges, int32_t* bin1, int32_t* bin2, const BinGrid& bg, ViewIndex ix1, ViewIndex ix2, cudaStream_t st);

void launch_h2_angles(const double* x2, const double* y2, int n, double ex, double ey, double* out, cudaStream_t st);
void launch_h1_wedges(const PassConst& pc, const double* x1, const double* y1, int n, double2* wedge, int32_t* flag, cudaStream_t st);

// formation: count pass (m, n per view-1 edge) and fill pass (rows + clustering + per-edge cluster count)
void launch_form_count(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                       const double* x2, const double* y2, const double* th2, const double* ang2, int E2, int32_t* m_out,
                       int32_t* n_out, cudaStream_t st);
void launch_form_fill(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                      const double* x2, const double* y2, const double* th2, const double* ang2, int E2, const int32_t* n_of,
                      const int64_t* row_off, RowTable rt, int32_t* n_clusters, cudaStream_t st);
// unique-hypothesis table from the clustered rows (+ view-independent reprojection terms)
void launch_uniq_fill(const PassConst& pc, const double* x1, const double* y1, const double* th1, RowTable rt, int64_t H,
                      const int64_t* row_off, const int64_t* u_off, UniqTable ut, cudaStream_t st);

void launch_scan_i32_i64(const int32_t* in, int64_t* out, int n, cudaStream_t st);   // out has n + 1 entries

void launch_wedge31(const PassConst& pc, const ViewConst* vc, const double* x1, const double* y1, const int32_t* n_of, int h1_begin,
                    int h1_end, int E1, double2* w31, DevCounters* ctr, cudaStream_t st);

// The validation kernel: per (unique hypothesis, view) reprojection + wedge + parallel test + support test.
struct ValidateArgs {
    // hypotheses
    UniqTable ut; int64_t U;
    const double *x1, *y1;                    // view-1 edges (for Gamma1)
    const double2* w31; int64_t w31_stride;   // [V][w31_stride] indexed by view-1 edge (or group)
    const ViewConst* vc;
    // GIVEN mode (stand-alone support test): wedge32 / tangent are inputs, [V][U]; group of a hypothesis from ut.h1
    const double2* w32_in; const double2* tgt_in;
    // per-view edge indexes
    const int32_t* vali_view; const int64_t* edge_off;
    BinGrid bg; ViewIndex ix1, ix2;
    double cos_thresh, parallel_deg, delta;
    int rows, cols, n_vali;
    // outputs
    int32_t* support; int64_t sup_stride;     // [V][sup_stride] indexed by unique hypothesis
    double2* w32_out; double2* tgt_out;       // optional (parity intermediates), [V][U]
    DevCounters* ctr;
};
void launch_validate(const PassConst& pc, const ValidateArgs& a, bool given, cudaStream_t st);

// vote (per unique hypothesis), ratio test (per view-1 edge), compaction
void launch_count_u(const int32_t* support, int64_t sup_stride, int64_t U, int n_vali, int32_t* count_u, cudaStream_t st);
void launch_select(const PassConst& pc, const int64_t* row_off, int h1_begin, int h1_end, const int32_t* count_u, RowTable rt,
                   int32_t* n_valid, int32_t* n_sel, cudaStream_t st);
void launch_select_fill(const int64_t* row_off, int h1_begin, int h1_end, RowTable rt, const int64_t* valid_off, const int64_t* pair_off,
                        int32_t* v_h1, int32_t* v_row, int32_t* v_count, int32_t* v_sel, int32_t* p_h1, int32_t* p_h2rep, int32_t* p_row,
                        int32_t* p_count, cudaStream_t st);
// out[p][v] = support[v][row_u[p_row[p]]]
void launch_gather_supports(const int32_t* p_row, const int32_t* row_u, int64_t P, int n_vali, const int32_t* support, int64_t sup_stride,
                            int32_t* out, cudaStream_t st);
// out[r][v] = support[v][row_u[r]]  (the reference's supported_indices layout, expanded from the unique table)
void launch_expand_support(const int32_t* support, int64_t sup_stride, const int32_t* row_u, int64_t H, int n_vali, int32_t* out,
                           cudaStream_t st);

void launch_finalize(const PassConst& pc, const double* x1, const double* y1, const double* th1, const double* x2, const double* y2,
                     const double* th2, const int32_t* p_h1, const int32_t* p_h2rep, int64_t P, double* gamma_cam, double* gamma_world,
                     double* tangent_world, double* h2c, int32_t* nan_flag, cudaStream_t st);

// claimed-edge bookkeeping (Calculate_Edge_Support_Ratios_And_Select_Next_Views, EdgeSketch_Core.cpp:785-797): one bit per edge
// of every view in the arena (view v starts at word edge_off[v] / 32; slots are 32-edge aligned)
struct ClaimArgs {
    const int32_t *p_h1, *p_h2rep, *p_sup, *nan_flag; int64_t P;
    const int64_t* row_off; const int32_t *rawat, *label;   // hypothesis table (cluster lookup through the reference's key quirk)
    const int32_t* vali_view; const int64_t* edge_off; int n_vali, view1, view2, E2;
    uint32_t* bits;
};
void launch_claim_pairs(const ClaimArgs& a, cudaStream_t st);
void launch_claim_count(const uint32_t* bits, const int64_t* edge_off, const int32_t* edge_n, int n_views, unsigned long long* out, cudaStream_t st);
void launch_claim_expand(const uint32_t* bits, int64_t off, int n, uint8_t* out, cudaStream_t st);

}  // namespace cvx
