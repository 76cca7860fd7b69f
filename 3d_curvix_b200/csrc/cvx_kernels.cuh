// =====================================================================================
// cvx_kernels.cuh -- launch interface between the C-ABI layer (cvx_api.cu) and the sm_100a kernels
// (cvx_kernels.cu).  All pointers are device pointers unless noted.
// =====================================================================================
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cvx_math.cuh"

namespace cvx {

// The per-pass acceleration structure of the validation views: a uniform G1 x G2 bin grid over the two epipolar angles
// (a31, a32) of every edge of a view -- the angular image of the pixel grid, since (a31, a32) is a smooth one-to-one map of
// the image plane away from the epipoles.  The reference enumerates wedge31 n wedge32 with two Theta(E3) scans and a
// set_intersection (EdgeSketch_Core.cpp:348-374); here the intersection of the two angle ranges is a small rectangle of
// cells.  Edges of view slot v are counting-sorted by cell, in two layouts: A with cell = b1 * G2 + b2 (the cells of one a31
// row are contiguous in a32) and B with cell = b2 * G1 + b1.  A query walks the layout in which its rectangle has fewer rows
// (a wedge about a nearby epipole is wide in its own angle: one contiguous range in the other layout).  Entry e of a view
// lives at [edge_off[view] + e]; cell_end[v][c] = END (exclusive) of cell c, cell G1*G2 holds NaN angles.
struct ViewGrid {        // per validation view: bin_k = int((a_k - lo_k) * s_k), clamped to [0, G_k)
    double lo1, hi1, s1, lo2, hi2, s2;
    int32_t G1, G2;      // bins along a31 / a32, sized per view so that a cell is about a third of a typical wedge (G1 * G2 < CS)
};
struct GridLayout {
    double4* rec;        // (a31, a32, cos theta, sin theta): the two epipolar angles (degrees, [0,180)) and the edge's unit tangent
    int32_t* idx;        // original edge index inside its view
    int32_t* cell_end;   // [V][CS]
};
struct GridIndex {
    GridLayout A, B;
    ViewGrid* vg;        // [V]
    unsigned long long* minmax;   // [V][4] scratch: ordered keys of min a31, max a31, min a32, max a32
    int32_t CS;          // cell_end stride per view (cell budget + 1, multiple of 4)
};

// Per-edge arrays of all views, concatenated; view v occupies [edge_off[v], edge_off[v] + edge_n[v]).
struct EdgeArena {
    double *x, *y, *th, *cs, *sn;   // static per view (uploaded / derived once per cvx_set_view_edges)
    double2* ang;                   // per pass scratch: (a31, a32)
    int32_t* cell;                  // per pass scratch: (b1 << 16) | b2 of the edge, -1 for NaN angles
    GridIndex gi;                   // per pass: the bin grid of every validation view
};

#define CVX_CTR_SLOTS 64
struct DevCounters {   // spread over slots to keep same-address atomics off the critical path; summed on the host
    unsigned long long n_degenerate[CVX_CTR_SLOTS], n_supported[CVX_CTR_SLOTS], n_parallel[CVX_CTR_SLOTS], n_inliers[CVX_CTR_SLOTS],
        n_visited[CVX_CTR_SLOTS], n_heavy[CVX_CTR_SLOTS];
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


// Scratch of the run-based clustering of the > 128-row edges (cvx_kernels.cu, "clustering by RUNS"): work lists of the short runs,
// and a second hypothesis table into which the long runs are copied as virtual edges for the per-edge clustering kernels.
struct RunScratch {
    int4 *desc_long, *desc_short;      // (view-1 edge, scratch row offset | first sorted position, rows, 0)
    int32_t* ctr;                      // [0] long runs, [1] short runs, [2] scratch rows used
    int32_t cap_long; int64_t cap_rows;
    RowTable vrt;                      // xc, yc, th, h2orig in; xbar, ybar, thbar, label, h2rep, urank out (thn / orient scratch)
    int32_t *v_n, *v_ncl, *v_order;    // rows / cluster count / clustering order (+ 8 class counts) of the virtual edges
    int64_t* v_off;                    // first scratch row of every virtual edge
    int32_t* v_oi;                     // scratch row -> row within its edge
};

void launch_edge_trig(const double* th, double* cs, double* sn, int64_t n, cudaStream_t st);
void launch_ingest_views(const double* sx, const double* sy, const double* st_, const int64_t* tab, int n_views, int64_t max_n, EdgeArena ea,
                         cudaStream_t st);

int pick_cell_stride(int64_t max_edges);   // CS: cell budget per view (~16 cells per edge of the largest view) + 1
// scratch of the segmented cell scan: 2 * n_vali * scan_tiles(CS) ints
int scan_tiles(int CS);
// what the grid sizing samples: wedges of view-1 edges in a validation view and of view-2 edges in it
struct GridSample {
    const double *x1, *y1, *x2, *y2; int E1, E2;    // fused path: edges of the two hypothesis views (wedge_row is evaluated on samples)
    const double2 *w31, *w32; int64_t n31, n32;     // given path: wedge arrays [V][n31], [V][n32]
    double delta; int rows, cols;
};
// stage all validation views of a pass: angles + ranges -> grid sizing -> cell histogram -> scan -> scatter into the grid
void launch_stage_views(const EdgeArena& ea, const int64_t* edge_off, const int32_t* edge_n, const int32_t* vali_view, int n_vali,
                        int max_edges, const ViewConst* vc, const GridSample& gs, int32_t* scan_tmp, cudaStream_t st, int* launches);
// stage externally supplied per-view arrays (stand-alone support test): a31/a32/cs/sn concatenated with view_off[V+1]
void launch_stage_given(const double* a31, const double* a32, const double* cs, const double* sn, const int64_t* view_off, int n_vali,
                        int64_t max_edges, int32_t* cell, GridIndex gi, const GridSample& gs, int32_t* scan_tmp, cudaStream_t st, int* launches);

void launch_h2_angles(const double* x2, const double* y2, int n, double ex, double ey, double* out, cudaStream_t st);
// The view-2 edges counting-sorted by the bin of their epipolar angle (B uniform bins over [0,180], bin B = NaN):
// the candidate search of a view-1 edge (getHYPO2_idx_Ore, PairEdgeHypo.cpp:88-102) reads one bin range instead of all of view 2.
struct AngleIndex {
    double* ang;        // sorted angles
    int32_t* idx;       // original view-2 edge index
    int32_t* bin_end;   // [B + 1] END (exclusive) of every bin
    int32_t B; double scale;   // scale = B / 180
};
int pick_angle_bins(int n);   // B for n edges
void launch_h2_index(const double* ang2, int n, int32_t* bin_scratch, AngleIndex ix, cudaStream_t st, int* launches);
void launch_h1_wedges(const PassConst& pc, const double* x1, const double* y1, int n, double2* wedge, int32_t* flag, cudaStream_t st);

// formation: count pass (m, n per view-1 edge) and fill pass (rows + clustering + per-edge cluster count)
void launch_form_count(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                       const double* x2, const double* y2, const double* th2, const double* ang2, int E2, AngleIndex ix, int32_t* m_out,
                       int32_t* n_out, cudaStream_t st);
void launch_form_fill(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                      const double* x2, const double* y2, const double* th2, const double* ang2, int E2, AngleIndex ix, const int32_t* m_of,
                      const int32_t* n_of, const int64_t* row_off, RowTable rt, int32_t* n_clusters, int32_t* big_list, int32_t* big_count,
                      const RunScratch* rs, cudaStream_t st, cudaStream_t side, cudaEvent_t ev_fork, cudaEvent_t ev_join);
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
    GridIndex gi;
    double cos_thresh, parallel_deg, delta;
    int rows, cols, n_vali;
    int strict;                               // reference libm call order for the view-2 wedge (cvx_config.strict_libm_order)
    // outputs
    int32_t* support; int64_t sup_stride;     // [V][sup_stride] indexed by unique hypothesis
    double2* w32_out; double2* tgt_out;       // optional (parity intermediates), [V][U]
    DevCounters* ctr;
};
void launch_validate(const PassConst& pc, const ValidateArgs& a, bool given, cudaStream_t st);

// vote (per unique hypothesis), ratio test (per view-1 edge), compaction
void launch_count_u(const int32_t* support, int64_t sup_stride, int64_t U, int n_vali, int32_t* count_u, cudaStream_t st);
void launch_select(const PassConst& pc, const int64_t* row_off, int h1_begin, int h1_end, const int32_t* count_u, RowTable rt,
                   int32_t* n_valid, int32_t* n_sel, int32_t* max_sel, cudaStream_t st);
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
    const uint8_t* keep;    // NULL, or 1 for the accepted rows the reference's paired_edge capacity keeps (cvx_get_reference_order)
};
void launch_claim_pairs(const ClaimArgs& a, cudaStream_t st);
void launch_claim_count(const uint32_t* bits, const int64_t* edge_off, const int32_t* edge_n, int n_views, unsigned long long* out, cudaStream_t st);
void launch_claim_expand(const uint32_t* bits, int64_t off, int n, uint8_t* out, cudaStream_t st);

}  // namespace cvx
