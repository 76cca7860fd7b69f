// =====================================================================================
// cvx_kernels.cuh -- launch interface between the C-ABI layer (cvx_api.cu) and the sm_100a kernels
// (cvx_kernels.cu).  All pointers are device pointers unless noted.
// =====================================================================================
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cvx_math.cuh"

namespace cvx {

// Per-edge arrays of all views, concatenated; view v occupies [edge_off[v], edge_off[v] + edge_n[v]).
struct EdgeArena {
    // static per view (uploaded / derived once per cvx_set_view_edges)
    double *x, *y, *th, *cs, *sn;
    // per pass scratch: epipolar angles about the two hypothesis-view epipoles and the bin id
    double *a31, *a32;
    int32_t* bin;
    // per pass staged copy, grouped by angular bin of a31 (the "bin grid" the support kernel walks)
    double *s_a31, *s_a32, *s_cs, *s_sn;
    int32_t* s_idx;
};

struct DevCounters {
    unsigned long long n_degenerate, n_supported, n_parallel, n_inliers;
};

// Hypothesis table (one row per epipolar-corrected candidate), grouped by view-1 edge.
struct RowTable {
    int32_t *h1, *h2orig, *local, *rawat, *label, *h2rep, *count, *flags;
    double *xc, *yc, *th, *xbar, *ybar, *thbar;
    double *thn, *orient;   // clustering scratch: normalised orientation, per-label average orientation
};

struct BinGrid {
    int32_t* bins;     // [V][BS] after staging: bins[v][b] = end of bin b (exclusive) relative to the view's base
    int32_t B, BS;     // B angular bins over [0,180], BS = B + 1 (last = NaN bin)
    double scale;      // B / 180
};

void launch_edge_trig(const double* th, double* cs, double* sn, int64_t n, cudaStream_t st);

// stage all validation views of a pass: angles -> histogram -> scan -> scatter
void launch_stage_views(const EdgeArena& ea, const int64_t* edge_off, const int32_t* edge_n, const int32_t* vali_view, int n_vali,
                        int max_edges, const ViewConst* vc, const BinGrid& bg, cudaStream_t st, int* launches);

// stage externally supplied per-view arrays (stand-alone support test): a31/a32/cs/sn concatenated with view_off[V+1]
void launch_stage_given(const double* a31, const double* a32, const double* cs, const double* sn, const int64_t* view_off, int n_vali,
                        int64_t max_edges, int32_t* bin_tmp, const BinGrid& bg, double* s_a31, double* s_a32, double* s_cs, double* s_sn,
                        int32_t* s_idx, int32_t* vali_iota, cudaStream_t st);

void launch_h2_angles(const double* x2, const double* y2, int n, double ex, double ey, double* out, cudaStream_t st);
void launch_h1_wedges(const PassConst& pc, const double* x1, const double* y1, int n, double2* wedge, int32_t* flag, cudaStream_t st);

// formation: count pass (m, n per view-1 edge) and fill pass (rows + clustering)
void launch_form_count(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                       const double* x2, const double* y2, const double* th2, const double* ang2, int E2, int32_t* m_out,
                       int32_t* n_out, cudaStream_t st);
void launch_form_fill(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                      const double* x2, const double* y2, const double* th2, const double* ang2, int E2, const int32_t* n_of,
                      const int64_t* row_off, RowTable rt, cudaStream_t st);

void launch_scan_i32_i64(const int32_t* in, int64_t* out, int n, cudaStream_t st);   // out has n + 1 entries

void launch_wedge31(const PassConst& pc, const ViewConst* vc, const double* x1, const double* y1, const int32_t* n_of, int h1_begin,
                    int h1_end, int E1, double2* w31, DevCounters* ctr, cudaStream_t st);
void launch_reproject(const PassConst& pc, const ViewConst* vc, const double* x1, const double* y1, const double* th1, RowTable rt,
                      int64_t row_begin, int64_t row_end, double2* w32, double2* tgt, DevCounters* ctr, cudaStream_t st);

struct SupportArgs {
    const int64_t* group_off;   // [G + 1] first row of each group
    int g_begin, g_end;         // groups processed by this launch
    int64_t row_base;           // rows of w32/tgt are stored relative to row_base with stride chunk_rows
    int64_t chunk_rows;
    const double2* w31; int64_t w31_stride;   // [V][w31_stride], indexed by group
    const double2* w32; const double2* tgt;   // [V][chunk_rows]
    const int32_t* vali_view; const int64_t* edge_off;   // view slot -> view, view -> arena offset
    BinGrid bg;
    const double *s_a31, *s_a32, *s_cs, *s_sn; const int32_t* s_idx;
    double cos_thresh, parallel_deg;
    int32_t* support; int64_t sup_stride;     // [V][sup_stride], indexed by absolute row
    DevCounters* ctr;
    int n_vali;
};
void launch_support(const SupportArgs& a, cudaStream_t st);

void launch_select(const PassConst& pc, const int64_t* row_off, int h1_begin, int h1_end, const int32_t* support, int64_t sup_stride,
                   RowTable rt, int32_t* n_valid, int32_t* n_sel, cudaStream_t st);
void launch_select_fill(const int64_t* row_off, int h1_begin, int h1_end, RowTable rt, const int64_t* valid_off, const int64_t* pair_off,
                        int32_t* v_h1, int32_t* v_row, int32_t* v_count, int32_t* v_sel, int32_t* p_h1, int32_t* p_h2rep, int32_t* p_row,
                        int32_t* p_count, cudaStream_t st);
void launch_gather_supports(const int32_t* p_row, int64_t P, int n_vali, const int32_t* support, int64_t sup_stride, int32_t* out,
                            cudaStream_t st);
void launch_transpose_support(const int32_t* support, int64_t sup_stride, int64_t H, int n_vali, int32_t* out, cudaStream_t st);

void launch_finalize(const PassConst& pc, const double* x1, const double* y1, const double* th1, const double* x2, const double* y2,
                     const double* th2, const int32_t* p_h1, const int32_t* p_h2rep, int64_t P, double* gamma_cam, double* gamma_world,
                     double* tangent_world, double* h2c, int32_t* nan_flag, cudaStream_t st);

}  // namespace cvx
