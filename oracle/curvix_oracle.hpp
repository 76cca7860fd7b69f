// =====================================================================================
// curvix_oracle.hpp -- TEST INFRASTRUCTURE ONLY (CPU oracle, never shipped, never timed as product).
//
// A CPU restatement (Eigen-free C++17 + OpenMP, plain double) of the 3D Curvix edge-sketch hot
// path of C-H-Chien/3D_Curvix.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.  The product (3d_curvix_b200/) never links it.
//
// PARITY STATUS: pinned against the reference itself.  The reference's unmodified sources compile against the stand-in
// Eigen / yaml-cpp headers of oracle/ref_shim into oracle/_ref/libcurvix_ref.so; tests/test_oracle_vs_ref.py shows this port
// equal to it bit for bit (getOreList, getOreListBar on 2-hit rows, getHYPO2_idx_Ore, epipolar correction, clustering,
// reprojected tangents, wedges, getSupportIdx, 3-D tangents; whole passes: accepted pairs, support matrices, cluster map) and
// to 1e-6 for linearTriangulation; tests/golden/ref_abc_*.npz are reference-generated vectors (tests/test_oracle_golden.py).
// What stays a restatement: Eigen's reduction order inside 3x3 products (oracle/ref_shim/README.md, test_sum_order_sensitivity).
//
// Operation order mimics a non-fast-math build: 3-term dot products are summed left to right,
// Matrix3d::inverse() is cofactor * (1/det), transcendental functions are scalar glibc calls.
// Build with -O2 -ffp-contract=off (see oracle/Makefile).
// =====================================================================================
#pragma once
#include <cstdint>

extern "C" {

// Mirrors cvx_config in include/curvix_b200.h (same field order; tests fill both through ctypes).
typedef struct orc_config {
    double delta;                       // YAML delta                         EdgeSketch_Core.cpp:40
    double delta_theta_deg;             // YAML delta_theta                   EdgeSketch_Core.cpp:41
    double parallel_angle_deg;          // YAML Parallel_Epipolar_Line_Angle  EdgeSketch_Core.cpp:44
    double lowes_ratio;                 // LOWES_RATIO_THRESHOLD              definitions.h:14
    double epip_tangency_displ_thresh;  // EPIP_TANGENCY_DISPL_THRESH         definitions.h:20
    double location_perturbation;       // LOCATION_PERTURBATION              definitions.h:21
    double orient_perturbation;         // ORIENT_PERTURBATION                definitions.h:22
    double cluster_dist_thresh;         // CLUSTER_DIST_THRESH                definitions.h:23
    double cluster_orient_thresh_deg;   // CLUSTER_ORIENT_THRESH              definitions.h:24
    double cluster_gauss_sigma;         // CLUSTER_ORIENT_GAUSS_SIGMA         definitions.h:26
    double pi_truncated;                // PI (3.1415926)                     definitions.h:54
    int32_t n_views;                    // Total_Num_Of_Images
    int32_t img_rows;
    int32_t img_cols;
    int32_t max_support_views;          // YAML Max_Num_Of_Support_Views
    int32_t max_cluster_size;           // MAX_CLUSTER_SIZE                   definitions.h:25
    int32_t reserved;
} orc_config;

typedef struct orc_counts {
    int64_t n_h1_total;        // edges in hypothesis view 1
    int64_t n_h1_processed;    // H1 edges that reached validation (n >= 1)
    int64_t n_candidates;      // sum of m (raw wedge candidates)
    int64_t n_hypotheses;      // sum of n (rows surviving epipolar correction)
    int64_t n_validations;     // sum of n * V
    int64_t n_valid_rows;      // rows with count >= Max_Num_Of_Support_Views
    int64_t n_pairs;           // accepted rows (after the ratio test and the per-H1 capacity)
    int64_t n_degenerate_wedges; // wedge rows flagged (reference UB made explicit)
    int64_t n_supported;       // validations with a support index != -2
    int64_t n_parallel;        // validations short-circuited by the parallel-wedge test
    int64_t n_inliers;         // sum over validations of |wedge31 ∩ wedge32|
    int64_t n_unique_hypotheses; // clusters (sum over processed view-1 edges)
} orc_counts;

typedef struct orc_ctx orc_ctx;

void  orc_default_config(orc_config* c);
orc_ctx* orc_create(const orc_config* c);
void  orc_destroy(orc_ctx* o);
// R: n x 9 row-major, T: n x 3, K: 9 doubles (k_per_view == 0) or n x 9 (k_per_view == 1)
int   orc_set_cameras(orc_ctx* o, int n, const double* R, const double* T, const double* K, int k_per_view);
int   orc_set_view_edges(orc_ctx* o, int view, int n, const double* x, const double* y, const double* theta,
                         const double* strength);
int   orc_set_hypothesis_pair(orc_ctx* o, int h1, int h2);
// mode 0 = faithful (reference loop nest, Theta(E3) scans, by-value copies), 1 = hoisted (per-view precompute)
// keep_intermediates != 0 stores wedge31 / wedge32 / reprojected tangents and per-validation margins.
int   orc_run_edge_sketch(orc_ctx* o, int h1_begin, int h1_end, int mode, int num_threads, int keep_intermediates);
void  orc_get_counts(orc_ctx* o, orc_counts* out);

// pass-level arrays
int   orc_get_h2_angles(orc_ctx* o, double* out /*E2*/);
int   orc_get_h1_wedges(orc_ctx* o, double* out /*E1 x 2*/, int32_t* flag /*E1*/);
// per-H1 arrays (size E1): m, n, row offset (exclusive scan of n over processed edges, -1 when not processed)
int   orc_get_h1_table(orc_ctx* o, int32_t* m, int32_t* n, int64_t* row_off, double* form_margin);
// per-row arrays (size n_hypotheses)
int   orc_get_rows(orc_ctx* o, int32_t* h1, int32_t* h2_orig, int32_t* local_idx, int32_t* raw_at_local,
                   double* xc, double* yc, double* theta_orig, double* xbar, double* ybar, double* thbar,
                   int32_t* label, int32_t* h2_rep, int32_t* count);
int   orc_get_support_matrix(orc_ctx* o, int32_t* out /*H x V row-major*/);
// needs keep_intermediates: all H x V (x2) row-major; wedge31 is (n_h1_processed x V x 2) in processed order
int   orc_get_intermediates(orc_ctx* o, double* wedge31, double* wedge32, double* tgt, double* vali_margin,
                            int32_t* wflag /*H x V*/);
// valid rows (count >= Nsup) in ascending (h1,row) order
int   orc_get_valid_rows(orc_ctx* o, int32_t* h1, int32_t* row, int32_t* count, int32_t* selected);
// accepted pairs in reference order (ascending h1, then std::sort order): supports is P x V
int   orc_get_pairs(orc_ctx* o, int32_t* h1, int32_t* h2rep, int32_t* row, int32_t* count, int32_t* supports);
// finalize: runs EdgeSketch_Core.cpp:575-733 (geometry part) + Stack_3D_Edges transform
int   orc_finalize(orc_ctx* o, double* gamma_cam /*P x 3*/, double* gamma_world, double* tangent_world,
                   double* h2_corrected_xy /*P x 2*/, int32_t* nan_flag /*P*/);
// cluster members of accepted pair p as looked up through hypo2_clusters (EdgeSketch_Core.hpp:193-202)
int   orc_get_pair_cluster(orc_ctx* o, int p, int32_t* members, int cap);

// Claimed edges of the accepted pairs of the last run, as Finalize inserts them into EdgeMapping (EdgeSketch_Core.cpp:682-726)
// and Calculate_Edge_Support_Ratios_And_Select_Next_Views collects them per view (:785-797).  masks[v] points at n_edges(v)
// bytes that are OR-ed with 1 for every claimed edge (accumulate over passes by reusing the buffers).  Needs orc_finalize
// semantics internally (pairs whose representative is dropped or NaN insert nothing).
int   orc_accumulate_claims(orc_ctx* o, uint8_t** masks /*n_views pointers*/);
// Next hypothesis views from the claimed ratios (EdgeSketch_Core.cpp:799-868): literal std::sort by ratio + history filter.
// Returns 1 when avg_ratio > stop_ratio (enable_aborting_3D_edge_sketch).
int   orc_select_next_views(int n_views, const int64_t* claimed, const int64_t* total, const int32_t* history, int n_history,
                            double stop_ratio, int32_t* best1, int32_t* best2, double* avg_ratio);

// stand-alone functions for unit tests / KATs
void  orc_pair_geometry(const double* R1, const double* T1, const double* R2, const double* T2, const double* K1,
                        const double* K2, double* R21, double* T21, double* F, double* e1pix, double* e2pix);
void  orc_wedge_row(double x, double y, const double* F, const double* e1pix, const double* e2pix, double delta,
                    int rows, int cols, double* lo, double* hi, int32_t* flag);
void  orc_triangulate(const double* p1, const double* p2, const double* R21, const double* T21, const double* K,
                      double* out3);
void  orc_tangent3d(const double* e1 /*x y th*/, const double* e2, const double* K1, const double* K2,
                    const double* R1, const double* T1, const double* R2, const double* T2, double* out3);
int   orc_correct_one(const orc_config* cfg, const double* F21, double x1, double y1, double x2, double y2, double th2,
                      double* cx, double* cy);
void  orc_reproject_tangent(double x1, double y1, double th1, double x2, double y2, double th2, const double* K1,
                            const double* K2, const double* R21, const double* T21, const double* R31, const double* T31,
                            double* t3);
double orc_ore_angle(double x, double y, double ex, double ey);
void  orc_inverse3(const double* A, double* out);
// stand-alone support test on pre-formed hypotheses (cfg5): group CSR, per-view staged edges
// ang31/ang32/cs/sn: per view concatenated arrays with offsets view_off[V+1]
int   orc_support_test(int n_groups, const int64_t* group_off, int V, const int64_t* view_off, const double* ang31,
                       const double* ang32, const double* cs, const double* sn, const double* wedge31 /*V x G x 2*/,
                       const double* wedge32 /*V x H x 2*/, const double* tgt /*V x H x 2*/, double thr,
                       double parallel_deg, int32_t* support /*V x H*/, int num_threads);
}
