// =====================================================================================
// ref_capi.hpp -- TEST INFRASTRUCTURE ONLY.  C interface around the UNMODIFIED 3D Curvix reference sources
// (/root/reference/Edge_Reconst/*.cpp), compiled against the stand-in headers of oracle/ref_shim/ into
// oracle/_ref/libcurvix_ref.so (git-ignored; recipe: oracle/ref_shim/Makefile).  It pins the port oracle
// (oracle/curvix_oracle.cpp) and serves as the timed CPU reference arm of bench.py.  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
// =====================================================================================
#pragma once
#include <cstdint>

extern "C" {

typedef struct ref_config {
    double delta;                 // YAML delta
    double delta_theta_deg;       // YAML delta_theta
    double parallel_angle_deg;    // YAML Parallel_Epipolar_Line_Angle
    double reproj_dist_thresh;    // YAML Reproj_Dist_Thresh (unused on the live path)
    double stop_ratio;            // YAML Ratio_Of_Claimed_Edges_to_Stop
    double fx, fy, cx, cy;
    int32_t n_views, img_rows, img_cols, max_support_views, num_threads, toed_thresh, max_passes, reserved;
    char dataset_path[512];       // only read by ref_select_next_views / ref_consolidate (the reference re-reads edge files there)
    char dataset_name[64];
    char scene_name[64];
} ref_config;

typedef struct ref_ctx ref_ctx;

const char* ref_build_info(void);   // compiler, flags, EIGEN_SHIM_SUM_ORDER
ref_ctx* ref_create(const ref_config* c);
void ref_destroy(ref_ctx* r);
int ref_set_cameras(ref_ctx* r, int n, const double* R /*n x 9 row-major*/, const double* T /*n x 3*/);
int ref_set_view_edges(ref_ctx* r, int view, int n, const double* x, const double* y, const double* theta, const double* strength);
// bench sampling: view-1 edges [0, n) only (applied at the next ref_set_pair); n < 0 = all
int ref_limit_h1(ref_ctx* r, int n);
// EdgeSketch_Core: hyp01/02_view_indx, Set_Hypothesis_Views_Camera, All_Edgels refill, Set_Hypothesis_Views_Edgels
int ref_set_pair(ref_ctx* r, int h1, int h2);
int ref_run(ref_ctx* r);        // Run_3D_Edge_Sketch
int ref_finalize(ref_ctx* r);   // Finalize_Edge_Pairs_and_Reconstruct_3D_Edges + Stack_3D_Edges (files go to ../../outputs if it exists)
int ref_clear(ref_ctx* r);      // Clear_Data
// Calculate_Edge_Support_Ratios_And_Select_Next_Views (re-reads the edge files under dataset_path)
int ref_select_next_views(ref_ctx* r, int32_t* best1, int32_t* best2, double* avg_ratio, int32_t* abort_flag);
// EdgeMapping::add3DToSupportingEdgesMapping on the context's map (rotation / translation of `image` from ref_set_cameras): lets a
// different producer of 3-D edges -- the port oracle, the CUDA path -- feed the reference's own host-side curve grouping
int ref_mapping_add(ref_ctx* r, const double* edge_3d, const double* tangent_3d_world, const double* supporting_edge /*2*/,
                    const double* supporting_edge_uncorrected /*3*/, int image, int edge_idx);
// many at once: n records of 12 doubles {X Y Z, tx ty tz, x y, ux uy utheta, 0} + image[n], edge_idx[n]
int ref_mapping_add_many(ref_ctx* r, int64_t n, const double* rec, const int32_t* image, const int32_t* edge_idx);
// EdgeMapping::Setup_Data_Parameters + consolidate_3D_edges (writes ../../outputs/final_curves.txt etc.)
int ref_consolidate(ref_ctx* r);
void ref_get_times(ref_ctx* r, double* pair_edges, double* finalize, double* next_views);

int64_t ref_n_pairs(ref_ctx* r);                     // rows of paired_edge_final
int ref_get_pairs(ref_ctx* r, int32_t* h1, int32_t* h2, int32_t* supports /*P x V*/);
int ref_get_3d(ref_ctx* r, double* gamma_cam /*P x 3*/, double* tangent_world /*P x 3*/, double* gamma_world /*P x 3: rows of all_3D_Edges added by the last pass*/);
// all_supported_indices: one (m x V) matrix per processed view-1 edge, in processing order (ascending with 1 thread)
int64_t ref_n_support_blocks(ref_ctx* r);
int ref_get_support_block_rows(ref_ctx* r, int64_t* rows /*n_blocks*/);
int ref_get_support_blocks(ref_ctx* r, double* out /*sum(rows) x V row-major; uninitialised cells are NaN*/);
// hypo2_clusters flattened: key (h1, h2) -> members
int64_t ref_n_cluster_keys(ref_ctx* r, int64_t* n_members_total);
int ref_get_cluster_map(ref_ctx* r, int32_t* key_h1, int32_t* key_h2, int64_t* off /*n_keys+1*/, int32_t* members);
// claimed (view, edge) sets of EdgeMapping::edge_3D_to_supporting_edges, OR-ed into masks[v][edge]
int ref_accumulate_claims(ref_ctx* r, uint8_t** masks);
int64_t ref_n_3d_keys(ref_ctx* r);                   // distinct 3-D points in the map

// ---- the helper classes, called directly (cameras: n x 9 row-major R, n x 3 T; K 9 row-major; edges n x 4 row-major)
void ref_relative_pose(const double* R1, const double* T1, const double* R2, const double* T2, double* R21, double* T21, double* R12, double* T12);
void ref_fundamental(const double* K1, const double* K2, const double* R21, const double* T21, double* F);
void ref_inverse3(const double* A, double* out);
// which: 0 getOreList(h1=ref_idx, h2=valid_idx), 1 getOreListVali(valid_idx, ref_idx)
void ref_ore_list(int which, int n_views, const double* R, const double* T, const double* K1, const double* K2, int valid_idx, int ref_idx,
                  int n, const double* edges, double delta, int rows, int cols, double* out /*n*/);
void ref_ore_list_bar(int n_views, const double* R, const double* T, const double* K1, const double* K2, int valid_idx, int ref_idx,
                      int n, const double* edges, double delta, int rows, int cols, double* out /*n x 2*/, double* epipole /*3*/);
// getHYPO2_idx_Ore: returns the count, indices into out (capacity n)
int ref_hypo2_idx_ore(int n, const double* angles, double lo, double hi, int32_t* out);
// getedgels_HYPO2_Ore + edgelsHYPO2_epipolar_correction: returns the number of surviving rows (10 doubles each)
int ref_epipolar_correction(int m, const double* edgels_h2 /*m x 4*/, const double* edgel_h1 /*4*/, const double* F21, const double* F12,
                            const double* idx_raw /*m*/, double* out /*m x 10*/);
// EdgeClusterer(n, rows, h1).performClustering(idx_raw, Edges_HYPO2 (only its row count matters), corrected)
int ref_cluster(int n, const double* h2_final /*n x 4*/, int h1_idx, int m, const double* idx_raw /*m*/, int e2_rows,
                const double* corrected /*n x 10*/, int32_t* labels /*n*/, double* hypo2_idx /*n*/, double* averaged /*n x 4*/,
                int32_t* n_clusters, int32_t* key_h2 /*cap n*/, int64_t* key_off /*n+1*/, int32_t* key_members /*cap n*n*/, int32_t* n_keys);
void ref_gamma3(int n_views, const double* R, const double* T, const double* K1, const double* K2, int h1, int h2, int valid_idx,
                int n, const double* edges_h1 /*n x 4*/, const double* edges_h2 /*n x 4*/, double* out /*n x 4*/);
int ref_support_idx(double delta_theta_deg, const double* tgt /*2*/, int e3, const double* tangents /*e3 x 2*/, int k, const double* inliers /*k*/);
void ref_triangulate(const double* p1, const double* p2, const double* R21, const double* T21, const double* K, double* out3);
void ref_tangent3d(const double* e1 /*x y th*/, const double* e2, const double* K1, const double* K2, const double* R1, const double* T1,
                   const double* R2, const double* T2, double* out3);
// helpers of the curve consolidation: util.cpp:123-131, 186-213, 241-252 (called from edge_mapping.cpp:262, 541-575)
void ref_closest_vector_to_line(const double* P1, const double* d1, const double* P2, double* out3);
void ref_shortest_align_euler_deg(const double* v1, const double* v2, double* out3);
void ref_euler_to_rotation(double roll, double pitch, double yaw, double* out9 /*row-major*/);
}
