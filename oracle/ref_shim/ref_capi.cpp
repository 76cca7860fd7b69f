// =====================================================================================
// ref_capi.cpp -- TEST INFRASTRUCTURE ONLY.  See ref_capi.hpp.  This file holds no reference code: it
// converts plain arrays to the reference's argument types and calls the reference's own classes.
// =====================================================================================
#include "ref_capi.hpp"

#include <omp.h>

#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "EdgeSketch_Core.hpp"   // the reference's header (-I /root/reference/Edge_Reconst), stand-in Eigen/yaml-cpp

#ifndef REF_BUILD_FLAGS
#define REF_BUILD_FLAGS "unknown"
#endif
#define REF_STR2(x) #x
#define REF_STR(x) REF_STR2(x)

namespace {

using Eigen::Matrix3d;
using Eigen::MatrixXd;
using Eigen::Vector3d;

Matrix3d m3(const double* a) {   // row-major 9 -> Matrix3d
    Matrix3d m;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) m(i, j) = a[i * 3 + j];
    return m;
}
Vector3d v3(const double* a) { return Vector3d(a[0], a[1], a[2]); }
void out3(const Matrix3d& m, double* a) {
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) a[i * 3 + j] = m(i, j);
}
void out3(const Vector3d& v, double* a) { a[0] = v(0); a[1] = v(1); a[2] = v(2); }
MatrixXd rows_to_mat(int n, int c, const double* a) {   // row-major n x c -> MatrixXd
    MatrixXd m(n, c);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < c; j++) m(i, j) = a[size_t(i) * c + j];
    return m;
}
void cams(int n, const double* R, const double* T, std::vector<Matrix3d>& aR, std::vector<Vector3d>& aT) {
    aR.clear(); aT.clear();
    for (int i = 0; i < n; i++) { aR.push_back(m3(R + 9 * i)); aT.push_back(v3(T + 3 * i)); }
}

}  // namespace

struct ref_ctx {
    ref_config cfg;
    std::unique_ptr<EdgeSketch_Core> core;
    std::shared_ptr<EdgeMapping> mapping;
    YAML::Node yaml;
    std::vector<MatrixXd> edges;   // the wrapper's copy: Clear_Data empties All_Edgels every pass (cmd/main.cpp re-reads the files)
    int h1_limit = -1;
    Eigen::Index world_rows_last = 0;
};

extern "C" {

const char* ref_build_info(void) {
    return "reference sources compiled unmodified against oracle/ref_shim; " __VERSION__ "; flags: " REF_BUILD_FLAGS
           "; EIGEN_SHIM_SUM_ORDER=" REF_STR(EIGEN_SHIM_SUM_ORDER);
}

ref_ctx* ref_create(const ref_config* c) {
    // The reference clamps its YAML thread count to omp_get_max_threads() (EdgeSketch_Core.cpp:78-79).  Launchers such as torchrun
    // export OMP_NUM_THREADS=1; the caller's explicit request takes precedence over that environment default.
    if (c->num_threads > 0) omp_set_num_threads(c->num_threads);
    ref_ctx* r = new ref_ctx();
    r->cfg = *c;
    YAML::Node& y = r->yaml;
    auto S = [](double v) { char b[64]; std::snprintf(b, sizeof b, "%.17g", v); return std::string(b); };
    y.set("Num_Of_OMP_Threads", std::to_string(c->num_threads));
    y.set("Init_Hypo1_View_Index", "0");
    y.set("Init_Hypo2_View_Index", "1");
    y.set("delta", S(c->delta));
    y.set("delta_theta", S(c->delta_theta_deg));
    y.set("Max_Num_Of_Support_Views", std::to_string(c->max_support_views));
    y.set("TOED_Thresh", std::to_string(c->toed_thresh));
    y.set("Parallel_Epipolar_Line_Angle", S(c->parallel_angle_deg));
    y.set("Reproj_Dist_Thresh", S(c->reproj_dist_thresh));
    y.set("Ratio_Of_Claimed_Edges_to_Stop", S(c->stop_ratio));
    y.set("Max_Num_Of_3D_Edge_Sketch_Passes", std::to_string(c->max_passes));
    y.set("Dataset_Path", c->dataset_path);
    y.set("Dataset_Name", c->dataset_name);
    y.set("Scene_Name", c->scene_name);
    y.set("Total_Num_Of_Images", std::to_string(c->n_views));
    y.set("Image_Rows", std::to_string(c->img_rows));
    y.set("Image_Cols", std::to_string(c->img_cols));
    y.set("Use_Multiple_K", "false");
    y.set("fx", S(c->fx));
    y.set("fy", S(c->fy));
    y.set("cx", S(c->cx));
    y.set("cy", S(c->cy));
    r->core.reset(new EdgeSketch_Core(y));
    r->core->K << c->fx, 0, c->cx, 0, c->fy, c->cy, 0, 0, 1;   // EdgeSketch_Core.cpp:112 (Read_Camera_Data reads files for R, T)
    r->mapping = std::make_shared<EdgeMapping>();
    r->edges.resize(size_t(c->n_views));
    return r;
}
void ref_destroy(ref_ctx* r) { delete r; }

int ref_set_cameras(ref_ctx* r, int n, const double* R, const double* T) {
    if (n != r->cfg.n_views) return -1;
    cams(n, R, T, r->core->All_R, r->core->All_T);
    return 0;
}
int ref_set_view_edges(ref_ctx* r, int view, int n, const double* x, const double* y, const double* theta, const double* strength) {
    if (view < 0 || view >= r->cfg.n_views) return -1;
    MatrixXd m(n, 4);
    for (int i = 0; i < n; i++) { m(i, 0) = x[i]; m(i, 1) = y[i]; m(i, 2) = theta[i]; m(i, 3) = strength ? strength[i] : 1.0; }
    r->edges[size_t(view)] = m;
    return 0;
}
int ref_limit_h1(ref_ctx* r, int n) { r->h1_limit = n; return 0; }

int ref_set_pair(ref_ctx* r, int h1, int h2) {
    EdgeSketch_Core& c = *r->core;
    c.hyp01_view_indx = h1;
    c.hyp02_view_indx = h2;
    c.Set_Hypothesis_Views_Camera();
    c.All_Edgels = r->edges;   // stands for Read_Edgels_Data (cmd/main.cpp:134)
    if (r->h1_limit >= 0 && r->h1_limit < c.All_Edgels[size_t(h1)].rows()) {
        MatrixXd sub = c.All_Edgels[size_t(h1)].topRows(r->h1_limit);
        c.All_Edgels[size_t(h1)] = sub;
    }
    c.Set_Hypothesis_Views_Edgels();
    return 0;
}
int ref_run(ref_ctx* r) { r->core->Run_3D_Edge_Sketch(); return 0; }
int ref_finalize(ref_ctx* r) {
    EdgeSketch_Core& c = *r->core;
    c.Finalize_Edge_Pairs_and_Reconstruct_3D_Edges(r->mapping);
    c.Stack_3D_Edges();
    r->world_rows_last = c.Gamma1s.rows();
    return 0;
}
int ref_clear(ref_ctx* r) { r->core->Clear_Data(); return 0; }
int ref_select_next_views(ref_ctx* r, int32_t* best1, int32_t* best2, double* avg_ratio, int32_t* abort_flag) {
    EdgeSketch_Core& c = *r->core;
    c.Calculate_Edge_Support_Ratios_And_Select_Next_Views(r->mapping);
    *best1 = c.hyp01_view_indx;
    *best2 = c.hyp02_view_indx;
    *avg_ratio = c.avg_ratio;
    *abort_flag = c.enable_aborting_3D_edge_sketch ? 1 : 0;
    return 0;
}
int ref_mapping_add(ref_ctx* r, const double* e3, const double* t3, const double* se, const double* su, int image, int edge_idx) {
    const EdgeSketch_Core& c = *r->core;
    if (image < 0 || size_t(image) >= c.All_R.size()) return -1;
    r->mapping->add3DToSupportingEdgesMapping(v3(e3), v3(t3), Eigen::Vector2d(se[0], se[1]), v3(su), image, edge_idx, c.All_R[size_t(image)],
                                              c.All_T[size_t(image)]);
    return 0;
}
int ref_mapping_add_many(ref_ctx* r, int64_t n, const double* rec, const int32_t* image, const int32_t* edge_idx) {
    for (int64_t i = 0; i < n; i++) {
        const double* q = rec + 12 * i;
        const int rc = ref_mapping_add(r, q, q + 3, q + 6, q + 8, image[i], edge_idx[i]);
        if (rc) return rc;
    }
    return 0;
}
int ref_consolidate(ref_ctx* r) {
    EdgeSketch_Core& c = *r->core;
    r->mapping->Setup_Data_Parameters(r->yaml);
    r->mapping->consolidate_3D_edges(c.All_R, c.All_T, c.K, c.Num_Of_Total_Imgs);
    return 0;
}
void ref_get_times(ref_ctx* r, double* pair_edges, double* finalize, double* next_views) {
    *pair_edges = r->core->pair_edges_time;
    *finalize = r->core->finalize_edge_pair_time;
    *next_views = r->core->find_next_hypothesis_view_time;
}

int64_t ref_n_pairs(ref_ctx* r) { return r->core->paired_edge_final.rows(); }
int ref_get_pairs(ref_ctx* r, int32_t* h1, int32_t* h2, int32_t* supports) {
    const MatrixXd& p = r->core->paired_edge_final;
    const Eigen::Index V = p.cols() - 2;
    for (Eigen::Index i = 0; i < p.rows(); i++) {
        h1[i] = int32_t(p(i, 0));
        h2[i] = int32_t(p(i, 1));
        for (Eigen::Index v = 0; v < V; v++) supports[i * V + v] = int32_t(p(i, 2 + v));
    }
    return 0;
}
int ref_get_3d(ref_ctx* r, double* gamma_cam, double* tangent_world, double* gamma_world) {
    const EdgeSketch_Core& c = *r->core;
    const Eigen::Index P = c.paired_edge_final.rows();
    for (Eigen::Index i = 0; i < P; i++)
        for (int k = 0; k < 3; k++) {
            if (gamma_cam) gamma_cam[i * 3 + k] = c.Gamma1s(i, k);
            if (tangent_world) tangent_world[i * 3 + k] = c.tangent3Ds(i, k);
        }
    if (gamma_world) {
        const Eigen::Index base = c.all_3D_Edges.rows() - r->world_rows_last;
        for (Eigen::Index i = 0; i < P && i < r->world_rows_last; i++)
            for (int k = 0; k < 3; k++) gamma_world[i * 3 + k] = c.all_3D_Edges(base + i, k);
    }
    return 0;
}
int64_t ref_n_support_blocks(ref_ctx* r) { return int64_t(r->core->all_supported_indices.size()); }
int ref_get_support_block_rows(ref_ctx* r, int64_t* rows) {
    for (size_t i = 0; i < r->core->all_supported_indices.size(); i++) rows[i] = r->core->all_supported_indices[i].rows();
    return 0;
}
int ref_get_support_blocks(ref_ctx* r, double* out) {
    size_t o = 0;
    for (const MatrixXd& m : r->core->all_supported_indices)
        for (Eigen::Index i = 0; i < m.rows(); i++)
            for (Eigen::Index j = 0; j < m.cols(); j++) out[o++] = m(i, j);
    return 0;
}
int64_t ref_n_cluster_keys(ref_ctx* r, int64_t* n_members_total) {
    int64_t t = 0;
    for (const auto& kv : r->core->hypo2_clusters) t += int64_t(kv.second.size());
    if (n_members_total) *n_members_total = t;
    return int64_t(r->core->hypo2_clusters.size());
}
int ref_get_cluster_map(ref_ctx* r, int32_t* key_h1, int32_t* key_h2, int64_t* off, int32_t* members) {
    int64_t k = 0, o = 0;
    off[0] = 0;
    for (const auto& kv : r->core->hypo2_clusters) {
        key_h1[k] = kv.first.first;
        key_h2[k] = kv.first.second;
        for (int m : kv.second) members[o++] = m;
        off[++k] = o;
    }
    return 0;
}
int ref_accumulate_claims(ref_ctx* r, uint8_t** masks) {
    for (const auto& kv : r->mapping->edge_3D_to_supporting_edges)
        for (const auto& s : kv.second) masks[s.image_number][s.edge_idx] |= 1;
    return 0;
}
int64_t ref_n_3d_keys(ref_ctx* r) { return int64_t(r->mapping->edge_3D_to_supporting_edges.size()); }

// ------------------------------------------------------------------------------------- helper classes
void ref_relative_pose(const double* R1, const double* T1, const double* R2, const double* T2, double* R21, double* T21, double* R12, double* T12) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    Matrix3d r21, r12;
    Vector3d t21, t12;
    util.getRelativePoses(m3(R1), v3(T1), m3(R2), v3(T2), r21, t21, r12, t12);
    out3(r21, R21); out3(t21, T21);
    if (R12) out3(r12, R12);
    if (T12) out3(t12, T12);
}
void ref_fundamental(const double* K1, const double* K2, const double* R21, const double* T21, double* F) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    out3(util.getFundamentalMatrix(m3(K1).inverse(), m3(K2).inverse(), m3(R21), v3(T21)), F);   // EdgeSketch_Core.cpp:131
}
void ref_inverse3(const double* A, double* out) { Matrix3d inv = m3(A).inverse(); out3(inv, out); }

void ref_ore_list(int which, int n_views, const double* R, const double* T, const double* K1, const double* K2, int valid_idx, int ref_idx,
                  int n, const double* edges, double delta, int rows, int cols, double* out) {
    std::vector<Matrix3d> aR; std::vector<Vector3d> aT;
    cams(n_views, R, T, aR, aT);
    GetOrientationList::get_OrientationList ore(delta, rows, cols);
    MatrixXd e = rows_to_mat(n, 4, edges);
    MatrixXd res = which == 0 ? ore.getOreList(ref_idx, valid_idx, e, aR, aT, m3(K1), m3(K2))
                              : ore.getOreListVali(e, aR, aT, m3(K1), m3(K2), valid_idx, ref_idx);
    for (int i = 0; i < n; i++) out[i] = res(i, 0);
}
void ref_ore_list_bar(int n_views, const double* R, const double* T, const double* K1, const double* K2, int valid_idx, int ref_idx,
                      int n, const double* edges, double delta, int rows, int cols, double* out, double* epipole) {
    std::vector<Matrix3d> aR; std::vector<Vector3d> aT;
    cams(n_views, R, T, aR, aT);
    GetOrientationList::get_OrientationList ore(delta, rows, cols);
    auto res = ore.getOreListBar(rows_to_mat(n, 4, edges), aR, aT, m3(K1), m3(K2), valid_idx, ref_idx);
    for (int i = 0; i < n; i++) { out[2 * i] = res.first(i, 0); out[2 * i + 1] = res.first(i, 1); }
    if (epipole) out3(res.second, epipole);
}
int ref_hypo2_idx_ore(int n, const double* angles, double lo, double hi, int32_t* out) {
    PairEdgeHypothesis::pair_edge_hypothesis ph(2.0);
    MatrixXd a = rows_to_mat(n, 1, angles);
    MatrixXd idx = ph.getHYPO2_idx_Ore(a, std::make_pair(lo, hi));
    for (Eigen::Index i = 0; i < idx.rows(); i++) out[i] = int32_t(idx(i, 0));
    return int(idx.rows());
}
int ref_epipolar_correction(int m, const double* edgels_h2, const double* edgel_h1, const double* F21, const double* F12,
                            const double* idx_raw, double* out) {
    PairEdgeHypothesis::pair_edge_hypothesis ph(2.0);
    MatrixXd res = ph.edgelsHYPO2_epipolar_correction(rows_to_mat(m, 4, edgels_h2), rows_to_mat(1, 4, edgel_h1), m3(F21), m3(F12), rows_to_mat(m, 1, idx_raw));
    for (Eigen::Index i = 0; i < res.rows(); i++)
        for (int j = 0; j < 10; j++) out[i * 10 + j] = res(i, j);
    return int(res.rows());
}
int ref_cluster(int n, const double* h2_final, int h1_idx, int m, const double* idx_raw, int e2_rows, const double* corrected,
                int32_t* labels, double* hypo2_idx, double* averaged, int32_t* n_clusters, int32_t* key_h2, int64_t* key_off,
                int32_t* key_members, int32_t* n_keys) {
    EdgeClusterer eng(n, rows_to_mat(n, 4, h2_final), h1_idx);
    MatrixXd Edges_HYPO2 = MatrixXd::Zero(e2_rows, 4);   // performClustering reads only Edges_HYPO2.rows() (EdgeClusterer.cpp:222)
    MatrixXd hi = eng.performClustering(rows_to_mat(m, 1, idx_raw), Edges_HYPO2, rows_to_mat(n, 10, corrected));
    for (int i = 0; i < n; i++) {
        labels[i] = eng.cluster_labels[size_t(i)];
        hypo2_idx[i] = hi(i, 0);
        for (int j = 0; j < 4; j++) averaged[i * 4 + j] = eng.Epip_Correct_H2_Edges(i, j);
    }
    *n_clusters = int32_t(eng.Num_Of_Clusters);
    int k = 0;
    int64_t o = 0;
    key_off[0] = 0;
    for (const auto& kv : eng.H2_Clusters) {
        key_h2[k] = kv.first.second;
        for (int mem : kv.second) key_members[o++] = mem;
        key_off[++k] = o;
    }
    *n_keys = k;
    return 0;
}
void ref_gamma3(int n_views, const double* R, const double* T, const double* K1, const double* K2, int h1, int h2, int valid_idx,
                int n, const double* edges_h1, const double* edges_h2, double* out) {
    std::vector<Matrix3d> aR; std::vector<Vector3d> aT;
    cams(n_views, R, T, aR, aT);
    GetReprojectedEdgel::get_Reprojected_Edgel g;
    MatrixXd res = g.getGamma3LocationAndTangent(h1, h2, rows_to_mat(n, 4, edges_h1), rows_to_mat(n, 4, edges_h2), aR, aT, valid_idx, m3(K1), m3(K2));
    for (int i = 0; i < n; i++)
        for (int j = 0; j < 4; j++) out[i * 4 + j] = res(i, j);
}
int ref_support_idx(double delta_theta_deg, const double* tgt, int e3, const double* tangents, int k, const double* inliers) {
    GetSupportedEdgels::get_SupportedEdgels s(delta_theta_deg);
    return s.getSupportIdx(Eigen::Vector2d(tgt[0], tgt[1]), rows_to_mat(e3, 2, tangents), rows_to_mat(k, 1, inliers));
}
void ref_triangulate(const double* p1, const double* p2, const double* R21, const double* T21, const double* K, double* o) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    std::vector<Eigen::Vector2d> pts = {Eigen::Vector2d(p1[0], p1[1]), Eigen::Vector2d(p2[0], p2[1])};
    std::vector<Matrix3d> Rs = {m3(R21)};
    std::vector<Vector3d> Ts = {v3(T21)};
    out3(util.linearTriangulation(2, pts, Rs, Ts, m3(K)), o);   // EdgeSketch_Core.cpp:656
}
void ref_tangent3d(const double* e1, const double* e2, const double* K1, const double* K2, const double* R1, const double* T1,
                   const double* R2, const double* T2, double* o) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    out3(util.get3DTangentFromTwo2Dtangents(v3(e1), v3(e2), m3(K1), m3(K2), m3(R1), v3(T1), m3(R2), v3(T2)), o);   // EdgeSketch_Core.cpp:675
}
// ---- the helpers of the curve consolidation (edge_mapping.cpp:262, 541-575)
void ref_closest_vector_to_line(const double* P1, const double* d1, const double* P2, double* o) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    out3(util.findClosestVectorFromPointToLine(v3(P1), v3(d1), v3(P2)), o);
}
void ref_shortest_align_euler_deg(const double* a, const double* b, double* o) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    out3(util.getShortestAlignEulerAnglesDegrees(v3(a), v3(b)), o);
}
void ref_euler_to_rotation(double roll, double pitch, double yaw, double* o9) {
    MultiviewGeometryUtil::multiview_geometry_util util;
    const Eigen::Matrix3d R = util.euler_to_rotation_matrix(roll, pitch, yaw);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) o9[3 * i + j] = R(i, j);
}

}  // extern "C"
