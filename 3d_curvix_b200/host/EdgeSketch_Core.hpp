// EdgeSketch_Core.hpp -- host-side drop-in for the reference class of the same name
// (Edge_Reconst/EdgeSketch_Core.{hpp,cpp} as driven by cmd/main.cpp:114-170).
//
// Same public method names, call order and public state as the reference class, so a pass loop written against the
// reference compiles against this one.  The three hot methods (Run_3D_Edge_Sketch,
// Finalize_Edge_Pairs_and_Reconstruct_3D_Edges, Stack_3D_Edges) and the claimed-edge bookkeeping of
// Calculate_Edge_Support_Ratios_And_Select_Next_Views go through the C-ABI of include/curvix_b200.h into the sm_100a
// kernels; there is no CPU implementation of the path in this class.  Eigen and yaml-cpp are not available in this image,
// so matrices are a small row-major container and the settings come from a flat-YAML reader.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "curvix_b200.h"
#include "dataset_io.hpp"
#include "edge_mapping.hpp"
#include "flat_yaml.hpp"

namespace curvix_host {

// rows x cols doubles, row-major; m(i, j) like an Eigen::MatrixXd
struct Matrix {
    int64_t rows = 0;
    int cols = 0;
    std::vector<double> d;
    Matrix() = default;
    Matrix(int64_t r, int c, double fill = 0.0) : rows(r), cols(c), d(size_t(r) * size_t(c), fill) {}
    double& operator()(int64_t i, int j) { return d[size_t(i) * size_t(cols) + size_t(j)]; }
    double operator()(int64_t i, int j) const { return d[size_t(i) * size_t(cols) + size_t(j)]; }
    const double* row(int64_t i) const { return d.data() + size_t(i) * size_t(cols); }
    void append_rows(const Matrix& o) {
        if (rows == 0) cols = o.cols;
        d.insert(d.end(), o.d.begin(), o.d.end());
        rows += o.rows;
    }
};

struct HostOptions {
    int device = 0;
    std::vector<int> devices;                    // more than one: every pass is sharded by view-1 edge range over these GPUs (one context and
                                                 // one host thread each), results gathered to the first with cvx_gather_* (NCCL)
    std::string outputs_dir = "../../outputs";   // OUTPUT_FOLDER_NAME, definitions.h:4
    std::string edge_cache_dir;                  // optional binary SoA cache of the parsed edge files
    int use_refined_poses = -1;                  // USE_REFINED_CAM_POSES (definitions.h:11): 1 / 0, -1 = refined when the files exist
    bool write_3d_edges = true;                  // WRITE_3D_EDGES (definitions.h)
    bool quiet = false;
};

}  // namespace curvix_host

class EdgeSketch_Core {
public:
    using Matrix = curvix_host::Matrix;

    EdgeSketch_Core(const curvix_host::FlatYaml& settings, const curvix_host::HostOptions& opt = curvix_host::HostOptions());
    ~EdgeSketch_Core();
    EdgeSketch_Core(const EdgeSketch_Core&) = delete;
    EdgeSketch_Core& operator=(const EdgeSketch_Core&) = delete;

    // ---- the reference's pass interface (EdgeSketch_Core.hpp:68-78,87-90)
    void Read_Camera_Data();
    void Set_Hypothesis_Views_Camera();
    void Read_Edgels_Data();
    void Set_Hypothesis_Views_Edgels();
    void Run_3D_Edge_Sketch();
    void Finalize_Edge_Pairs_and_Reconstruct_3D_Edges(std::shared_ptr<EdgeMapping> edgeMapping);
    void Stack_3D_Edges();
    void Calculate_Edge_Support_Ratios_And_Select_Next_Views(std::shared_ptr<EdgeMapping> edgeMapping);
    void Clear_Data();

    // members of the cluster of an accepted (view-1 edge, view-2 edge) pair, as hypo2_clusters answers it (hpp:193-202)
    std::vector<int> get_edges_in_same_cluster(int hypo1_edge, int hypo2_edge) const;

    // ---- public state the reference's callers read (cmd/main.cpp:123-170)
    int hyp01_view_indx = 0, hyp02_view_indx = 0;
    int Max_3D_Edge_Sketch_Passes = 0;
    bool enable_aborting_3D_edge_sketch = false;
    int Num_Of_Total_Imgs = 0;
    std::vector<double> All_R;   // n x 9 row-major (world -> camera)
    std::vector<double> All_T;   // n x 3
    double K[9] = {0};           // single intrinsic matrix (Use_Multiple_K = false)
    std::vector<double> All_K;   // n x 9 when Use_Multiple_K
    double pair_edges_time = 0.0, finalize_edge_pair_time = 0.0, find_next_hypothesis_view_time = 0.0;
    double read_edges_time = 0.0, stack_time = 0.0;   // not in the reference's total (file reading is untimed there)
    double avg_ratio = 0.0;
    std::vector<int> history_hypothesis_views_index;

    // ---- produced state (EdgeSketch_Core.hpp:44-48,130,160-164)
    Matrix paired_edge_final;    // P x N: view-1 edge, view-2 edge, support index per validation view (ascending view), -2 = none
    Matrix Gamma1s;              // P x 3, camera-1 frame
    Matrix Gamma1s_world;        // P x 3, world frame (what Stack_3D_Edges writes)
    Matrix tangent3Ds;           // P x 3, world frame
    Matrix all_3D_Edges;         // all passes, world frame
    std::vector<curvix_host::EdgeTable> All_Edgels;
    std::string Post_File_Name_Str;
    cvx_counts last_counts;      // statistics of the last pass (hypotheses, validations, ...)

    // settings (names as in the reference's YAML / members)
    int Num_Of_OMP_Threads = 1, Max_Num_Of_Support_Views = 4, Edge_Detection_Thresh = 1, Img_Rows = 0, Img_Cols = 0;
    double Edge_Loc_Pertubation = 0.4, Orien_Thresh = 15, Parallel_Epipolar_Line_Angle_Deg = 15, Reproj_Dist_Thresh = 2;
    double Stop_3D_Edge_Sketch_by_Ratio_Of_Claimed_Edges = 0.7;
    double fx = 0, fy = 0, cx = 0, cy = 0;
    bool Use_Multiple_K = false;
    std::string Dataset_Path, Dataset_Name, Scene_Name;

    cvx_ctx* gpu() { return gpu_; }

private:
    void check(int rc, const char* what) const;
    std::string scene_dir() const;

    curvix_host::HostOptions opt_;
    cvx_ctx* gpu_ = nullptr;                     // the first device: where results are gathered and the claims are read
    std::vector<cvx_ctx*> gpus_;                 // all devices (gpus_[0] == gpu_)
    bool multi() const { return gpus_.size() > 1; }
    void run_sharded();                          // Run_3D_Edge_Sketch over gpus_
    bool edges_uploaded_ = false;
    // multi-device pass: what rank 0 received (rank order = ascending view-1 edge), and every rank's cluster tables
    std::vector<double> mg_gc_, mg_gw_, mg_tw_, mg_h2c_;
    std::vector<int32_t> mg_nf_;
    std::vector<int32_t> shard_bounds_;
    std::vector<std::vector<int64_t>> rk_row_off_;
    std::vector<std::vector<int32_t>> rk_row_n_, rk_rawat_, rk_label_;
    // per-pass host copies used by Finalize (EdgeMapping insertions) and the cluster lookup
    std::vector<int64_t> order_;                 // device pair index of every paired_edge_final row
    std::vector<int64_t> row_off_;               // first hypothesis row of every view-1 edge (-1: none)
    std::vector<int32_t> row_n_, raw_at_local_, label_;
    std::vector<int64_t> valid_rows_of_pair_;    // device pair indices of the rows written by Finalize (Gamma1s order)
};
