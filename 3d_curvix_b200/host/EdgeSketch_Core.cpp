// EdgeSketch_Core.cpp -- see EdgeSketch_Core.hpp.  Every reference citation is a path under the reference root.
#include "EdgeSketch_Core.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <numeric>
#include <thread>

using curvix_host::EdgeTable;
using curvix_host::FlatYaml;
using curvix_host::HostOptions;

namespace {
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
bool file_exists(const std::string& p) {
    FILE* f = std::fopen(p.c_str(), "rb");
    if (f) std::fclose(f);
    return f != nullptr;
}
[[noreturn]] void die_file(const std::string& path) {   // LOG_FILE_ERROR + exit(1), file_reader.cpp:57-60
    std::cerr << "ERROR: file " << path << " does not exist or cannot be read" << std::endl;
    std::exit(1);
}
}  // namespace

// ---- constructor: the 22 settings of EdgeSketch_Core.cpp:37-59, then the device context
EdgeSketch_Core::EdgeSketch_Core(const FlatYaml& y, const HostOptions& opt) : opt_(opt) {
    Num_Of_OMP_Threads = y.as_int("Num_Of_OMP_Threads", 1);   // kept for the settings file's sake: there is no OpenMP region here
    hyp01_view_indx = y.as_int("Init_Hypo1_View_Index");
    hyp02_view_indx = y.as_int("Init_Hypo2_View_Index");
    Edge_Loc_Pertubation = y.as_double("delta");
    Orien_Thresh = y.as_double("delta_theta");
    Max_Num_Of_Support_Views = y.as_int("Max_Num_Of_Support_Views");
    Edge_Detection_Thresh = y.as_int("TOED_Thresh");
    Parallel_Epipolar_Line_Angle_Deg = y.as_double("Parallel_Epipolar_Line_Angle");
    Reproj_Dist_Thresh = y.as_double("Reproj_Dist_Thresh", 2.0);
    Stop_3D_Edge_Sketch_by_Ratio_Of_Claimed_Edges = y.as_double("Ratio_Of_Claimed_Edges_to_Stop");
    Max_3D_Edge_Sketch_Passes = y.as_int("Max_Num_Of_3D_Edge_Sketch_Passes");
    Dataset_Path = y.str("Dataset_Path");
    Dataset_Name = y.str("Dataset_Name");
    Scene_Name = y.str("Scene_Name");
    Num_Of_Total_Imgs = y.as_int("Total_Num_Of_Images");
    Img_Rows = y.as_int("Image_Rows");
    Img_Cols = y.as_int("Image_Cols");
    Use_Multiple_K = y.as_bool("Use_Multiple_K", false);
    fx = y.as_double("fx"); fy = y.as_double("fy"); cx = y.as_double("cx"); cy = y.as_double("cy");
    std::memset(&last_counts, 0, sizeof(last_counts));

    cvx_config cfg;
    cvx_default_config(&cfg);   // definitions.h constants (ratio test, correction thresholds, clustering, truncated PI)
    cfg.delta = Edge_Loc_Pertubation;
    cfg.delta_theta_deg = Orien_Thresh;
    cfg.parallel_angle_deg = Parallel_Epipolar_Line_Angle_Deg;
    cfg.n_views = Num_Of_Total_Imgs;
    cfg.img_rows = Img_Rows;
    cfg.img_cols = Img_Cols;
    cfg.max_support_views = Max_Num_Of_Support_Views;
    std::vector<int> devs = opt_.devices.empty() ? std::vector<int>{opt_.device} : opt_.devices;
    for (int d : devs) {
        cvx_ctx* c = nullptr;
        const int rc = cvx_create(&cfg, d, &c);
        if (rc != CVX_OK) {
            std::cerr << "ERROR: cvx_create on device " << d << " failed with status " << rc
                      << (rc == CVX_ERR_NO_DEVICE ? " (no CUDA device; this build has no CPU path)" : "") << std::endl;
            std::exit(1);
        }
        gpus_.push_back(c);
    }
    gpu_ = gpus_[0];
    if (multi()) {   // one NCCL communicator over the contexts: the id from the library, one cvx_comm_init per device, concurrently
        char id[128];
        check(cvx_comm_unique_id(id), "cvx_comm_unique_id");
        std::vector<std::thread> th;
        std::vector<int> rcs(gpus_.size(), 0);
        for (size_t d = 0; d < gpus_.size(); d++)
            th.emplace_back([&, d] { rcs[d] = cvx_comm_init(gpus_[d], int(gpus_.size()), int(d), id); });
        for (auto& t : th) t.join();
        for (size_t d = 0; d < gpus_.size(); d++)
            if (rcs[d] != CVX_OK) { std::cerr << "ERROR: cvx_comm_init failed on rank " << d << ": " << cvx_last_error(gpus_[d]) << std::endl; std::exit(1); }
    }
}

EdgeSketch_Core::~EdgeSketch_Core() {
    for (cvx_ctx* c : gpus_) if (c) cvx_destroy(c);
}

void EdgeSketch_Core::check(int rc, const char* what) const {   // the reference has no error path inside the sketch: log + exit(1)
    if (rc == CVX_OK) return;
    std::cerr << "ERROR: " << what << " failed (" << rc << "): " << cvx_last_error(gpu_) << std::endl;
    std::exit(1);
}

std::string EdgeSketch_Core::scene_dir() const { return Dataset_Path + Dataset_Name + "/" + Scene_Name; }   // file_reader.cpp:10

// ---- Read_Camera_Data, EdgeSketch_Core.cpp:95-113
void EdgeSketch_Core::Read_Camera_Data() {
    const std::string rnt = scene_dir() + "/RnT/";
    bool refined = opt_.use_refined_poses == 1;
    if (opt_.use_refined_poses < 0) refined = file_exists(rnt + "refined_R.txt") && file_exists(rnt + "refined_T.txt");   // definitions.h:11
    const std::string rp = rnt + (refined ? "refined_R.txt" : "R_matrix.txt"), tp = rnt + (refined ? "refined_T.txt" : "T_matrix.txt");
    if (!curvix_host::read_R_file(rp, &All_R)) die_file(rp);
    if (!curvix_host::read_T_file(tp, &All_T)) die_file(tp);
    if (int(All_R.size() / 9) < Num_Of_Total_Imgs || int(All_T.size() / 3) < Num_Of_Total_Imgs) {
        std::cerr << "ERROR: " << rp << " / " << tp << " hold fewer than Total_Num_Of_Images cameras" << std::endl;
        std::exit(1);
    }
    All_R.resize(size_t(Num_Of_Total_Imgs) * 9);
    All_T.resize(size_t(Num_Of_Total_Imgs) * 3);
    const double k[9] = {fx, 0, cx, 0, fy, cy, 0, 0, 1};   // EdgeSketch_Core.cpp:112
    std::memcpy(K, k, sizeof(k));
    if (Use_Multiple_K) {
        const std::string kp = rnt + "K_matrix.txt";
        if (!curvix_host::read_R_file(kp, &All_K) || int(All_K.size() / 9) < Num_Of_Total_Imgs) die_file(kp);
        All_K.resize(size_t(Num_Of_Total_Imgs) * 9);
        for (cvx_ctx* c : gpus_) check(cvx_set_cameras(c, Num_Of_Total_Imgs, All_R.data(), All_T.data(), All_K.data(), 1), "cvx_set_cameras");
    } else {
        for (cvx_ctx* c : gpus_) check(cvx_set_cameras(c, Num_Of_Total_Imgs, All_R.data(), All_T.data(), K, 0), "cvx_set_cameras");
    }
    if (!opt_.quiet) std::cout << "Camera poses: " << (refined ? "refined" : "un-refined") << " (" << rp << ")" << std::endl;
}

// ---- Set_Hypothesis_Views_Camera, EdgeSketch_Core.cpp:115-143: the matrices are derived inside cvx_set_hypothesis_pair;
// what stays on the host is the history list and the output file suffix.
void EdgeSketch_Core::Set_Hypothesis_Views_Camera() {
    history_hypothesis_views_index.push_back(hyp01_view_indx);
    history_hypothesis_views_index.push_back(hyp02_view_indx);
    Post_File_Name_Str = Dataset_Name + "_" + Scene_Name + "_hypo1_" + std::to_string(hyp01_view_indx) + "_hypo2_" +
                         std::to_string(hyp02_view_indx) + ".txt";
}

// ---- Read_Edgels_Data, EdgeSketch_Core.hpp:87-90.  The reference re-parses all N text files here every pass (and again in
// the next-view selection); the edges cannot change between passes, so they are parsed and uploaded once and stay resident.
void EdgeSketch_Core::Read_Edgels_Data() {
    if (edges_uploaded_) return;
    const double t0 = now_s();
    All_Edgels.assign(size_t(Num_Of_Total_Imgs), EdgeTable());
    for (int v = 0; v < Num_Of_Total_Imgs; v++) {
        EdgeTable& t = All_Edgels[size_t(v)];
        const std::string path = curvix_host::edge_file_path(scene_dir(), v, Edge_Detection_Thresh);
        std::string cache;
        bool have = false;
        if (!opt_.edge_cache_dir.empty()) {
            cache = opt_.edge_cache_dir + "/" + Dataset_Name + "_" + Scene_Name + "_Edge_" + std::to_string(v) + "_t" +
                    std::to_string(Edge_Detection_Thresh) + ".cvxe";
            have = curvix_host::read_edge_cache(cache, &t);
        }
        if (!have) {
            if (!curvix_host::read_edge_file(path, &t)) die_file(path);
            if (!cache.empty()) curvix_host::write_edge_cache(cache, t);
        }
        for (cvx_ctx* c : gpus_)
            check(cvx_set_view_edges(c, v, t.size(), t.x.data(), t.y.data(), t.theta.data(), t.strength.data()), "cvx_set_view_edges");
    }
    for (cvx_ctx* c : gpus_) check(cvx_claims_reset(c), "cvx_claims_reset");
    edges_uploaded_ = true;
    read_edges_time += now_s() - t0;
}

// ---- Set_Hypothesis_Views_Edgels, EdgeSketch_Core.cpp:145-166 (+ the per-view staging hoisted out of Run_3D_Edge_Sketch)
void EdgeSketch_Core::Set_Hypothesis_Views_Edgels() {
    const double t0 = now_s();
    for (cvx_ctx* c : gpus_) check(cvx_set_hypothesis_pair(c, hyp01_view_indx, hyp02_view_indx), "cvx_set_hypothesis_pair");   // asynchronous
    pair_edges_time += now_s() - t0;   // in the reference this work sits inside Run_3D_Edge_Sketch's loops
}

// ---- Run_3D_Edge_Sketch, EdgeSketch_Core.cpp:168-573
// The pass over several GPUs: the view-1 edges are independent (EdgeSketch_Core.cpp:209, disjoint paired_edge rows :515-523), so every
// device takes a contiguous range of them balanced by hypothesis rows; one host thread per context.  The accepted rows, 3-D edges and
// tangents of every range are gathered to the first device (cvx_gather_*: NCCL) in rank order = ascending view-1 edge = the
// single-GPU row order; the claimed-edge bits are OR-ed over the ranks; the cluster tables stay per rank.
void EdgeSketch_Core::run_sharded() {
    const int D = int(gpus_.size()), N = Num_Of_Total_Imgs, V = N - 2;
    shard_bounds_.assign(size_t(D) + 1, 0);
    check(cvx_split_h1_by_rows(gpu_, D, shard_bounds_.data()), "cvx_split_h1_by_rows");
    const int E1 = All_Edgels[size_t(hyp01_view_indx)].size();
    std::vector<cvx_counts> cnt(static_cast<size_t>(D));
    std::vector<std::vector<int64_t>> ord(static_cast<size_t>(D));
    rk_row_off_.assign(size_t(D), {}); rk_row_n_.assign(size_t(D), {}); rk_rawat_.assign(size_t(D), {}); rk_label_.assign(size_t(D), {});
    std::vector<std::string> err(static_cast<size_t>(D));
    auto worker = [&](int d) {
        cvx_ctx* c = gpus_[size_t(d)];
        auto ck = [&](int rc, const char* what) { if (rc != CVX_OK && err[size_t(d)].empty()) err[size_t(d)] = std::string(what) + ": " + cvx_last_error(c); return rc == CVX_OK; };
        if (!ck(cvx_run_edge_sketch(c, shard_bounds_[size_t(d)], shard_bounds_[size_t(d) + 1]), "cvx_run_edge_sketch")) return;
        ck(cvx_get_counts(c, &cnt[size_t(d)]), "cvx_get_counts");
        ck(cvx_claims_accumulate(c), "cvx_claims_accumulate");          // runs the finalize kernel; claims of this range
        if (d == 0) ck(cvx_gather_reset(c), "cvx_gather_reset");
    };
    {   // phase 1: compute
        std::vector<std::thread> th;
        for (int d = 0; d < D; d++) th.emplace_back(worker, d);
        for (auto& t : th) t.join();
    }
    for (int d = 0; d < D; d++) if (!err[size_t(d)].empty()) { std::cerr << "ERROR (device rank " << d << "): " << err[size_t(d)] << std::endl; std::exit(1); }
    {   // phase 2: the exchange (collective calls, one thread per rank) + the per-rank tables
        std::vector<std::thread> th;
        for (int d = 0; d < D; d++)
            th.emplace_back([&, d] {
                cvx_ctx* c = gpus_[size_t(d)];
                auto ck = [&](int rc, const char* what) { if (rc != CVX_OK && err[size_t(d)].empty()) err[size_t(d)] = std::string(what) + ": " + cvx_last_error(c); };
                if (d != 0) ck(cvx_gather_reset(c), "cvx_gather_reset");
                ck(cvx_gather_post(c, d, 0), "cvx_gather_post");
                ck(cvx_gather_wait(c, 1, 1), "cvx_gather_wait");
                ck(cvx_claims_allreduce(c), "cvx_claims_allreduce");
                const int64_t P = cnt[size_t(d)].n_pairs, H = cnt[size_t(d)].n_hypotheses;
                ord[size_t(d)].assign(size_t(std::max<int64_t>(P, 1)), 0);
                int64_t n_out = 0;
                ck(cvx_get_reference_order(c, ord[size_t(d)].data(), &n_out), "cvx_get_reference_order");
                ord[size_t(d)].resize(size_t(n_out));
                rk_row_off_[size_t(d)].assign(size_t(E1), -1); rk_row_n_[size_t(d)].assign(size_t(E1), 0);
                rk_rawat_[size_t(d)].assign(size_t(H), 0); rk_label_[size_t(d)].assign(size_t(H), 0);
                ck(cvx_get_h1_table(c, nullptr, rk_row_n_[size_t(d)].data(), rk_row_off_[size_t(d)].data()), "cvx_get_h1_table");
                ck(cvx_get_rows(c, nullptr, nullptr, nullptr, rk_rawat_[size_t(d)].data(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                rk_label_[size_t(d)].data(), nullptr, nullptr), "cvx_get_rows");
            });
        for (auto& t : th) t.join();
    }
    for (int d = 0; d < D; d++) if (!err[size_t(d)].empty()) { std::cerr << "ERROR (device rank " << d << "): " << err[size_t(d)] << std::endl; std::exit(1); }
    // rank 0: the gathered units, concatenated in rank order
    std::memset(&last_counts, 0, sizeof(last_counts));
    int64_t Ptot = 0;
    for (int d = 0; d < D; d++) {
        const cvx_counts& c = cnt[size_t(d)];
        last_counts.n_h1_total = c.n_h1_total; last_counts.n_h1_processed += c.n_h1_processed; last_counts.n_candidates += c.n_candidates;
        last_counts.n_hypotheses += c.n_hypotheses; last_counts.n_validations += c.n_validations; last_counts.n_valid_rows += c.n_valid_rows;
        last_counts.n_pairs += c.n_pairs; last_counts.n_degenerate_wedges += c.n_degenerate_wedges; last_counts.n_supported += c.n_supported;
        last_counts.n_parallel += c.n_parallel; last_counts.n_inliers += c.n_inliers; last_counts.n_unique_hypotheses += c.n_unique_hypotheses;
        Ptot += c.n_pairs;
    }
    std::vector<int32_t> h1(size_t(Ptot), 0), h2(size_t(Ptot), 0);
    std::vector<int64_t> soff;
    std::vector<int32_t> sent;
    mg_gc_.assign(size_t(Ptot) * 3, 0.0); mg_gw_.assign(size_t(Ptot) * 3, 0.0); mg_tw_.assign(size_t(Ptot) * 3, 0.0); mg_h2c_.assign(size_t(Ptot) * 2, 0.0);
    mg_nf_.assign(size_t(Ptot), 0);
    std::vector<int32_t> sup(size_t(Ptot) * size_t(V), -2);
    order_.clear();
    int64_t base = 0;
    for (int d = 0; d < D; d++) {
        cvx_gathered_unit u;
        check(cvx_gather_get(gpu_, 0, d, &u), "cvx_gather_get");
        const int64_t P = u.meta[0], S = u.meta[7];
        if (P != cnt[size_t(d)].n_pairs) { std::cerr << "ERROR: gathered unit of rank " << d << " has " << P << " rows, expected " << cnt[size_t(d)].n_pairs << std::endl; std::exit(1); }
        if (P > 0) {
            soff.assign(size_t(P) + 1, 0); sent.assign(size_t(S) * 2, 0);
            check(cvx_gather_fetch(gpu_, 0, d, h1.data() + base, h2.data() + base, nullptr, mg_nf_.data() + base, soff.data(), sent.data(),
                                   mg_gw_.data() + base * 3, mg_tw_.data() + base * 3, mg_gc_.data() + base * 3, mg_h2c_.data() + base * 2), "cvx_gather_fetch");
            for (int64_t r = 0; r < P; r++)
                for (int64_t e = soff[size_t(r)]; e < soff[size_t(r) + 1]; e++)
                    sup[size_t(base + r) * size_t(V) + size_t(sent[size_t(e) * 2])] = sent[size_t(e) * 2 + 1];
        }
        for (int64_t i : ord[size_t(d)]) order_.push_back(base + i);
        base += P;
    }
    const int64_t n_out = int64_t(order_.size());
    paired_edge_final = Matrix(n_out, N, -2.0);   // :596-601
    for (int64_t i = 0; i < n_out; i++) {
        const int64_t p = order_[size_t(i)];
        paired_edge_final(i, 0) = h1[size_t(p)];
        paired_edge_final(i, 1) = h2[size_t(p)];
        for (int v = 0; v < V; v++) paired_edge_final(i, 2 + v) = sup[size_t(p) * size_t(V) + size_t(v)];
    }
}

void EdgeSketch_Core::Run_3D_Edge_Sketch() {
    const double t0 = now_s();
    if (multi()) { run_sharded(); pair_edges_time += now_s() - t0; return; }
    check(cvx_run_edge_sketch(gpu_, 0, -1), "cvx_run_edge_sketch");
    check(cvx_get_counts(gpu_, &last_counts), "cvx_get_counts");
    const int N = Num_Of_Total_Imgs, V = N - 2;
    const int64_t P = last_counts.n_pairs;
    const size_t Ps = static_cast<size_t>(P);
    std::vector<int32_t> h1(Ps, 0), h2(Ps, 0), sup(Ps * size_t(V), 0);
    check(cvx_get_pairs(gpu_, h1.data(), h2.data(), nullptr, nullptr, sup.data()), "cvx_get_pairs");
    // the reference's per-edge std::sort by support count + capacity of N rows per view-1 edge (:480-526)
    order_.assign(size_t(std::max<int64_t>(P, 1)), 0);
    int64_t n_out = 0;
    check(cvx_get_reference_order(gpu_, order_.data(), &n_out), "cvx_get_reference_order");
    order_.resize(size_t(n_out));
    paired_edge_final = Matrix(n_out, N, -2.0);   // :596-601
    for (int64_t i = 0; i < n_out; i++) {
        const int64_t p = order_[size_t(i)];
        paired_edge_final(i, 0) = h1[size_t(p)];
        paired_edge_final(i, 1) = h2[size_t(p)];
        for (int v = 0; v < V; v++) paired_edge_final(i, 2 + v) = sup[size_t(p) * size_t(V) + size_t(v)];
    }
    // hypo2_clusters (hpp:47-48, merged at :559-566): kept as the hypothesis-table columns it is built from
    const int E1 = All_Edgels[size_t(hyp01_view_indx)].size();
    const int64_t H = last_counts.n_hypotheses;
    row_off_.assign(size_t(E1), -1); row_n_.assign(size_t(E1), 0);
    raw_at_local_.assign(size_t(H), 0); label_.assign(size_t(H), 0);
    check(cvx_get_h1_table(gpu_, nullptr, row_n_.data(), row_off_.data()), "cvx_get_h1_table");
    check(cvx_get_rows(gpu_, nullptr, nullptr, nullptr, raw_at_local_.data(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                       label_.data(), nullptr, nullptr), "cvx_get_rows");
    pair_edges_time += now_s() - t0;
}

std::vector<int> EdgeSketch_Core::get_edges_in_same_cluster(int hypo1_edge, int hypo2_edge) const {
    // key = (view-1 edge, HYPO2_idx_raw(local row)), value = HYPO2_idx_raw of every row of that cluster (EdgeClusterer.cpp:216-236)
    const std::vector<int64_t>* off = &row_off_;
    const std::vector<int32_t>*rn = &row_n_, *raw = &raw_at_local_, *lab = &label_;
    if (multi()) {   // the tables of the rank that owns this view-1 edge
        size_t d = 0;
        while (d + 1 < gpus_.size() && hypo1_edge >= shard_bounds_[d + 1]) d++;
        off = &rk_row_off_[d]; rn = &rk_row_n_[d]; raw = &rk_rawat_[d]; lab = &rk_label_[d];
    }
    if (hypo1_edge >= 0 && size_t(hypo1_edge) < off->size() && (*off)[size_t(hypo1_edge)] >= 0) {
        const int64_t r0 = (*off)[size_t(hypo1_edge)], r1 = r0 + (*rn)[size_t(hypo1_edge)];
        for (int64_t r = r0; r < r1; r++) {
            if ((*raw)[size_t(r)] != hypo2_edge) continue;
            std::vector<int> members;
            for (int64_t q = r0; q < r1; q++)
                if ((*lab)[size_t(q)] == (*lab)[size_t(r)]) members.push_back((*raw)[size_t(q)]);
            return members;
        }
    }
    return {hypo2_edge};   // hpp:200-201
}

// ---- Finalize_Edge_Pairs_and_Reconstruct_3D_Edges, EdgeSketch_Core.cpp:575-733
void EdgeSketch_Core::Finalize_Edge_Pairs_and_Reconstruct_3D_Edges(std::shared_ptr<EdgeMapping> edgeMapping) {
    const double t0 = now_s();
    const int N = Num_Of_Total_Imgs;
    const int64_t P = last_counts.n_pairs, n_rows = paired_edge_final.rows;
    if (!opt_.quiet) std::cout << "Number of edge pairs: " << n_rows << std::endl;   // :603-604
    const size_t Ps = static_cast<size_t>(P);
    std::vector<double> gc(Ps * 3, 0.0), gw(Ps * 3, 0.0), tw(Ps * 3, 0.0), h2c(Ps * 2, 0.0);
    std::vector<int32_t> nf(Ps, 0);
    if (multi()) {   // finalize kernels and claims ran on every rank inside the pass; the geometry came with the gather
        gc = mg_gc_; gw = mg_gw_; tw = mg_tw_; h2c = mg_h2c_; nf = mg_nf_;
    } else {
        check(cvx_finalize(gpu_, gc.data(), gw.data(), tw.data(), h2c.data(), nf.data()), "cvx_finalize");
        check(cvx_claims_accumulate(gpu_), "cvx_claims_accumulate");   // device-side counterpart of the insertions below (:682-726)
    }
    const EdgeTable& E1 = All_Edgels[size_t(hyp01_view_indx)];
    const EdgeTable& E2 = All_Edgels[size_t(hyp02_view_indx)];
    Gamma1s = Matrix(0, 3); Gamma1s_world = Matrix(0, 3); tangent3Ds = Matrix(0, 3);
    Gamma1s.d.reserve(size_t(n_rows) * 3); Gamma1s_world.d.reserve(size_t(n_rows) * 3); tangent3Ds.d.reserve(size_t(n_rows) * 3);
    valid_rows_of_pair_.clear();
    for (int64_t i = 0; i < n_rows; i++) {
        const int64_t p = order_[size_t(i)];
        const int f = nf[size_t(p)];
        if (f & 2) { std::cout << "No valid matches found for edge " << i << std::endl; continue; }          // :639-642
        if (f & 1) { std::cerr << "NaN values detected in edge_pt_3D for pair_idx: " << i << std::endl; continue; }   // :658-663
        const int a = int(paired_edge_final(i, 0)), b = int(paired_edge_final(i, 1));
        const double* G = &gc[size_t(p) * 3]; const double* W = &gw[size_t(p) * 3]; const double* T = &tw[size_t(p) * 3];
        Gamma1s.d.insert(Gamma1s.d.end(), G, G + 3); Gamma1s.rows++;
        Gamma1s_world.d.insert(Gamma1s_world.d.end(), W, W + 3); Gamma1s_world.rows++;
        tangent3Ds.d.insert(tangent3Ds.d.end(), T, T + 3); tangent3Ds.rows++;
        valid_rows_of_pair_.push_back(p);
        if (!edgeMapping) continue;
        // :678-681 the two hypothesis edges
        const double pt1[2] = {E1.x[size_t(a)], E1.y[size_t(a)]};
        const double un1[3] = {E1.x[size_t(a)], E1.y[size_t(a)], E1.theta[size_t(a)]};
        const double pt2[2] = {h2c[size_t(p) * 2], h2c[size_t(p) * 2 + 1]};
        const double un2[3] = {E2.x[size_t(b)], E2.y[size_t(b)], E2.theta[size_t(b)]};
        edgeMapping->add3DToSupportingEdgesMapping(W, T, pt1, un1, hyp01_view_indx, a);
        edgeMapping->add3DToSupportingEdgesMapping(W, T, pt2, un2, hyp02_view_indx, b);
        // :684-692 the other members of the representative's cluster
        for (int m : get_edges_in_same_cluster(a, b)) {
            if (m < 0 || m >= E2.size() || m == b) continue;
            const double unm[3] = {E2.x[size_t(m)], E2.y[size_t(m)], E2.theta[size_t(m)]};
            edgeMapping->add3DToSupportingEdgesMapping(W, T, pt2, unm, hyp02_view_indx, m);
        }
        // :694-724 the supporting edge of every validation view
        int col = 2;
        for (int v = 0; v < N; v++) {
            if (v == hyp01_view_indx || v == hyp02_view_indx) continue;
            const int s = int(paired_edge_final(i, col));
            if (s != -2) {
                const EdgeTable& Ev = All_Edgels[size_t(v)];
                const double ptv[2] = {Ev.x[size_t(s)], Ev.y[size_t(s)]};
                const double unv[3] = {Ev.x[size_t(s)], Ev.y[size_t(s)], Ev.theta[size_t(s)]};
                edgeMapping->add3DToSupportingEdgesMapping(W, T, ptv, unv, v, s);
            }
            col++;
        }
    }
    finalize_edge_pair_time += now_s() - t0;
}

// ---- Stack_3D_Edges, EdgeSketch_Core.cpp:735-774.  The camera -> world transform R1^T (Gamma - T1) (:742-748) is done by the
// finalize kernel; what remains is the two `ofstream << MatrixXd` writes and the concatenation.
void EdgeSketch_Core::Stack_3D_Edges() {
    const double t0 = now_s();
    if (opt_.write_3d_edges) {
        const std::string p_edges = opt_.outputs_dir + "/3D_edges_" + Post_File_Name_Str;
        const std::string p_tgts = opt_.outputs_dir + "/3D_tangents_" + Post_File_Name_Str;
        if (!opt_.quiet) std::cout << "Writing 3D edge locations to: " << p_edges << std::endl;
        if (!curvix_host::write_eigen_matrix(p_edges, Gamma1s_world.d.data(), Gamma1s_world.rows, 3)) die_file(p_edges);
        if (!opt_.quiet) std::cout << "Writing 3D edge tangents to: " << p_tgts << std::endl;
        if (!curvix_host::write_eigen_matrix(p_tgts, tangent3Ds.d.data(), tangent3Ds.rows, 3)) die_file(p_tgts);
    }
    all_3D_Edges.append_rows(Gamma1s_world);   // :766-773
    stack_time += now_s() - t0;                // untimed in the reference; part of BASELINE.json's seconds-per-object metric
}

// ---- Calculate_Edge_Support_Ratios_And_Select_Next_Views, EdgeSketch_Core.cpp:776-871
void EdgeSketch_Core::Calculate_Edge_Support_Ratios_And_Select_Next_Views(std::shared_ptr<EdgeMapping> /*edgeMapping*/) {
    const double t0 = now_s();
    const int N = Num_Of_Total_Imgs;
    // :779-797 -- the reference re-reads every edge file and walks the 3-D -> 2-D support map into one std::set per view; the
    // same sets are the claim bits the context accumulated after every pass
    std::vector<int64_t> claimed(static_cast<size_t>(N), 0), total(static_cast<size_t>(N), 0);
    check(cvx_claims_get_counts(gpu_, claimed.data(), total.data()), "cvx_claims_get_counts");
    std::vector<double> ratios;
    std::vector<std::pair<int, double>> frameSupportCounts;
    for (int i = 0; i < N; i++) {   // :802-811
        const double ratio = total[size_t(i)] > 0 ? double(claimed[size_t(i)]) / double(total[size_t(i)]) : 0.0;   // 0/0 is NaN there
        ratios.push_back(ratio);
        frameSupportCounts.push_back({i, ratio});
    }
    avg_ratio = std::accumulate(ratios.begin(), ratios.end(), 0.0) / double(ratios.size());   // :813
    if (!opt_.quiet) std::cout << "Average ratio of claimed edges: " << avg_ratio << std::endl;
    // :818-821 literally std::sort (not stable): ties between equal ratios resolve as libstdc++'s introsort resolves them
    std::sort(frameSupportCounts.begin(), frameSupportCounts.end(),
              [](const std::pair<int, double>& a, const std::pair<int, double>& b) { return a.second < b.second; });
    auto used = [&](int v) { return std::count(history_hypothesis_views_index.begin(), history_hypothesis_views_index.end(), v) > 0; };
    auto at = [&](size_t k, int keep) { return k < frameSupportCounts.size() ? frameSupportCounts[k].first : keep; };   // the reference reads past the end here
    int best1 = frameSupportCounts[0].first, best2 = frameSupportCounts[1].first;
    size_t k = 0;
    while (true) {   // :828-854
        const bool u1 = used(best1), u2 = used(best2);
        if (!u1 && !u2) break;
        if (u1 && u2) { best1 = at(2 + k, best1); best2 = at(3 + k, best2); }
        else if (u1) { best1 = at(1 + k, best1); best2 = at(2 + k, best2); }
        else { best2 = at(2 + k, best2); }
        k++;
        if (k > frameSupportCounts.size() - 2) {
            std::cout << "Warning: Unable to find unused views. Reusing some previous views." << std::endl;
            break;
        }
    }
    hyp01_view_indx = best1;
    hyp02_view_indx = best2;
    if (!opt_.quiet) std::cout << "Selected next hypothesis views: " << best1 << " and " << best2 << std::endl;
    enable_aborting_3D_edge_sketch = avg_ratio > Stop_3D_Edge_Sketch_by_Ratio_Of_Claimed_Edges;   // :862-868
    if (enable_aborting_3D_edge_sketch && !opt_.quiet) std::cout << "Reached Stop_3D_Edge_Sketch_by_Ratio_Of_Claimed_Edges value" << std::endl;
    find_next_hypothesis_view_time += now_s() - t0;
}

// ---- Clear_Data, EdgeSketch_Core.cpp:1013-1033: the per-pass host copies; device buffers are reused, edges stay resident
void EdgeSketch_Core::Clear_Data() {
    order_.clear(); row_off_.clear(); row_n_.clear(); raw_at_local_.clear(); label_.clear(); valid_rows_of_pair_.clear();
}
