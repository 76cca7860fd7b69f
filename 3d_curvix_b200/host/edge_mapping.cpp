// edge_mapping.cpp -- EdgeMapping::consolidate_3D_edges of the drop-in (see edge_mapping.hpp).  Reference: Edge_Reconst/edge_mapping.cpp.
#include "edge_mapping.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <iostream>
#include <limits>
#include <set>
#include <unordered_set>

#include "../../include/curvix_b200.h"
#include "curve_io.hpp"
#include "dataset_io.hpp"
#include "flat_yaml.hpp"

namespace {

using curvix_host::append_eigen_row3;

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct V3 { double x, y, z; };
inline V3 v3(const double* p) { return V3{p[0], p[1], p[2]}; }
inline V3 sub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 neg(V3 a) { return V3{-a.x, -a.y, -a.z}; }

// `ostream << double` at the default precision
inline void put_g(std::string* s, double v) {
    char t[40];
    s->append(t, size_t(std::snprintf(t, sizeof t, "%.6g", v)));
}
inline bool flush_to(const std::string& path, const std::string& text) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fwrite(text.data(), 1, text.size(), f);
    return std::fclose(f) == 0;
}

// a link of the 3-D edge graph: the two node indices with `first` the lexicographically smaller POINT (edge_mapping.cpp:174-178), hashed as
// HashEigenVector3dPair does (reference edge_mapping.hpp:48-52); equality on the node pair (the reference's fuzzy equality can only merge
// two keys that already share a bucket, i.e. the same two points)
struct LinkKey {
    int first, second;
    size_t hash;
    bool operator==(const LinkKey& o) const { return (first == o.first && second == o.second) || (first == o.second && second == o.first); }
};
struct LinkHash { size_t operator()(const LinkKey& k) const { return k.hash; } };
using LinkMap = std::unordered_map<LinkKey, int, LinkHash>;

struct Curve {                      // reference edge_mapping.hpp:118-129
    std::vector<int> edge_indices;
    bool b_loops_back_on_left = false, b_loops_back_on_right = false;
    int to_be_merged_left_edge_index = -1, to_be_merged_right_edge_index = -1;
    int to_be_merged_left_curve_index = -1, to_be_merged_right_curve_index = -1;
    int consolidation_set_from_left = -1, consolidation_set_from_right = -1;
};
struct GraphNode {                  // reference ConnectivityGraphNode, edge_mapping.hpp:106-113
    V3 location, orientation;
    int left = -1, right = -1;
    V3 left_orientation{0, 0, 0}, right_orientation{0, 0, 0};
    int curve_index = -1;
};

bool in_first_or_last_two(const std::vector<int>& v, int num) {   // edge_mapping.cpp:1189-1210
    if (v.empty()) return false;
    if (v.size() == 1) return v[0] == num;
    if (v[0] == num || v[1] == num) return true;
    return v[v.size() - 1] == num || v[v.size() - 2] == num;
}

}  // namespace

void EdgeMapping::Setup_Data_Parameters(const curvix_host::FlatYaml& y) {
    thresh_EDG = y.as_int("TOED_Thresh");
    Dataset_Path = y.str("Dataset_Path");
    Dataset_Name = y.str("Dataset_Name");
    Scene_Name = y.str("Scene_Name");
    Num_Of_Images = y.as_int("Total_Num_Of_Images");
}

bool EdgeMapping::consolidate_3D_edges(cvx_ctx* gpu, const std::vector<double>& all_R, const std::vector<double>& all_T, const double K[9],
                                       int Num_Of_Total_Imgs, const std::string& outputs_dir, bool write_edge_graphs, bool quiet) {
    const double t_begin = now_s();
    report = ConsolidationReport();
    auto fail = [&](const std::string& m) { last_error = m; return false; };
    if (!gpu) return fail("consolidate_3D_edges: no device context (there is no CPU path)");
    const int N = Num_Of_Total_Imgs;
    const std::string scene_dir = Dataset_Path + Dataset_Name + "/" + Scene_Name;

    // ---- nodes in the iteration order of the reference's edge_3D_to_supporting_edges (buildEdgeNodeGraph, edge_mapping.cpp:420-438)
    const std::vector<size_t> order = reference_order();
    const int64_t n = int64_t(order.size());
    std::vector<double> loc(size_t(n) * 3), tan0(size_t(n) * 3), ori(size_t(n) * 3);
    for (int64_t i = 0; i < n; i++) {
        const Edge3D& e = edge_3D_to_supporting_edges[order[size_t(i)]];
        double sum[3] = {0, 0, 0};
        for (const auto& s : e.supports) for (int k = 0; k < 3; k++) sum[k] += s.tangents_3D_world[k];             // :427-430
        const double z = (sum[0] * sum[0] + sum[1] * sum[1]) + sum[2] * sum[2];
        if (!e.supports.empty() && z > 0) { const double d = std::sqrt(z); for (int k = 0; k < 3; k++) sum[k] /= d; }   // normalize(), :431-433
        for (int k = 0; k < 3; k++) {
            loc[size_t(i) * 3 + k] = e.point[k];
            tan0[size_t(i) * 3 + k] = e.supports.empty() ? 0.0 : e.supports.front().tangents_3D_world[k];
            ori[size_t(i) * 3 + k] = sum[k];
        }
    }
    // ---- 2-D edge -> the 3-D edges it supports (map_Uncorrected2DEdge_To_SupportingData, :36-62): records in node order x support order
    std::vector<int64_t> view_off(size_t(N) + 1, 0);
    std::vector<std::vector<int32_t>> cv_triples(static_cast<size_t>(N));
    {
        std::vector<int> n_edges(size_t(N), 0);
        for (int v = 0; v < N; v++) {                                      // read_edgels (:89-104): only the row counts are needed here
            curvix_host::EdgeTable t;
            if (!curvix_host::read_edge_file(curvix_host::edge_file_path(scene_dir, v, thresh_EDG), &t))
                return fail("cannot read " + curvix_host::edge_file_path(scene_dir, v, thresh_EDG));
            n_edges[size_t(v)] = t.size();
        }
        for (int v = 0; v < N; v++) view_off[size_t(v) + 1] = view_off[size_t(v)] + n_edges[size_t(v)];
        for (int v = 0; v < N; v++) {                                      // read_curvelets (file_reader.cpp:81-110): the int stream in triples
            const std::string p = scene_dir + "/curvelets/Curvelets_" + std::to_string(v) + "_t" + std::to_string(thresh_EDG) + ".txt";
            std::vector<double> raw;
            if (!curvix_host::read_doubles(p, &raw)) return fail("cannot read " + p);
            for (size_t k = 0; k + 3 <= raw.size(); k += 3) { cv_triples[size_t(v)].push_back(int32_t(raw[k])); cv_triples[size_t(v)].push_back(int32_t(raw[k + 1])); }
        }
    }
    const int64_t n_keys = view_off[size_t(N)];
    std::vector<int64_t> key_off(size_t(n_keys) + 1, 0);
    int64_t n_rec = 0;
    for (int64_t i = 0; i < n; i++)
        for (const auto& s : edge_3D_to_supporting_edges[order[size_t(i)]].supports) {
            if (s.image_number < 0 || s.image_number >= N || s.edge_idx < 0 || s.edge_idx >= view_off[size_t(s.image_number) + 1] - view_off[size_t(s.image_number)])
                return fail("a supporting edge names a view / edge outside the dataset");
            key_off[size_t(view_off[size_t(s.image_number)] + s.edge_idx) + 1]++;
            n_rec++;
        }
    for (int64_t k = 0; k < n_keys; k++) key_off[size_t(k) + 1] += key_off[size_t(k)];
    std::vector<int32_t> key_nodes(size_t(std::max<int64_t>(n_rec, 1)));
    {
        std::vector<int64_t> fill(key_off.begin(), key_off.end() - 1);
        for (int64_t i = 0; i < n; i++)
            for (const auto& s : edge_3D_to_supporting_edges[order[size_t(i)]].supports)
                key_nodes[size_t(fill[size_t(view_off[size_t(s.image_number)] + s.edge_idx)]++)] = int32_t(i);
    }
    std::vector<int32_t> cv_self, cv_nb;
    for (int v = 0; v < N; v++) {
        const int64_t ne = view_off[size_t(v) + 1] - view_off[size_t(v)];
        const auto& t = cv_triples[size_t(v)];
        for (size_t k = 0; k + 2 <= t.size(); k += 2) {
            const int a = t[k], b = t[k + 1];
            cv_self.push_back(a >= 0 && a < ne ? int32_t(view_off[size_t(v)] + a) : -1);
            cv_nb.push_back(b >= 0 && b < ne ? int32_t(view_off[size_t(v)] + b) : -1);
        }
    }
    report.nodes = n; report.records = n_rec; report.curvelets = int64_t(cv_self.size());

    // ---- device: links, prunings (cvx_graph_*)
    cvx_graph* g = nullptr;
    auto ck = [&](int rc, const char* what) { if (rc != CVX_OK) { last_error = std::string(what) + ": " + cvx_last_error(gpu); if (g) cvx_graph_destroy(g); g = nullptr; return false; } return true; };
    if (!ck(cvx_graph_create(gpu, n, loc.data(), tan0.data(), &g), "cvx_graph_create")) return false;
    int64_t L0 = 0, L1 = 0, L2 = 0;
    if (!ck(cvx_graph_build_links(g, n_keys, key_off.data(), key_nodes.data(), int64_t(cv_self.size()), cv_self.data(), cv_nb.data(), &L0), "cvx_graph_build_links")) return false;
    if (!ck(cvx_graph_prune(g, prune_lambda1, prune_lambda2, prune_lambda3, prune_by_proj_prox_thresh, N, all_R.data(), all_T.data(), K, report.stats, &L1, &L2), "cvx_graph_prune")) return false;
    std::vector<int32_t> la(size_t(std::max<int64_t>(L0, 1))), lb(size_t(std::max<int64_t>(L0, 1)));
    std::vector<int64_t> first_pos(size_t(std::max<int64_t>(L0, 1)));
    std::vector<uint8_t> keep3d(size_t(std::max<int64_t>(L0, 1))), keepproj(size_t(std::max<int64_t>(L0, 1)));
    if (!ck(cvx_graph_get_links(g, 0, la.data(), lb.data(), nullptr), "cvx_graph_get_links")) return false;
    if (!ck(cvx_graph_get_link_details(g, first_pos.data(), keep3d.data(), keepproj.data()), "cvx_graph_get_link_details")) return false;
    report.links = L0; report.links_after_3d = L1; report.links_after_projection = L2;
    if (!quiet) {
        std::cout << "[GRAPH STATS BEFORE PRUNING]\n";                                                        // :291-294
        std::cout << "mu_t     (meters) : " << report.stats[0] << ", sigma_t:     " << report.stats[1] << "\n";
        std::cout << "mu_n     (meters) : " << report.stats[2] << ", sigma_n:     " << report.stats[3] << "\n";
        std::cout << "mu_theta (degrees): " << report.stats[4] << ", sigma_theta: " << report.stats[5] << "\n";
    }
    // ---- the three graphs as the reference holds them: unordered_maps filled in its insertion order, read in their iteration order
    auto make_key = [&](int a, int b) {
        const double *pa = &loc[size_t(a) * 3], *pb = &loc[size_t(b) * 3];
        const bool swap = (pa[0] > pb[0]) || (pa[0] == pb[0] && pa[1] > pb[1]) || (pa[0] == pb[0] && pa[1] == pb[1] && pa[2] > pb[2]);   // :174-178
        LinkKey k;
        k.first = swap ? b : a; k.second = swap ? a : b;
        k.hash = hash_point(&loc[size_t(k.first) * 3]) ^ (hash_point(&loc[size_t(k.second) * 3]) << 1);
        return k;
    };
    std::vector<int64_t> by_pos(static_cast<size_t>(L0));
    for (int64_t i = 0; i < L0; i++) by_pos[size_t(i)] = i;
    std::sort(by_pos.begin(), by_pos.end(), [&](int64_t x, int64_t y) { return first_pos[size_t(x)] < first_pos[size_t(y)]; });
    LinkMap graph, pruned, pruned_by_proj;
    for (int64_t i : by_pos) graph.emplace(make_key(la[size_t(i)], lb[size_t(i)]), int(i));                  // :183 (first insertion of every pair)
    for (const auto& kv : graph) if (keep3d[size_t(kv.second)]) pruned.emplace(kv.first, kv.second);          // :297-323
    for (const auto& kv : pruned) if (keepproj[size_t(kv.second)]) pruned_by_proj.emplace(kv.first, kv.second);   // :355-396
    auto write_graph = [&](const LinkMap& m, const char* name) {                                             // write_edge_graph, :195-221
        std::string s;
        for (const auto& kv : m) {
            const int a = kv.first.first, b = kv.first.second;
            double t1[3], t2[3];
            for (int k = 0; k < 3; k++) { t1[k] = tan0[size_t(a) * 3 + k]; t2[k] = tan0[size_t(b) * 3 + k]; }
            for (double* t : {t1, t2}) { const double z = (t[0] * t[0] + t[1] * t[1]) + t[2] * t[2]; if (z > 0) { const double d = std::sqrt(z); t[0] /= d; t[1] /= d; t[2] /= d; } }
            for (int k = 0; k < 3; k++) { put_g(&s, loc[size_t(a) * 3 + k]); s.push_back('\t'); }
            for (int k = 0; k < 3; k++) { put_g(&s, loc[size_t(b) * 3 + k]); s.push_back('\t'); }
            for (int k = 0; k < 3; k++) { put_g(&s, t1[k]); s.push_back('\t'); }
            for (int k = 0; k < 3; k++) { put_g(&s, t2[k]); s.push_back(k == 2 ? '\n' : '\t'); }
        }
        flush_to(outputs_dir + "/" + name + ".txt", s);
    };
    if (write_edge_graphs) { write_graph(graph, "3D_edge_graph"); write_graph(pruned, "3D_edge_pruned_graph"); write_graph(pruned_by_proj, "3D_edge_pruned_graph_by_proj"); }
    // ---- neighbour lists in the reference's push order (buildEdgeNodeGraph, :440-455), then the alignment on the device
    std::vector<std::vector<int>> nbr(static_cast<size_t>(n));
    for (const auto& kv : pruned_by_proj) { nbr[size_t(kv.first.first)].push_back(kv.first.second); nbr[size_t(kv.first.second)].push_back(kv.first.first); }
    std::vector<int64_t> adj_off(size_t(n) + 1, 0);
    for (int64_t i = 0; i < n; i++) adj_off[size_t(i) + 1] = adj_off[size_t(i)] + int64_t(nbr[size_t(i)].size());
    std::vector<int32_t> adj(size_t(std::max<int64_t>(adj_off[size_t(n)], 1)));
    for (int64_t i = 0; i < n; i++) std::copy(nbr[size_t(i)].begin(), nbr[size_t(i)].end(), adj.begin() + adj_off[size_t(i)]);
    if (!ck(cvx_graph_set_neighbors(g, adj_off.data(), adj.data()), "cvx_graph_set_neighbors")) return false;
    report.graph_s = now_s() - t_begin;
    {   // 3D_edges_before_smoothing.txt (:481-484)
        std::string s;
        for (int64_t i = 0; i < n; i++) { append_eigen_row3(&s, &loc[size_t(i) * 3]); s.push_back(' '); append_eigen_row3(&s, &ori[size_t(i) * 3]); s.push_back('\n'); }
        flush_to(outputs_dir + "/3D_edges_before_smoothing.txt", s);
    }
    if (!quiet) std::cout << "Start aligning edges..." << std::endl;
    const double t_align = now_s();
    std::vector<double> aloc(size_t(n) * 3), aori(size_t(n) * 3);
    float kernel_ms = 0.f;
    if (!ck(cvx_graph_align(g, ori.data(), num_of_iterations, init_force_step_size, init_torque_step_size, aloc.data(), aori.data(), &kernel_ms), "cvx_graph_align")) return false;
    cvx_graph_destroy(g);
    g = nullptr;
    report.align_s = now_s() - t_align; report.align_kernel_ms = kernel_ms;
    {   // 3D_edges_after_smoothing.txt (:583-585) and pruned_graph_aligned.txt (:590-613)
        std::string s;
        for (int64_t i = 0; i < n; i++) { append_eigen_row3(&s, &aloc[size_t(i) * 3]); s.push_back(' '); append_eigen_row3(&s, &aori[size_t(i) * 3]); s.push_back('\n'); }
        flush_to(outputs_dir + "/3D_edges_after_smoothing.txt", s);
        s.clear();
        for (int64_t i = 0; i < n; i++)
            for (int j : nbr[size_t(i)]) {
                for (int k = 0; k < 3; k++) { put_g(&s, aloc[size_t(i) * 3 + k]); s.push_back('\t'); }
                for (int k = 0; k < 3; k++) { put_g(&s, aloc[size_t(j) * 3 + k]); s.push_back('\t'); }
                for (int k = 0; k < 3; k++) { put_g(&s, aori[size_t(i) * 3 + k]); s.push_back('\t'); }
                for (int k = 0; k < 3; k++) { put_g(&s, aori[size_t(j) * 3 + k]); s.push_back('\t'); }
                s.push_back('\n');
            }
        flush_to(outputs_dir + "/pruned_graph_aligned.txt", s);
    }
    const double t_curves = now_s();
    // ---- createConnectivityGraph (:639-747): sequential, order dependent (orientation flips propagate from node to node)
    std::vector<V3> node_loc(static_cast<size_t>(n)), node_ori(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; i++) { node_loc[size_t(i)] = v3(&aloc[size_t(i) * 3]); node_ori[size_t(i)] = v3(&aori[size_t(i) * 3]); }
    std::vector<char> fixed(size_t(n), 0);
    std::vector<GraphNode> cg(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; i++) {
        GraphNode gn;
        gn.location = node_loc[size_t(i)];
        gn.orientation = node_ori[size_t(i)];                                // copied BEFORE the flips below (:650-652)
        const std::vector<int>& nb = nbr[size_t(i)];
        if (!nb.empty()) {
            for (int j : nb)                                                 // :661-668 the first neighbour that is already fixed decides this node's sign
                if (fixed[size_t(j)]) {
                    if (!(dot(node_ori[size_t(j)], node_ori[size_t(i)]) > 0)) node_ori[size_t(i)] = neg(node_ori[size_t(i)]);
                    break;
                }
            fixed[size_t(i)] = 1;
            std::vector<std::pair<int, V3>> updated;
            std::vector<double> proj;
            for (int j : nb) {                                               // :672-688
                if (!(dot(node_ori[size_t(i)], node_ori[size_t(j)]) > 0)) node_ori[size_t(j)] = neg(node_ori[size_t(j)]);
                updated.emplace_back(j, node_ori[size_t(j)]);
                proj.push_back(dot(sub(node_loc[size_t(j)], node_loc[size_t(i)]), node_ori[size_t(i)]));   // getLineVariable, util.cpp:177-179
                fixed[size_t(j)] = 1;
            }
            int left = -1, right = -1;
            double min_pos = std::numeric_limits<double>::max(), max_neg = -std::numeric_limits<double>::max();
            for (size_t j = 0; j < proj.size(); j++) {                       // :691-706
                if (proj[j] >= 0 && proj[j] < min_pos) { min_pos = proj[j]; left = int(j); }
                if (proj[j] < 0 && proj[j] > max_neg) { max_neg = proj[j]; right = int(j); }
            }
            if (left >= 0) { gn.left = updated[size_t(left)].first; gn.left_orientation = updated[size_t(left)].second; }
            if (right >= 0) { gn.right = updated[size_t(right)].first; gn.right_orientation = updated[size_t(right)].second; }
        }
        cg[size_t(i)] = gn;
    }
    // the reference reads connectivity_graph (unordered_map<int, node>, filled for i = 0 .. n-1) in its iteration order
    std::vector<int> cg_order;
    {
        std::unordered_map<int, int> m;
        for (int i = 0; i < int(n); i++) m[i] = i;
        cg_order.reserve(size_t(n));
        for (const auto& kv : m) cg_order.push_back(kv.first);
    }
    {   // writeConnectivityGraphToFile (:1090-1124)
        std::string s = "# Connectivity Graph\n# Format: NodeIndex Location(x,y,z); LeftNeighborIndex LeftOrientation(x,y,z); RightNeighborIndex RightOrientation(x,y,z)\n";
        for (int i : cg_order) {
            const GraphNode& q = cg[size_t(i)];
            s += std::to_string(i) + " ";
            put_g(&s, q.location.x); s += " "; put_g(&s, q.location.y); s += " "; put_g(&s, q.location.z); s += " ;";
            put_g(&s, q.orientation.x); s += " "; put_g(&s, q.orientation.y); s += " "; put_g(&s, q.orientation.z); s += " ";
            s += std::to_string(q.left) + " ";
            if (q.left >= 0) { put_g(&s, q.left_orientation.x); s += " "; put_g(&s, q.left_orientation.y); s += " "; put_g(&s, q.left_orientation.z); s += " ;"; }
            else s += "0 0 0 ";
            s += std::to_string(q.right) + " ";
            if (q.right >= 0) { put_g(&s, q.right_orientation.x); s += " "; put_g(&s, q.right_orientation.y); s += " "; put_g(&s, q.right_orientation.z); }
            else s += "0 0 0";
            s += "\n";
        }
        flush_to(outputs_dir + "/connectivity_graph.txt", s);
    }
    // ---- buildCurvesFromConnectivityGraph (:749-967)
    std::vector<Curve> curves;
    {
        std::vector<char> assigned(size_t(n), 0);
        int curve_count = 0;
        for (int start : cg_order) {
            if (assigned[size_t(start)]) continue;
            Curve c;
            c.edge_indices.push_back(start);
            assigned[size_t(start)] = 1;
            for (int dir = 0; dir < 2; dir++) {                              // dir 0: follow right neighbours (:768-799), dir 1: left (:802-833)
                int cur = start;
                while (true) {
                    const int nxt = dir == 0 ? cg[size_t(cur)].right : cg[size_t(cur)].left;
                    cg[size_t(cur)].curve_index = curve_count;
                    if (nxt < 0 || nxt >= int(n)) break;
                    if (assigned[size_t(nxt)]) {
                        const bool own = std::find(c.edge_indices.begin(), c.edge_indices.end(), nxt) != c.edge_indices.end();
                        if (dir == 0) {
                            if (own) c.b_loops_back_on_right = true;
                            else { c.to_be_merged_right_edge_index = nxt; c.to_be_merged_right_curve_index = cg[size_t(nxt)].curve_index; }
                        } else {
                            if (own) c.b_loops_back_on_left = true;
                            else { c.to_be_merged_left_edge_index = nxt; c.to_be_merged_left_curve_index = cg[size_t(nxt)].curve_index; }
                        }
                        break;
                    }
                    if (dir == 0) c.edge_indices.push_back(nxt); else c.edge_indices.insert(c.edge_indices.begin(), nxt);
                    assigned[size_t(nxt)] = 1;
                    cur = nxt;
                }
            }
            if (c.to_be_merged_left_curve_index == -1 && c.to_be_merged_right_curve_index == -1) {          // :835-838
                c.consolidation_set_from_left = curve_count; c.consolidation_set_from_right = curve_count;
            } else {                                                                                       // :839-866
                const int tl = c.to_be_merged_left_curve_index, tr = c.to_be_merged_right_curve_index;
                if (tl >= 0 && tl < int(curves.size())) {
                    if (in_first_or_last_two(curves[size_t(tl)].edge_indices, c.to_be_merged_left_edge_index))
                        c.consolidation_set_from_left = curves[size_t(tl)].consolidation_set_from_left == -1 ? curves[size_t(tl)].consolidation_set_from_right
                                                                                                           : curves[size_t(tl)].consolidation_set_from_left;
                    else c.consolidation_set_from_left = curve_count;
                }
                if (tr >= 0 && tr < int(curves.size())) {
                    if (in_first_or_last_two(curves[size_t(tr)].edge_indices, c.to_be_merged_right_edge_index))
                        c.consolidation_set_from_right = curves[size_t(tr)].consolidation_set_from_right == -1 ? curves[size_t(tr)].consolidation_set_from_left
                                                                                                             : curves[size_t(tr)].consolidation_set_from_right;
                    else c.consolidation_set_from_right = curve_count;
                }
            }
            curve_count++;
            curves.push_back(c);
        }
    }
    // ---- sets of curves to be consolidated (:871-944)
    std::vector<std::vector<int>> sets(curves.size());
    {
        std::set<std::pair<int, int>> pairs_to_join;
        for (int id = 0; id < int(curves.size()); id++) {
            const Curve& c = curves[size_t(id)];
            const int fl = c.consolidation_set_from_left, fr = c.consolidation_set_from_right;
            if (fl == -1 && fr == -1) { if (!quiet) std::cout << "Something's wrong in curve " << id << std::endl; continue; }   // the reference indexes [-1] here (UB); skipped
            if (fl == fr) sets[size_t(fl)].push_back(id);
            else if (fl >= 0 && (fr == -1 || fr == id)) sets[size_t(fl)].push_back(id);
            else if ((fl == -1 || fl == id) && fr >= 0) sets[size_t(fr)].push_back(id);
            else { sets[size_t(fl)].push_back(id); sets[size_t(fr)].push_back(id); pairs_to_join.insert(fl < fr ? std::make_pair(fl, fr) : std::make_pair(fr, fl)); }
        }
        auto join = [&](const std::set<std::pair<int, int>>& ps) {                                      // :899-905, :925-931 (reverse order)
            for (auto it = ps.rbegin(); it != ps.rend(); ++it) {
                std::vector<int> a = sets[size_t(it->first)];
                const std::vector<int>& b = sets[size_t(it->second)];
                a.insert(a.end(), b.begin(), b.end());
                sets[size_t(it->first)] = a;
                sets[size_t(it->second)].clear();
            }
        };
        auto tidy = [&]() {                                                                              // :908-917
            for (auto& s : sets) { if (s.empty()) continue; std::sort(s.begin(), s.end()); s.erase(std::unique(s.begin(), s.end()), s.end()); }
            sets.erase(std::remove_if(sets.begin(), sets.end(), [](const std::vector<int>& s) { return s.empty(); }), sets.end());
        };
        join(pairs_to_join);
        tidy();
        while (true) {                                                                                   // :919-943 sets that still share a curve id
            std::unordered_map<int, std::vector<int>> where;                                             // check_duplicate_curve_ids, :1212-1245
            for (int i = 0; i < int(sets.size()); i++) for (int x : sets[size_t(i)]) where[x].push_back(i);
            std::set<std::pair<int, int>> dup;
            for (int i = 0; i < int(sets.size()); i++)
                for (int x : sets[size_t(i)]) {
                    const std::vector<int>& w = where[x];
                    if (w.size() > 1) dup.insert(w[0] < w[1] ? std::make_pair(w[0], w[1]) : std::make_pair(w[1], w[0]));
                }
            if (dup.empty()) break;
            join(dup);
            tidy();
        }
    }
    if (!quiet) {
        std::cout << "Initial number of curves from connectivity graph = " << curves.size() << std::endl;
        std::cout << "Number of sets of curves for consolidation       = " << sets.size() << std::endl;
    }
    {   // writeCurveIndiciesForMerging (:1166-1186)
        std::string s;
        for (const auto& set : sets) { for (int id : set) s += std::to_string(id) + " "; s += "\n"; }
        flush_to(outputs_dir + "/curve_indices_for_merging.txt", s);
    }
    // ---- merge_multiple_curves / merge_curve_pair (:969-1054): greedy by end-point distance
    auto P = [&](int node) { return cg[size_t(node)].location; };
    auto merge_pair = [&](const std::vector<int>& c1, const std::vector<int>& c2in) {
        const V3 d1 = sub(P(c1.back()), P(c1.front())), d2 = sub(P(c2in.back()), P(c2in.front()));
        std::vector<int> c2 = dot(d1, d2) < 0 ? std::vector<int>(c2in.rbegin(), c2in.rend()) : c2in;     // make_curve_orientation_consistent
        const double dAB = norm(sub(P(c1.back()), P(c2.front()))), dBA = norm(sub(P(c2.back()), P(c1.front())));
        const double dRBA = norm(sub(P(c2.front()), P(c1.front()))), dARB = norm(sub(P(c1.back()), P(c2.back())));
        double best = dAB;
        int choice = 0;
        if (dBA < best) { best = dBA; choice = 1; }
        if (dRBA < best) { best = dRBA; choice = 2; }
        if (dARB < best) { best = dARB; choice = 3; }
        std::vector<int> m;
        if (choice == 0) { m = c1; m.insert(m.end(), c2.begin(), c2.end()); }
        else if (choice == 1) { m = c2; m.insert(m.end(), c1.begin(), c1.end()); }
        else if (choice == 2) { m.assign(c2.rbegin(), c2.rend()); m.insert(m.end(), c1.begin(), c1.end()); }
        else { m = c1; m.insert(m.end(), c2.rbegin(), c2.rend()); }
        return m;
    };
    std::vector<Curve> final_curves(sets.size());
    for (size_t si = 0; si < sets.size(); si++) {
        std::vector<std::vector<int>> rest;
        for (int id : sets[si]) rest.push_back(curves[size_t(id)].edge_indices);
        std::vector<int> merged = rest.front();
        rest.erase(rest.begin());
        while (!rest.empty()) {
            int best_i = -1;
            double best_d = std::numeric_limits<double>::max();
            for (size_t i = 0; i < rest.size(); i++) {
                const std::vector<int>& c = rest[i];
                const double d0 = norm(sub(P(merged.front()), P(c.front()))), d1 = norm(sub(P(merged.front()), P(c.back())));
                const double d2 = norm(sub(P(merged.back()), P(c.front()))), d3 = norm(sub(P(merged.back()), P(c.back())));
                const double d = std::min(std::min(d0, d1), std::min(d2, d3));
                if (d < best_d) { best_d = d; best_i = int(i); }
            }
            merged = merge_pair(merged, rest[size_t(best_i)]);
            rest.erase(rest.begin() + best_i);
        }
        final_curves[si].edge_indices = merged;
    }
    if (!quiet) std::cout << "Complete tracing 3D edges to form 3D curves" << std::endl;
    // ---- writeCurvesToFile (:1126-1164)
    auto write_curves = [&](const std::vector<Curve>& cs, const char* name, bool info) {
        std::string s = "# Curves from Connectivity Graph\n# Format: CurveID 3DEdgeIndex Location(x,y,z) Orientation(x,y,z)\n";
        for (size_t id = 0; id < cs.size(); id++) {
            const Curve& c = cs[id];
            for (int node : c.edge_indices) {
                const double l[3] = {cg[size_t(node)].location.x, cg[size_t(node)].location.y, cg[size_t(node)].location.z};
                const double o[3] = {cg[size_t(node)].orientation.x, cg[size_t(node)].orientation.y, cg[size_t(node)].orientation.z};
                s += std::to_string(id) + " " + std::to_string(node) + " ";
                append_eigen_row3(&s, l); s.push_back(' '); append_eigen_row3(&s, o);
                if (info) {
                    s += " " + std::to_string(int(c.b_loops_back_on_left)) + " " + std::to_string(int(c.b_loops_back_on_right));
                    s += " " + std::to_string(c.to_be_merged_left_edge_index) + " " + std::to_string(c.to_be_merged_right_edge_index);
                    s += " " + std::to_string(c.to_be_merged_left_curve_index) + " " + std::to_string(c.to_be_merged_right_curve_index);
                    s += " " + std::to_string(c.consolidation_set_from_left) + " " + std::to_string(c.consolidation_set_from_right);
                }
                s.push_back('\n');
            }
            s.push_back('\n');
        }
        const bool ok = flush_to(outputs_dir + "/" + name + ".txt", s);
        if (!quiet) std::cout << "Wrote " << cs.size() << " curves to " << outputs_dir << "/" << name << ".txt" << std::endl;
        return ok;
    };
    const bool ok1 = write_curves(curves, "curves_from_connectivity_graph", true);
    const bool ok2 = write_curves(final_curves, "final_curves", false);
    report.curves_from_graph = int64_t(curves.size()); report.final_curves = int64_t(final_curves.size());
    report.curves_s = now_s() - t_curves; report.total_s = now_s() - t_begin;
    if (!ok1 || !ok2) return fail("cannot write the curve files under " + outputs_dir);
    return true;
}
