// edge_mapping.hpp -- the consumer of the edge-sketch path in the drop-in: the 3-D edge -> supporting 2-D edges table and the grouping of
// the 3-D edges into curves, with the reference's class and member names (Edge_Reconst/edge_mapping.{hpp,cpp}).
//
//   add3DToSupportingEdgesMapping   edge_mapping.cpp:20-33     host (insertions from Finalize_Edge_Pairs_and_Reconstruct_3D_Edges, :682-726)
//   consolidate_3D_edges            edge_mapping.cpp:1056-1088 the data-parallel front half on the DEVICE through the C-ABI (cvx_graph_*: links from
//                                   curvelets, the two prunings, the alignment iterations; csrc/cvx_graph.cu), the order-dependent back half
//                                   (connectivity graph, curve tracing, merging, writers: :639-1164) on the host, edge_mapping.cpp of this directory
//
// The reference keys its maps on the exact doubles of a 3-D point and several of its stages depend on the ITERATION ORDER of those
// unordered_maps (node numbering, neighbour order, the order in which curves are traced).  The drop-in reproduces those orders by keeping the
// same kind of container with the same hash (value of std::hash<double> combined as HashEigenVector3d / HashEigenVector3dPair do,
// edge_mapping.hpp:18-52) and the same insertion sequence, so that on identical inputs the nodes, neighbour lists and curves carry the
// reference's indices and the output files can be compared line by line.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <unordered_map>
#include <vector>

struct cvx_ctx;
namespace curvix_host { class FlatYaml; }

class EdgeMapping {
public:
    struct SupportingEdgeData {
        double edge[2];               // (corrected) 2-D location handed in by the caller
        double edge_uncorrected[3];   // x, y, theta of the original edge
        int image_number;
        int edge_idx;
        double tangents_3D_world[3];
    };
    struct Edge3D {
        double point[3];
        std::vector<SupportingEdgeData> supports;
    };
    struct ConsolidationReport {      // what consolidate_3D_edges did (printed by curvix_main, read by the tests)
        int64_t nodes = 0, records = 0, curvelets = 0, links = 0, links_after_3d = 0, links_after_projection = 0;
        int64_t curves_from_graph = 0, final_curves = 0;
        double stats[6] = {0, 0, 0, 0, 0, 0};   // mu_t, sigma_t, mu_n, sigma_n, mu_theta, sigma_theta (edge_mapping.cpp:291-294)
        double graph_s = 0, align_s = 0, align_kernel_ms = 0, curves_s = 0, total_s = 0;
    };

    // Same role and argument order as the reference method (rotation / translation of the image are a function of
    // image_number and are not stored again per record).
    void add3DToSupportingEdgesMapping(const double edge_3D[3], const double tangents_3D_world[3], const double supporting_edge[2],
                                       const double supporting_edge_uncorrected[3], int image_number, int edge_idx) {
        Key k;
        std::memcpy(k.v, edge_3D, 24);
        auto it = index_.find(k);
        size_t slot;
        if (it == index_.end()) {   // the reference keys its unordered_map on the exact doubles of the 3-D point
            slot = edge_3D_to_supporting_edges.size();
            index_.emplace(k, slot);
            Edge3D e;
            std::memcpy(e.point, edge_3D, 24);
            edge_3D_to_supporting_edges.push_back(std::move(e));
        } else {
            slot = it->second;
        }
        SupportingEdgeData d;
        d.edge[0] = supporting_edge[0]; d.edge[1] = supporting_edge[1];
        std::memcpy(d.edge_uncorrected, supporting_edge_uncorrected, 24);
        d.image_number = image_number; d.edge_idx = edge_idx;
        std::memcpy(d.tangents_3D_world, tangents_3D_world, 24);
        edge_3D_to_supporting_edges[slot].supports.push_back(d);
    }

    // 3-D edges in first-insertion order; reference_order() lists them in the iteration order of the reference's map
    std::vector<Edge3D> edge_3D_to_supporting_edges;
    std::vector<size_t> reference_order() const {
        std::vector<size_t> o;
        o.reserve(index_.size());
        for (const auto& kv : index_) o.push_back(kv.second);
        return o;
    }

    size_t num_3D_edges() const { return edge_3D_to_supporting_edges.size(); }
    size_t num_support_links() const {
        size_t n = 0;
        for (const auto& e : edge_3D_to_supporting_edges) n += e.supports.size();
        return n;
    }

    // One line per link: X Y Z  image edge_idx  x y  tx ty tz  ux uy utheta (input of consolidate_3D_edges; read_support_links restores it)
    bool write_support_links(const std::string& path) const {
        FILE* f = std::fopen(path.c_str(), "w");
        if (!f) return false;
        for (const auto& e : edge_3D_to_supporting_edges)
            for (const auto& s : e.supports)
                std::fprintf(f, "%.17g %.17g %.17g %d %d %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", e.point[0], e.point[1], e.point[2],
                             s.image_number, s.edge_idx, s.edge[0], s.edge[1], s.tangents_3D_world[0], s.tangents_3D_world[1], s.tangents_3D_world[2],
                             s.edge_uncorrected[0], s.edge_uncorrected[1], s.edge_uncorrected[2]);
        return std::fclose(f) == 0;
    }
    bool read_support_links(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "r");
        if (!f) return false;
        double r[13];
        while (std::fscanf(f, "%lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf %lf", r, r + 1, r + 2, r + 3, r + 4, r + 5, r + 6, r + 7, r + 8, r + 9, r + 10,
                           r + 11, r + 12) == 13)
            add3DToSupportingEdgesMapping(r, r + 7, r + 5, r + 10, int(r[3]), int(r[4]));
        std::fclose(f);
        return true;
    }

    // EdgeMapping::Setup_Data_Parameters (edge_mapping.cpp:64-70)
    void Setup_Data_Parameters(const curvix_host::FlatYaml& settings);
    // EdgeMapping::consolidate_3D_edges (edge_mapping.cpp:1056-1088).  all_R n x 9 row-major, all_T n x 3, K 9; files go to outputs_dir
    // (the reference's ../../outputs): 3D_edges_before_smoothing.txt, 3D_edges_after_smoothing.txt, pruned_graph_aligned.txt,
    // connectivity_graph.txt, curve_indices_for_merging.txt, curves_from_connectivity_graph.txt, final_curves.txt and, with
    // write_edge_graphs, the three 3D_edge*graph*.txt files of WRITE_3D_EDGE_GRAPH (definitions.h:9).
    bool consolidate_3D_edges(cvx_ctx* gpu, const std::vector<double>& all_R, const std::vector<double>& all_T, const double K[9],
                              int Num_Of_Total_Imgs, const std::string& outputs_dir, bool write_edge_graphs = true, bool quiet = false);
    ConsolidationReport report;
    std::string last_error;

    // parameters of the reference's definitions.h:28-38 (compile-time there)
    double prune_lambda1 = 1, prune_lambda2 = 1, prune_lambda3 = 1, prune_by_proj_prox_thresh = 6;
    int num_of_iterations = 1000;
    double init_force_step_size = 0.01, init_torque_step_size = 0.01;

    int thresh_EDG = 1, Num_Of_Images = 0;
    std::string Dataset_Path, Dataset_Name, Scene_Name;

    // HashEigenVector3d (edge_mapping.hpp:18-26 of the reference) on three doubles
    static size_t hash_point(const double v[3]) {
        size_t seed = 0;
        for (int i = 0; i < 3; i++) seed ^= std::hash<double>()(v[i]) + 0x9e3779b9 + (seed << 6) + (seed >> 2);
        return seed;
    }

private:
    struct Key {
        double v[3];
        bool operator==(const Key& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2]; }   // Eigen's operator==
    };
    struct KeyHash {
        size_t operator()(const Key& k) const { return hash_point(k.v); }
    };
    std::unordered_map<Key, size_t, KeyHash> index_;
};
