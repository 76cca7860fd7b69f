// edge_mapping.hpp -- the consumer-side interface of the edge-sketch path, host only.
//
// In the reference, EdgeMapping (Edge_Reconst/edge_mapping.{hpp,cpp}, 1.5 kLoC) receives every reconstructed 3-D edge
// with its supporting 2-D edges through add3DToSupportingEdgesMapping (edge_mapping.cpp:20-33, called from
// EdgeSketch_Core.cpp:682-726) and later groups them into curves (consolidate_3D_edges -> final_curves.txt).  The curve
// grouping stays host code of the reference and is out of scope here (BASELINE.json north_star).
// What the path itself needs from this class is the insertion API and the 3-D edge -> supporting edges table that
// Calculate_Edge_Support_Ratios_And_Select_Next_Views walks (EdgeSketch_Core.cpp:785-797); this header provides exactly
// that, with the reference's member names, so that the drop-in EdgeSketch_Core below compiles against either class.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

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

    // Same role and argument order as the reference method (rotation / translation of the image are a function of
    // image_number and are not stored again per record).
    void add3DToSupportingEdgesMapping(const double edge_3D[3], const double tangents_3D_world[3], const double supporting_edge[2],
                                       const double supporting_edge_uncorrected[3], int image_number, int edge_idx) {
        Key k;
        std::memcpy(k.bits, edge_3D, 24);
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

    // 3-D edges in first-insertion order (the reference iterates an unordered_map; consumers must not rely on order).
    std::vector<Edge3D> edge_3D_to_supporting_edges;

    size_t num_3D_edges() const { return edge_3D_to_supporting_edges.size(); }
    size_t num_support_links() const {
        size_t n = 0;
        for (const auto& e : edge_3D_to_supporting_edges) n += e.supports.size();
        return n;
    }

    // One line per link: X Y Z  image edge_idx  x y  tx ty tz   (input of the reference's host-side curve grouping)
    bool write_support_links(const std::string& path) const {
        FILE* f = std::fopen(path.c_str(), "w");
        if (!f) return false;
        for (const auto& e : edge_3D_to_supporting_edges)
            for (const auto& s : e.supports)
                std::fprintf(f, "%.17g %.17g %.17g %d %d %.17g %.17g %.17g %.17g %.17g\n", e.point[0], e.point[1], e.point[2], s.image_number,
                             s.edge_idx, s.edge[0], s.edge[1], s.tangents_3D_world[0], s.tangents_3D_world[1], s.tangents_3D_world[2]);
        return std::fclose(f) == 0;
    }

private:
    struct Key {
        uint64_t bits[3];
        bool operator==(const Key& o) const { return bits[0] == o.bits[0] && bits[1] == o.bits[1] && bits[2] == o.bits[2]; }
    };
    struct KeyHash {
        size_t operator()(const Key& k) const {
            uint64_t h = 0x9e3779b97f4a7c15ull;
            for (uint64_t b : k.bits) { h ^= b + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); }
            return size_t(h);
        }
    };
    std::unordered_map<Key, size_t, KeyHash> index_;
};
