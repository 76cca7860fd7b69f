// curve_io.hpp -- outputs/final_curves.txt and outputs/curves_from_connectivity_graph.txt in the reference's layout, host only.
//
// Writer: EdgeMapping::writeCurvesToFile (Edge_Reconst/edge_mapping.cpp:1126-1164) with b_write_curve_info = false: two '#'
// header lines; per node "CurveID NodeIndex <location.transpose()> <orientation.transpose()>" where each transposed Vector3d goes
// through Eigen's default IOFormat (6 significant digits, the three coefficients right-aligned to the widest of the three, single
// spaces); one blank line after every curve.
// Reader: the loop of visualization/plot_curves.m:17-38 -- skip two header lines, skip empty lines, sscanf '%f', keep lines with
// at least 8 numbers; a new curve starts when the curve id changes.
// The curve grouping itself (EdgeMapping::consolidate_3D_edges) is host code of the reference and stays there (BASELINE.json
// north_star); INTEGRATION.md shows how the reference's EdgeMapping consumes this path's output and writes the file.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace curvix_host {

struct CurveNode {
    int node_index;
    double location[3];
    double orientation[3];
};
struct Curve3D {
    int curve_id;
    std::vector<CurveNode> nodes;
};

// `ostream << Vector3d::transpose()` at the default precision
inline void append_eigen_row3(std::string* out, const double v[3]) {
    char txt[3][32];
    int len[3], width = 0;
    for (int k = 0; k < 3; k++) {
        len[k] = std::snprintf(txt[k], sizeof txt[k], "%.6g", v[k]);
        if (len[k] > width) width = len[k];
    }
    for (int k = 0; k < 3; k++) {
        if (k) out->push_back(' ');
        out->append(size_t(width - len[k]), ' ');
        out->append(txt[k], size_t(len[k]));
    }
}

inline bool write_curves_file(const std::string& path, const std::vector<Curve3D>& curves) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::string buf = "# Curves from Connectivity Graph\n# Format: CurveID 3DEdgeIndex Location(x,y,z) Orientation(x,y,z)\n";   // :1137-1138
    for (size_t c = 0; c < curves.size(); c++) {
        for (const CurveNode& n : curves[c].nodes) {
            buf += std::to_string(c) + " " + std::to_string(n.node_index) + " ";                                             // :1154 (curve_id is the position)
            append_eigen_row3(&buf, n.location);
            buf.push_back(' ');
            append_eigen_row3(&buf, n.orientation);
            buf.push_back('\n');
        }
        buf.push_back('\n');                                                                                                 // :1159
        if (buf.size() > (1u << 20)) { std::fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
    }
    std::fwrite(buf.data(), 1, buf.size(), f);
    return std::fclose(f) == 0;
}

inline bool read_curves_file(const std::string& path, std::vector<Curve3D>* curves) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    curves->clear();
    char line[1024];
    int lineno = 0;
    while (std::fgets(line, sizeof line, f)) {
        if (++lineno <= 2) continue;                        // the two header lines (plot_curves.m:18-19)
        double v[8];
        int n = 0;
        char* p = line;
        while (n < 8) {
            char* e;
            const double x = std::strtod(p, &e);
            if (e == p) break;
            v[n++] = x;
            p = e;
        }
        if (n < 8) continue;                                // empty line, or fewer than 8 numbers (plot_curves.m:31-34)
        const int id = int(v[0]);
        if (curves->empty() || curves->back().curve_id != id) { Curve3D c; c.curve_id = id; curves->push_back(c); }
        CurveNode nd;
        nd.node_index = int(v[1]);
        for (int k = 0; k < 3; k++) { nd.location[k] = v[2 + k]; nd.orientation[k] = v[5 + k]; }
        curves->back().nodes.push_back(nd);
    }
    std::fclose(f);
    return true;
}

}  // namespace curvix_host
