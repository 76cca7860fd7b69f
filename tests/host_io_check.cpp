// CPU-side check of the host drop-in's I/O pieces (no GPU, no CUDA library): flat-YAML reader, edge/R/T readers, the binary
// edge cache and the Eigen-format writer.  Driven by tests/test_host_io.py.
#include <cstdio>
#include <iostream>
#include <string>

#include "../3d_curvix_b200/host/curve_io.hpp"
#include "../3d_curvix_b200/host/dataset_io.hpp"
#include "../3d_curvix_b200/host/edge_mapping.hpp"
#include "../3d_curvix_b200/host/flat_yaml.hpp"

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const std::string cmd = argv[1], arg = argv[2];
    using namespace curvix_host;
    if (cmd == "yaml") {
        FlatYaml y = FlatYaml::load(arg);
        std::printf("%d %d %d %.17g %.17g %d %d %.17g %.17g %.17g %d\n", y.as_int("Num_Of_OMP_Threads"), y.as_int("Init_Hypo1_View_Index"),
                    y.as_int("Init_Hypo2_View_Index"), y.as_double("delta"), y.as_double("delta_theta"), y.as_int("Max_Num_Of_Support_Views"),
                    y.as_int("TOED_Thresh"), y.as_double("Parallel_Epipolar_Line_Angle"), y.as_double("Ratio_Of_Claimed_Edges_to_Stop"),
                    y.as_double("fx"), y.as_int("Max_Num_Of_3D_Edge_Sketch_Passes"));
        std::printf("%s|%s|%s|%d|%d|%d|%d\n", y.str("Dataset_Path").c_str(), y.str("Dataset_Name").c_str(), y.str("Scene_Name").c_str(),
                    y.as_int("Total_Num_Of_Images"), y.as_int("Image_Rows"), y.as_int("Image_Cols"), int(y.as_bool("Use_Multiple_K")));
        return 0;
    }
    if (cmd == "edges") {   // parse text -> write cache -> read cache -> dump
        EdgeTable t, u;
        if (!read_edge_file(arg, &t)) return 3;
        if (!write_edge_cache(arg + ".cvxe", t) || !read_edge_cache(arg + ".cvxe", &u)) return 4;
        std::printf("%d\n", u.size());
        for (int i = 0; i < u.size(); i++) std::printf("%.17g %.17g %.17g %.17g\n", u.x[i], u.y[i], u.theta[i], u.strength[i]);
        return 0;
    }
    if (cmd == "rt") {
        std::vector<double> R, T;
        if (!read_R_file(arg + "/R_matrix.txt", &R) || !read_T_file(arg + "/T_matrix.txt", &T)) return 3;
        std::printf("%zu %zu\n", R.size() / 9, T.size() / 3);
        for (double v : R) std::printf("%.17g\n", v);
        for (double v : T) std::printf("%.17g\n", v);
        return 0;
    }
    if (cmd == "eigen") {   // reads "rows cols v..." from stdin, writes the matrix to arg
        long rows; int cols;
        if (!(std::cin >> rows >> cols)) return 3;
        std::vector<double> m(size_t(rows) * size_t(cols));
        for (double& v : m) std::cin >> v;
        return write_eigen_matrix(arg, m.data(), rows, cols) ? 0 : 4;
    }
    if (cmd == "curves") {   // read arg (reference layout), report, write it back to arg + ".out"
        std::vector<Curve3D> c;
        if (!read_curves_file(arg, &c)) return 3;
        size_t nodes = 0;
        for (const auto& k : c) nodes += k.nodes.size();
        std::printf("%zu %zu\n", c.size(), nodes);
        return write_curves_file(arg + ".out", c) ? 0 : 4;
    }
    if (cmd == "mapping") {   // same 3-D point twice + another one -> 2 edges, 3 links
        EdgeMapping em;
        const double p[3] = {1, 2, 3}, q[3] = {1, 2, 3.0000001}, t[3] = {0, 0, 1}, e[2] = {5, 6}, u[3] = {5, 6, 0.1};
        em.add3DToSupportingEdgesMapping(p, t, e, u, 0, 7);
        em.add3DToSupportingEdgesMapping(p, t, e, u, 1, 8);
        em.add3DToSupportingEdgesMapping(q, t, e, u, 2, 9);
        std::printf("%zu %zu\n", em.num_3D_edges(), em.num_support_links());
        return em.write_support_links(arg) ? 0 : 4;
    }
    return 2;
}
