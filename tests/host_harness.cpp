// TEST INFRASTRUCTURE: compiles the product's per-element FP64 math (3d_curvix_b200/csrc/cvx_math.cuh) for the host
// so that its transcription can be checked against the oracle on a CPU-only box.  Not part of the product; the
// product never runs this math on the CPU.
#include <cstring>

#include "../3d_curvix_b200/csrc/cvx_graph_math.cuh"
#include "../3d_curvix_b200/csrc/cvx_math.cuh"

using namespace cvx;

extern "C" {
int hh_wedge_row(double x, double y, const double* F, const double* e1, const double* e2, double delta, int rows, int cols, double* lo,
                 double* hi) {
    return wedge_row(x, y, F, e1[0], e1[1], e2[0], e2[1], delta, rows, cols, lo, hi);
}
int hh_wedge_row_fast(double x, double y, const double* F, const double* e1, const double* e2, double delta, int rows, int cols, double* lo,
                      double* hi) {
    return wedge_row_fast(x, y, F, e1[0], e1[1], e2[0], e2[1], delta, rows, cols, lo, hi);
}
double hh_ore_angle(double x, double y, double ex, double ey) { return ore_angle(x, y, ex, ey); }
int hh_correct_one(const double* F21, double x1, double y1, double x2, double y2, double th2, double loc, double tang, double orient,
                   double* cx, double* cy) {
    double c[3];
    epip_line_coeffs(F21, x1, y1, c);
    return correct_one(c, x2, y2, th2, loc, tang, orient, cx, cy) ? 1 : 0;
}
void hh_reproject(double x1, double y1, double th1, double x2, double y2, double th2, const double* K1inv, const double* K2inv,
                  const double* R21, const double* T21, const double* R31, const double* T31, double* t3) {
    PassConst pc;
    std::memset(&pc, 0, sizeof(pc));
    std::memcpy(pc.K1inv, K1inv, 72); std::memcpy(pc.K2inv, K2inv, 72); std::memcpy(pc.R21, R21, 72); std::memcpy(pc.T21, T21, 24);
    RowGeom g;
    row_geometry(x1, y1, th1, x2, y2, th2, pc, &g);
    reproject_tangent(g, R31, T31, &t3[0], &t3[1]);
}
void hh_triangulate(const double* p1, const double* p2, const double* R21, const double* T21, const double* Kinv, double* out3) {
    triangulate2(p1[0], p1[1], p2[0], p2[1], R21, T21, Kinv, out3);
}
void hh_tangent3d(const double* e1, const double* e2, const double* K1inv, const double* K2inv, const double* R1, const double* R21,
                  double* out3) {
    PassConst pc;
    std::memset(&pc, 0, sizeof(pc));
    std::memcpy(pc.K1inv, K1inv, 72); std::memcpy(pc.K2inv, K2inv, 72); std::memcpy(pc.R21, R21, 72); std::memcpy(pc.R1, R1, 72);
    tangent3d_world(e1[0], e1[1], e1[2], e2[0], e2[1], e2[2], pc, out3);
}
// ---- 3-D edge graph arithmetic (cvx_graph_math.cuh): the helpers of align3DEdgesUsingEdgeNodes / the link metrics
void hh_closest_vector_to_line(const double* P1, const double* d1, const double* P2, double* o) {
    const cvxg::D3 r = cvxg::closest_vector_to_line(cvxg::ld3(P1, 0), cvxg::ld3(d1, 0), cvxg::ld3(P2, 0));
    o[0] = r.x; o[1] = r.y; o[2] = r.z;
}
void hh_shortest_align_euler_deg(const double* a, const double* b, double* o) {
    const cvxg::D3 r = cvxg::shortest_align_euler_deg(cvxg::ld3(a, 0), cvxg::ld3(b, 0));
    o[0] = r.x; o[1] = r.y; o[2] = r.z;
}
void hh_euler_to_rotation(double roll, double pitch, double yaw, double* o9) {
    const cvxg::M33 R = cvxg::euler_to_rotation(roll, pitch, yaw);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) o9[3 * i + j] = R.m[i][j];
}
}
