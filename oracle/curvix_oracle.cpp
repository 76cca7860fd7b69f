// =====================================================================================
// curvix_oracle.cpp -- TEST INFRASTRUCTURE ONLY.  See curvix_oracle.hpp for the contract.
//
// CPU restatement of the 3D Curvix edge-sketch hot path.  Every function cites the reference
// file:line it follows (paths relative to the reference root, Edge_Reconst/ unless noted).
// Parity status: pinned against oracle/_ref (the reference's own sources), see curvix_oracle.hpp.
// =====================================================================================
#include "curvix_oracle.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <unordered_map>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

const double kInf = std::numeric_limits<double>::infinity();
const double kNaN = std::numeric_limits<double>::quiet_NaN();

// ----------------------------------------------------------------------------- small linear algebra
struct V3 { double v[3]; };
struct M3 { double m[9]; };  // row-major

inline double dot3(double a0, double a1, double a2, double b0, double b1, double b2) {
    return (a0 * b0 + a1 * b1) + a2 * b2;  // left-to-right, no contraction
}
inline M3 mat_mul(const M3& a, const M3& b) {
    M3 c;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            c.m[i * 3 + j] = dot3(a.m[i * 3], a.m[i * 3 + 1], a.m[i * 3 + 2], b.m[j], b.m[3 + j], b.m[6 + j]);
    return c;
}
inline V3 mat_vec(const M3& a, const V3& x) {
    V3 r;
    for (int i = 0; i < 3; i++) r.v[i] = dot3(a.m[i * 3], a.m[i * 3 + 1], a.m[i * 3 + 2], x.v[0], x.v[1], x.v[2]);
    return r;
}
inline M3 transpose(const M3& a) {
    M3 t;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) t.m[i * 3 + j] = a.m[j * 3 + i];
    return t;
}
inline V3 cross(const V3& a, const V3& b) {
    V3 c;
    c.v[0] = a.v[1] * b.v[2] - a.v[2] * b.v[1];
    c.v[1] = a.v[2] * b.v[0] - a.v[0] * b.v[2];
    c.v[2] = a.v[0] * b.v[1] - a.v[1] * b.v[0];
    return c;
}
inline double norm3(const V3& a) { return std::sqrt((a.v[0] * a.v[0] + a.v[1] * a.v[1]) + a.v[2] * a.v[2]); }

// Eigen 3.3 Matrix3d::inverse(): cofactors * (1/det)  (Eigen/src/LU/InverseImpl.h, compute_inverse_size3_helper)
inline double cofactor3(const M3& a, int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return a.m[i1 * 3 + j1] * a.m[i2 * 3 + j2] - a.m[i1 * 3 + j2] * a.m[i2 * 3 + j1];
}
inline M3 inverse3(const M3& a) {
    const double c00 = cofactor3(a, 0, 0), c10 = cofactor3(a, 1, 0), c20 = cofactor3(a, 2, 0);
    const double det = (c00 * a.m[0] + c10 * a.m[3]) + c20 * a.m[6];
    const double invdet = 1.0 / det;
    M3 r;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) r.m[j * 3 + i] = cofactor3(a, i, j) * invdet;
    return r;
}

// ----------------------------------------------------------------------------- margins
// Every accept/reject comparison on the path records |a-b| so that a GPU disagreement can be
// attributed to a near-threshold case.
struct Margin {
    double m = kInf;
    inline void rec(double a, double b) {
        double d = std::fabs(a - b);
        if (d < m) m = d;  // NaN never lowers the margin
    }
};
inline bool lt(double a, double b, Margin* mg) { if (mg) mg->rec(a, b); return a < b; }
inline bool le(double a, double b, Margin* mg) { if (mg) mg->rec(a, b); return a <= b; }
inline bool gt(double a, double b, Margin* mg) { if (mg) mg->rec(a, b); return a > b; }
inline bool ge(double a, double b, Margin* mg) { if (mg) mg->rec(a, b); return a >= b; }

// ----------------------------------------------------------------------------- two-view geometry
// util.cpp:33-66 (skew, fundamental matrix, relative pose) + the epipoles of getOrientationList.cpp:47-50
struct PairGeom {
    M3 R21; V3 T21; M3 F;  // F = K2^-T [T21]x R21 K1^-1 maps a view-1 pixel to its epipolar line in view 2
    V3 e1pix;              // epipole in view 1 (pixels): K1 * (R21^T T21 normalised)   getOrientationList.cpp:47,49
    V3 e2pix;              // epipole in view 2 (pixels): K2 * (T21 / T21z)             getOrientationList.cpp:48,50
};

inline M3 rel_R21(const M3& R1, const M3& R2) { return mat_mul(R2, transpose(R1)); }  // util.cpp:50-56
inline V3 rel_T21(const M3& R1, const M3& R2, const V3& T1, const V3& T2) {           // util.cpp:58-66
    M3 nR1t = transpose(R1), nR2t = transpose(R2);
    for (int i = 0; i < 9; i++) { nR1t.m[i] = -1 * nR1t.m[i]; nR2t.m[i] = -1 * nR2t.m[i]; }
    V3 C1 = mat_vec(nR1t, T1), C2 = mat_vec(nR2t, T2);
    V3 d = {{C1.v[0] - C2.v[0], C1.v[1] - C2.v[1], C1.v[2] - C2.v[2]}};
    return mat_vec(R2, d);
}
inline M3 fundamental(const M3& invK1, const M3& invK2, const M3& R21, const V3& T21) {  // util.cpp:44-48
    M3 Tx = {{0., -T21.v[2], T21.v[1], T21.v[2], 0., -T21.v[0], -T21.v[1], T21.v[0], 0.}};
    return mat_mul(mat_mul(transpose(invK2), mat_mul(Tx, R21)), invK1);
}
PairGeom pair_geometry(const M3& R1, const V3& T1, const M3& R2, const V3& T2, const M3& K1, const M3& K2) {
    PairGeom g;
    g.R21 = rel_R21(R1, R2);
    g.T21 = rel_T21(R1, R2, T1, T2);
    g.F = fundamental(inverse3(K1), inverse3(K2), g.R21, g.T21);
    V3 rt = mat_vec(transpose(g.R21), g.T21);  // e_i^T R21^T T21
    V3 met1 = {{rt.v[0] / rt.v[2], rt.v[1] / rt.v[2], 1.0}};
    V3 met2 = {{g.T21.v[0] / g.T21.v[2], g.T21.v[1] / g.T21.v[2], 1.0}};
    g.e1pix = mat_vec(K1, met1);
    g.e2pix = mat_vec(K2, met2);
    return g;
}

// ----------------------------------------------------------------------------- a2: epipolar angle of an edge
// getOrientationList.cpp:391-425 (getOreList) / :427-461 (getOreListVali): identical bodies.
inline double ore_angle(double x, double y, const V3& e2pix) {
    const double dy = y - e2pix.v[1];
    const double dx = x - e2pix.v[0];
    const double a = std::atan(dy / dx);
    double deg = (a * 180) / M_PI;
    if (deg < 0) deg = deg + 180;
    return deg;
}

// ----------------------------------------------------------------------------- a3: epipolar wedge
// getOrientationList.cpp:33-353 (getOreListBar), one row.  flag bit0: line 1 crosses != 2 image
// borders, bit1: same for line 2, bit2: a NaN reached the result.  The reference reads stale /
// uninitialised memory in those cases (SURVEY.md a3 "UB"); the oracle defines them as "no wedge".
struct Wedge { double lo, hi; int flag; };

inline bool pick_border_point(double A, double B, double C, int rows, int cols, const V3& e2, double* px, double* py,
                              Margin* mg) {
    // candidates, getOrientationList.cpp:144-170
    const double xs[4] = {0.0, double(cols), -1 * C / A, -1 * (C + B * rows) / A};
    const double ys[4] = {-1 * C / B, -1 * (C + A * cols) / B, 0.0, double(rows)};
    int hits = 0;
    double sx[2] = {0, 0}, sy[2] = {0, 0}, sd[2] = {0, 0};
    for (int k = 0; k < 4; k++) {
        const double v = (k < 2) ? ys[k] : xs[k];
        const double hi = (k < 2) ? double(rows) : double(cols);
        // inclusive range test :203,212,221,229
        const bool in = ge(v, 0.0, mg) && le(v, hi, mg);
        if (in) {
            if (hits < 2) {
                sx[hits] = xs[k]; sy[hits] = ys[k];
                const double ddx = xs[k] - e2.v[0], ddy = ys[k] - e2.v[1];
                sd[hits] = ddx * ddx + ddy * ddy;
            }
            hits++;
        }
    }
    if (hits != 2) return false;
    if (lt(sd[0], sd[1], mg)) { *px = sx[0]; *py = sy[0]; }  // :237
    else { *px = sx[1]; *py = sy[1]; }
    return true;
}

inline void border_slopes(double px, double py, double ore_deg, const V3& e2, double delta, int cols, double* sp,
                          double* sn) {
    const double delta_x = std::fabs(std::cos(ore_deg / 180 * M_PI) * delta);  // :243
    double slope_p, slope_n;
    if (px == 0 || px == cols) {  // :246
        const double fdx = px - e2.v[0];
        const double fdyp = py - e2.v[1] - delta_x;
        const double fdyn = py - e2.v[1] + delta_x;
        slope_p = fdyp / fdx;
        slope_n = fdyn / fdx;
    } else {
        const double fdxp = px - e2.v[0] - delta_x;
        const double fdxn = px - e2.v[0] + delta_x;
        const double fdy = py - e2.v[1];
        slope_p = fdy / fdxp;
        slope_n = fdy / fdxn;
    }
    *sp = std::atan(slope_p) / M_PI * 180;
    *sn = std::atan(slope_n) / M_PI * 180;
}

Wedge wedge_row(double x, double y, const PairGeom& g, double delta, int rows, int cols, Margin* mg) {
    Wedge w; w.flag = 0;
    const double dy = y - g.e1pix.v[1];  // :54
    const double dx = x - g.e1pix.v[0];  // :55
    const double slope = dy / dx;        // :57
    const double d2 = dy * dy + dx * dx; // :58
    const double d12sq = d2 - delta * delta;  // :59
    const double d12 = std::sqrt(d12sq);      // :60
    const double th = std::asin(delta / d12); // :62
    const double ang = std::atan(slope);      // :63
    const double a1 = ang + th, a2 = ang - th;  // :66-67
    double dx1 = std::fabs(std::cos(a1) * d12), dy1 = std::fabs(std::sin(a1) * d12);  // :73-74
    double dx2 = std::fabs(std::cos(a2) * d12), dy2 = std::fabs(std::sin(a2) * d12);  // :75-76
    if (lt(dx, 0.0, mg)) { dx1 = -dx1; dx2 = -dx2; }  // :87
    if (lt(dy, 0.0, mg)) { dy1 = -dy1; dy2 = -dy2; }  // :92
    const double p1x = dx1 + g.e1pix.v[0], p1y = dy1 + g.e1pix.v[1];  // :102-103
    const double p2x = dx2 + g.e1pix.v[0], p2y = dy2 + g.e1pix.v[1];  // :105-106
    const double* F = g.F.m;
    const double A1 = dot3(F[0], F[1], F[2], p1x, p1y, 1.0), B1 = dot3(F[3], F[4], F[5], p1x, p1y, 1.0),
                 C1 = dot3(F[6], F[7], F[8], p1x, p1y, 1.0);  // :109-113
    const double A2 = dot3(F[0], F[1], F[2], p2x, p2y, 1.0), B2 = dot3(F[3], F[4], F[5], p2x, p2y, 1.0),
                 C2 = dot3(F[6], F[7], F[8], p2x, p2y, 1.0);  // :115-119
    double ore1 = std::atan(B1 / A1) / M_PI * 180;  // :120-121
    double ore2 = std::atan(B2 / A2) / M_PI * 180;  // :123-124
    if (lt(ore1, 0.0, mg)) ore1 = ore1 + 180;       // :128
    if (lt(ore2, 0.0, mg)) ore2 = ore2 + 180;       // :132
    double ore_lo, ore_hi;
    if (lt(ore1, ore2, mg)) { ore_lo = ore1; ore_hi = ore2; } else { ore_lo = ore2; ore_hi = ore1; }  // :135-139
    double q1x, q1y, q2x, q2y;
    if (!pick_border_point(A1, B1, C1, rows, cols, g.e2pix, &q1x, &q1y, mg)) w.flag |= 1;  // :200-241
    if (!pick_border_point(A2, B2, C2, rows, cols, g.e2pix, &q2x, &q2y, mg)) w.flag |= 2;  // :266-307
    if (w.flag) { w.lo = kNaN; w.hi = kNaN; return w; }
    double s[4];
    border_slopes(q1x, q1y, ore_lo, g.e2pix, delta, cols, &s[0], &s[1]);  // :243-264 (line 1 uses the MIN angle)
    border_slopes(q2x, q2y, ore_hi, g.e2pix, delta, cols, &s[2], &s[3]);  // :309-330 (line 2 uses the MAX angle)
    for (int k = 0; k < 4; k++)
        if (lt(s[k], 0.0, mg)) s[k] = s[k] + 180;  // :337-348
    if (std::isnan(s[0]) || std::isnan(s[1]) || std::isnan(s[2]) || std::isnan(s[3])) {
        w.flag |= 4; w.lo = kNaN; w.hi = kNaN; return w;
    }
    double lo = s[0], hi = s[0];
    for (int k = 1; k < 4; k++) { if (s[k] < lo) lo = s[k]; if (s[k] > hi) hi = s[k]; }  // :349-350
    w.lo = lo; w.hi = hi;
    return w;
}

// ----------------------------------------------------------------------------- a6: epipolar correction
// util.cpp:68-106 + PairEdgeHypo.cpp:225-293
inline double normal_dist(const double c[3], double x, double y, double* ex, double* ey) {  // util.cpp:81-88
    const double a1 = c[0], b1 = c[1], c1 = c[2];
    *ex = x - a1 * (a1 * x + b1 * y + c1) / (a1 * a1 + b1 * b1);
    *ey = y - b1 * (a1 * x + b1 * y + c1) / (a1 * a1 + b1 * b1);
    return std::sqrt((x - *ex) * (x - *ex) + (y - *ey) * (y - *ey));
}
inline double tangential_dist(const double c[3], double x, double y, double theta, double* xi, double* yi) {  // util.cpp:96-106
    const double a_e = std::tan(theta);
    const double b_e = -1;
    const double c_e = -(a_e * x - y);
    const double a1 = c[0], b1 = c[1], c1 = c[2];
    *xi = (b1 * c_e - b_e * c1) / (a1 * b_e - a_e * b1);
    *yi = (c1 * a_e - c_e * a1) / (a1 * b_e - a_e * b1);
    return std::sqrt((*xi - x) * (*xi - x) + (*yi - y) * (*yi - y));
}
inline void epip_line_coeffs(const M3& F, double x1, double y1, double out[3]) {  // util.cpp:68-78
    const double l0 = dot3(F.m[0], F.m[1], F.m[2], x1, y1, 1.0), l1 = dot3(F.m[3], F.m[4], F.m[5], x1, y1, 1.0),
                 l2 = dot3(F.m[6], F.m[7], F.m[8], x1, y1, 1.0);
    out[0] = -l0 / l1; out[1] = -1.0; out[2] = -l2 / l1;
}
// returns true when the H2 edge survives; corrected location in (cx, cy).  PairEdgeHypo.cpp:235-287
inline bool correct_one(const double c[3], double x, double y, double theta, const orc_config& cfg, double* cx,
                        double* cy, Margin* mg) {
    double ex, ey;
    const double nd = normal_dist(c, x, y, &ex, &ey);
    if (lt(nd, cfg.location_perturbation, mg)) { *cx = ex; *cy = ey; return true; }  // :243
    double xi, yi;
    double d = tangential_dist(c, x, y, theta, &xi, &yi);
    if (lt(d, cfg.epip_tangency_displ_thresh, mg)) { *cx = xi; *cy = yi; return true; }  // :255
    const double p = c[0] * std::cos(theta) + c[1] * std::sin(theta);     // :263
    const double dp = -c[0] * std::sin(theta) + c[1] * std::cos(theta);   // :264
    if (mg) { mg->rec(p, 0.0); mg->rec(dp, 0.0); }
    if (p > 0 && dp < 0) theta -= cfg.orient_perturbation;       // :267-270
    else if (p < 0 && dp < 0) theta -= cfg.orient_perturbation;
    else if (p > 0 && dp > 0) theta += cfg.orient_perturbation;
    else if (p < 0 && dp > 0) theta += cfg.orient_perturbation;
    d = tangential_dist(c, x, y, theta, &xi, &yi);               // :274
    if (lt(d, cfg.epip_tangency_displ_thresh, mg)) { *cx = xi; *cy = yi; return true; }  // :276
    return false;
}

// ----------------------------------------------------------------------------- a7: clustering
// EdgeClusterer.cpp:39-50,52-58,60-125,127-293
struct ClusterOut {
    std::vector<double> xbar, ybar, thbar;  // rows overwritten by the cluster average
    std::vector<int> label, rep_local;      // rep_local: local row index of the cluster representative
};

struct Clusterer {
    int n;
    std::vector<double> x, y, th;
    std::vector<int> label;
    std::vector<double> avg_orient;  // indexed by label (cluster_avg_orientations)
    const orc_config& cfg;
    Margin* mg;

    Clusterer(int n_, const orc_config& c, Margin* m) : n(n_), cfg(c), mg(m) {}

    double normalize(double o) {  // :39-50
        const double wrap = -90.0 + (cfg.cluster_orient_thresh_deg / 2.0);
        double r = o;
        if (lt(o * (180.0 / M_PI), wrap, mg)) r += M_PI;
        return r;
    }
    int size_of(int lab) { int s = 0; for (int i = 0; i < n; i++) if (label[i] == lab) s++; return s; }  // :52-58
    void gauss_avg(int l1, int l2, double* gx, double* gy, double* gth) {  // :60-125
        double sx = 0, sy = 0; int cnt = 0;
        for (int i = 0; i < n; i++)
            if (label[i] == l1 || (l2 != -1 && label[i] == l2)) { sx += x[i]; sy += y[i]; cnt++; }
        if (cnt == 0) { *gx = 0; *gy = 0; *gth = 0; return; }
        const double cx = sx / cnt, cy = sy / cnt;
        double tot = 0.0;
        for (int i = 0; i < n; i++)
            if (label[i] == l1 || (l2 != -1 && label[i] == l2)) {
                const double dx = x[i] - cx, dy = y[i] - cy;
                tot += std::sqrt(dx * dx + dy * dy);
            }
        const double mean_shift = tot / cnt;
        double swx = 0, swy = 0, swo = 0, tw = 0;
        for (int i = 0; i < n; i++)
            if (label[i] == l1 || (l2 != -1 && label[i] == l2)) {
                const double dx = x[i] - cx, dy = y[i] - cy;
                const double dist = std::sqrt(dx * dx + dy * dy);
                const double q = (dist - mean_shift) / cfg.cluster_gauss_sigma;
                const double w = std::exp(-0.5 * (q * q));
                swx += w * x[i]; swy += w * y[i]; swo += w * th[i]; tw += w;
            }
        *gx = swx / tw; *gy = swy / tw; *gth = swo / tw;
    }
    void run(ClusterOut* out) {  // :127-293
        label.resize(n); avg_orient.assign(n, 0.0);
        for (int i = 0; i < n; i++) { label[i] = i; th[i] = normalize(th[i]); avg_orient[i] = th[i]; }  // :131-135
        const double orient_thr = cfg.cluster_orient_thresh_deg * (M_PI / 180.0);  // util.hpp deg_to_rad
        bool merged = true;
        while (merged) {  // :139-189
            merged = false;
            for (int i = 0; i < n; i++) {
                double min_dist = std::numeric_limits<double>::max();
                int nearest = -1;
                for (int j = 0; j < n; j++) {
                    if (label[i] != label[j]) {
                        const double ddx = x[i] - x[j], ddy = y[i] - y[j];
                        const double dist = std::sqrt(ddx * ddx + ddy * ddy);
                        const double oi = avg_orient[label[i]], oj = avg_orient[label[j]];
                        if (mg) { mg->rec(dist, cfg.cluster_dist_thresh); if (dist < cfg.cluster_dist_thresh) { mg->rec(std::fabs(oi - oj), orient_thr); if (nearest >= 0) mg->rec(dist, min_dist); } }
                        if (dist < min_dist && dist < cfg.cluster_dist_thresh && std::fabs(oi - oj) < orient_thr) {  // :154
                            min_dist = dist; nearest = j;
                        }
                    }
                }
                if (nearest != -1) {
                    const int old_l = label[nearest], new_l = label[i];
                    if (size_of(old_l) + size_of(new_l) <= cfg.max_cluster_size) {  // :166
                        double gx, gy, gth;
                        gauss_avg(old_l, new_l, &gx, &gy, &gth);
                        avg_orient[new_l] = normalize(gth);  // :175-176
                        for (int k = 0; k < n; k++) if (label[k] == old_l) label[k] = new_l;  // :179-183
                        merged = true;
                        break;
                    }
                }
            }
        }
        // :195-198 label -> members in ascending local index; std::map iterates labels in ascending order
        std::map<int, std::vector<int>> l2c;
        for (int i = 0; i < n; i++) l2c[label[i]].push_back(i);
        out->xbar.assign(n, 0); out->ybar.assign(n, 0); out->thbar.assign(n, 0);
        out->label = label; out->rep_local.assign(n, -1);
        for (auto& kv : l2c) {  // :250-290
            const std::vector<int>& cl = kv.second;
            double gx, gy, gth;
            gauss_avg(label[cl[0]], -1, &gx, &gy, &gth);
            double min_dist = std::numeric_limits<double>::max();
            int closest = -1;
            for (size_t k = 0; k < cl.size(); k++) {
                const int idx = cl[k];
                const double ddx = x[idx] - gx, ddy = y[idx] - gy;
                const double dist = std::sqrt(ddx * ddx + ddy * ddy);
                if (mg && closest >= 0) mg->rec(dist, min_dist);
                if (dist < min_dist) { min_dist = dist; closest = idx; }  // :270
            }
            for (size_t k = 0; k < cl.size(); k++) {
                const int idx = cl[k];
                out->xbar[idx] = gx; out->ybar[idx] = gy; out->thbar[idx] = gth; out->rep_local[idx] = closest;
            }
        }
    }
};

// ----------------------------------------------------------------------------- a8: reprojection
// getReprojectedEdgel.cpp:32-43 (getTGT_Meters)
inline V3 tgt_meters(double x, double y, double theta, const M3& Kinv) {
    V3 pt = {{x, y, 1.0}};
    V3 G = mat_vec(Kinv, pt);
    V3 tp = {{std::cos(theta) + x, std::sin(theta) + y, 1.0}};
    V3 tm = mat_vec(Kinv, tp);
    return V3{{tm.v[0] - G.v[0], tm.v[1] - G.v[1], tm.v[2] - G.v[2]}};
}
// getReprojectedEdgel.cpp:132-176, one row; writes the reprojected unit tangent (t3x, t3y)
inline void reproject_tangent(double x1, double y1, double th1, double x2, double y2, double th2, const M3& K1inv,
                              const M3& K2inv, const M3& R21, const V3& T21, const M3& R31, const V3& T31,
                              double* t3x, double* t3y) {
    V3 p1 = {{x1, y1, 1.0}};
    V3 G1 = mat_vec(K1inv, p1);
    V3 t1 = tgt_meters(x1, y1, th1, K1inv);
    V3 p2 = {{x2, y2, 1.0}};
    V3 G2 = mat_vec(K2inv, p2);
    V3 RG1 = mat_vec(R21, G1);
    const double rho1 = (T21.v[0] - T21.v[2] * G2.v[0]) / (RG1.v[2] * G2.v[0] - RG1.v[0]);  // :160
    V3 R31G1 = mat_vec(R31, G1);
    const double rho3 = rho1 * R31G1.v[2] + T31.v[2];  // :161
    M3 R31s; for (int i = 0; i < 9; i++) R31s.m[i] = R31.m[i] * rho1;
    V3 g3u = mat_vec(R31s, G1);
    const double ir3 = 1 / rho3;
    V3 G3 = {{ir3 * (g3u.v[0] + T31.v[0]), ir3 * (g3u.v[1] + T31.v[1]), ir3 * (g3u.v[2] + T31.v[2])}};  // :162
    V3 t2 = tgt_meters(x2, y2, th2, K2inv);  // :164
    V3 n1 = cross(t1, G1);                    // :165
    V3 n2 = mat_vec(transpose(R21), cross(t2, G2));  // :166
    V3 c = cross(n1, n2);
    const double cn = norm3(c);
    V3 Tv1 = {{c.v[0] / cn, c.v[1] / cn, c.v[2] / cn}};  // :167
    V3 Tv3 = mat_vec(R31, Tv1);                           // :168
    V3 Tt = {{Tv3.v[0] - Tv3.v[2] * G3.v[0], Tv3.v[1] - Tv3.v[2] * G3.v[1], Tv3.v[2] - Tv3.v[2] * G3.v[2]}};  // :169
    const double tn = norm3(Tt);
    *t3x = Tt.v[0] / tn; *t3y = Tt.v[1] / tn;  // :170
}

// ----------------------------------------------------------------------------- a15: triangulation + tangent
// symmetric 4x4 cyclic Jacobi: eigenvector of the smallest eigenvalue (stands in for
// ATA.jacobiSvd(ComputeFullV).matrixV().col(3), util.cpp:307-308; tolerance 1e-6 relative)
void smallest_eigvec4(double a[4][4], double out[4]) {
    double v[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < 4; p++) for (int q = p + 1; q < 4; q++) off += a[p][q] * a[p][q];
        const double dg = (a[0][0] * a[0][0] + a[1][1] * a[1][1]) + (a[2][2] * a[2][2] + a[3][3] * a[3][3]);
        if (off <= 1e-40 * dg) break;   // off-diagonal norm below 1e-20 of the diagonal: converged far below 1 ulp (JacobiSVD itself stops at 2 eps)
        for (int p = 0; p < 3; p++)
            for (int q = p + 1; q < 4; q++) {
                if (a[p][q] == 0.0) continue;
                const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; k++) {
                    const double akp = a[k][p], akq = a[k][q];
                    a[k][p] = c * akp - s * akq; a[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 4; k++) {
                    const double apk = a[p][k], aqk = a[q][k];
                    a[p][k] = c * apk - s * aqk; a[q][k] = s * apk + c * aqk;
                }
                for (int k = 0; k < 4; k++) {
                    const double vkp = v[k][p], vkq = v[k][q];
                    v[k][p] = c * vkp - s * vkq; v[k][q] = s * vkp + c * vkq;
                }
            }
    }
    int best = 0;
    for (int k = 1; k < 4; k++) if (a[k][k] < a[best][best]) best = k;
    for (int k = 0; k < 4; k++) out[k] = v[k][best];
}
// util.cpp:254-318 with N = 2
V3 triangulate2(double x1, double y1, double x2, double y2, const M3& R21, const V3& T21, const M3& Kinv) {
    V3 h1 = {{x1, y1, 1.0}}, h2 = {{x2, y2, 1.0}};
    V3 g1 = mat_vec(Kinv, h1), g2 = mat_vec(Kinv, h2);
    double A[4][4];
    A[0][0] = 0.0; A[0][1] = -1.0; A[0][2] = g1.v[1]; A[0][3] = 0.0;
    A[1][0] = 1.0; A[1][1] = 0.0; A[1][2] = -g1.v[0]; A[1][3] = 0.0;
    const double* r = R21.m;
    const double mx = g2.v[0], my = g2.v[1];
    A[2][0] = my * r[6] - r[3]; A[2][1] = my * r[7] - r[4]; A[2][2] = my * r[8] - r[5]; A[2][3] = my * T21.v[2] - T21.v[1];
    A[3][0] = r[0] - mx * r[6]; A[3][1] = r[1] - mx * r[7]; A[3][2] = r[2] - mx * r[8]; A[3][3] = T21.v[0] - mx * T21.v[2];
    double ATA[4][4];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            double s = A[0][i] * A[0][j];
            for (int k = 1; k < 4; k++) s += A[k][i] * A[k][j];
            ATA[i][j] = s;
        }
    double g[4];
    smallest_eigvec4(ATA, g);
    return V3{{g[0] / g[3], g[1] / g[3], g[2] / g[3]}};
}
// util.cpp:134-167
V3 tangent3d_world(double x1, double y1, double th1, double x2, double y2, double th2, const M3& K1inv,
                   const M3& K2inv, const M3& R1, const M3& R21) {
    V3 G1 = mat_vec(K1inv, V3{{x1, y1, 1.0}});
    V3 G2 = mat_vec(K2inv, V3{{x2, y2, 1.0}});
    V3 t1m = mat_vec(K1inv, V3{{std::cos(th1), std::sin(th1), 0.0}});
    V3 t2m = mat_vec(K2inv, V3{{std::cos(th2), std::sin(th2), 0.0}});
    V3 n1 = cross(t1m, G1);
    V3 n2 = mat_vec(transpose(R21), cross(t2m, G2));
    V3 c = cross(n1, n2);
    const double cn = norm3(c);
    V3 T3 = {{c.v[0] / cn, c.v[1] / cn, c.v[2] / cn}};
    return mat_vec(transpose(R1), T3);
}
inline V3 to_world(const V3& p, const M3& R, const V3& T) {  // util.hpp:98-102
    V3 d = {{p.v[0] - T.v[0], p.v[1] - T.v[1], p.v[2] - T.v[2]}};
    return mat_vec(transpose(R), d);
}

// ----------------------------------------------------------------------------- a12: support test
// getSupportedEdgels.cpp:32-47.  inliers in ascending index order.
inline int support_idx(double t3x, double t3y, const double* cs, const double* sn, const int* inl, int n_inl,
                       double thr, Margin* mg) {
    double prev = 0; int sup = -2;
    for (int k = 0; k < n_inl; k++) {
        const int i = inl[k];
        const double d = std::fabs(t3x * cs[i] + t3y * sn[i]);
        if (mg) { mg->rec(d, thr); if (d >= thr) mg->rec(d, prev); }
        if (d >= thr && d > prev) { prev = d; sup = i; }
    }
    return sup;
}

struct ViewEdges { std::vector<double> x, y, th, s; int n = 0; };

}  // namespace

// ===================================================================================== context
struct orc_ctx {
    orc_config cfg;
    int n_views = 0;
    std::vector<M3> R, K; std::vector<V3> T; bool multi_K = false;
    std::vector<ViewEdges> views;
    int h1 = -1, h2 = -1;
    // pass constants: EdgeSketch_Core.cpp:115-143
    PairGeom g21;            // (h1 -> h2)
    M3 F21;                  // = g21.F (K_HYPO1, K_HYPO2)
    M3 K1, K2, K1inv, K2inv;
    std::vector<int> vali_views;
    std::vector<double> ore_h2;       // getOreList   EdgeSketch_Core.cpp:161
    std::vector<double> wedge_h1;     // getOreListBar EdgeSketch_Core.cpp:163
    std::vector<int32_t> wedge_h1_flag;
    // results of Run_3D_Edge_Sketch
    std::vector<int32_t> m_of_h1, n_of_h1; std::vector<int64_t> off_of_h1; std::vector<double> form_margin;
    std::vector<int32_t> row_h1, row_h2orig, row_local, row_rawat, row_label, row_h2rep, row_count;
    std::vector<double> row_xc, row_yc, row_th, row_xbar, row_ybar, row_thbar;
    std::vector<int32_t> support;  // H x V
    bool have_inter = false;
    std::vector<double> i_wedge31, i_wedge32, i_tgt, i_margin; std::vector<int32_t> i_wflag;
    std::vector<int32_t> vr_h1, vr_row, vr_count, vr_sel;
    std::vector<int32_t> p_h1, p_h2rep, p_row, p_count;
    std::unordered_map<int64_t, std::vector<int>> hypo2_clusters;  // key = h1 * 2^32 + h2
    orc_counts counts;
    const M3& Kof(int v) const { return multi_K ? K[v] : K[0]; }
};

namespace {

struct H1Result {
    int m = 0, n = 0; bool processed = false; double form_margin = kInf;
    std::vector<int32_t> h2orig, local, rawat, label, h2rep, count;
    std::vector<double> xc, yc, th, xbar, ybar, thbar;
    std::vector<int32_t> support;  // n x V
    std::vector<double> wedge31, wedge32, tgt, margin; std::vector<int32_t> wflag;
    std::vector<std::pair<int, int>> valid;   // (row, count) ascending row
    std::vector<int> selected_order;          // rows in paired_edge order (after std::sort + ratio + capacity)
    std::vector<char> valid_selected;
    std::vector<std::pair<int, std::vector<int>>> clusters;  // (orig key, members)
    int64_t n_degenerate = 0, n_supported = 0, n_parallel = 0, n_inliers = 0, n_clusters = 0;
};

// getHYPO2_idx_Ore, PairEdgeHypo.cpp:88-102 (faithful: copies the list, strict both sides)
std::vector<int> scan_wedge_faithful(std::vector<double> ore /*by value*/, double lo, double hi, Margin* mg) {
    std::vector<int> idx;
    std::vector<double> lst(ore.begin(), ore.end());
    for (size_t i = 0; i < lst.size(); i++) {
        if (mg) { mg->rec(lst[i], lo); mg->rec(lst[i], hi); }
        if (lst[i] > lo && lst[i] < hi) idx.push_back(int(i));
    }
    return idx;
}

// Per-view data hoisted out of the H1 loop (mode 1); in faithful mode it is rebuilt per (H1, view)
struct ViewPrep {
    PairGeom g31, g32;            // (h1 -> v), (h2 -> v)
    std::vector<double> ore31, ore32, cs, sn;
};

void build_view_prep(const orc_ctx& o, int v, ViewPrep* p, bool with_lists) {
    p->g31 = pair_geometry(o.R[o.h1], o.T[o.h1], o.R[v], o.T[v], o.K1, o.Kof(v));
    p->g32 = pair_geometry(o.R[o.h2], o.T[o.h2], o.R[v], o.T[v], o.K2, o.Kof(v));
    if (!with_lists) return;
    const ViewEdges& e = o.views[v];
    p->ore31.resize(e.n); p->ore32.resize(e.n); p->cs.resize(e.n); p->sn.resize(e.n);
    for (int i = 0; i < e.n; i++) {
        p->ore31[i] = ore_angle(e.x[i], e.y[i], p->g31.e2pix);   // getOreListVali EdgeSketch_Core.cpp:337
        p->ore32[i] = ore_angle(e.x[i], e.y[i], p->g32.e2pix);   // getOreListVali EdgeSketch_Core.cpp:342
        p->cs[i] = std::cos(e.th[i]); p->sn[i] = std::sin(e.th[i]);  // EdgeSketch_Core.cpp:321-322
    }
}

// One iteration of LOOP 1, EdgeSketch_Core.cpp:209-552
void process_h1(const orc_ctx& o, int H1, int mode, bool keep, const std::vector<ViewPrep>* hoisted, H1Result* r) {
    const orc_config& cfg = o.cfg;
    const ViewEdges& E1 = o.views[o.h1];
    const ViewEdges& E2 = o.views[o.h2];
    const int V = int(o.vali_views.size());
    Margin fm;
    // Skip_this_Edge, EdgeSketch_Core.hpp:92-110
    const double x1 = E1.x[H1], y1 = E1.y[H1], th1 = E1.th[H1];
    if (lt(x1, 10, &fm) || gt(x1, cfg.img_cols - 10, &fm) || lt(y1, 10, &fm) || gt(y1, cfg.img_rows - 10, &fm)) {
        r->form_margin = fm.m; return;
    }
    const double lo = o.wedge_h1[2 * H1], hi = o.wedge_h1[2 * H1 + 1];
    // candidates :223-226
    std::vector<int> cand;
    if (mode == 0) {
        cand = scan_wedge_faithful(o.ore_h2, lo, hi, &fm);
        std::vector<int> again = scan_wedge_faithful(o.ore_h2, lo, hi, nullptr);  // getedgels_HYPO2_Ore, same scan
        (void)again;
    } else {
        for (int i = 0; i < E2.n; i++) {
            fm.rec(o.ore_h2[i], lo); fm.rec(o.ore_h2[i], hi);
            if (o.ore_h2[i] > lo && o.ore_h2[i] < hi) cand.push_back(i);
        }
    }
    r->m = int(cand.size());
    if (cand.empty()) { r->form_margin = fm.m; return; }
    // epipolar correction :247
    double coeffs[3];
    epip_line_coeffs(o.F21, x1, y1, coeffs);
    for (int k = 0; k < r->m; k++) {
        const int i2 = cand[k];
        double cx, cy;
        if (correct_one(coeffs, E2.x[i2], E2.y[i2], E2.th[i2], cfg, &cx, &cy, &fm)) {
            r->h2orig.push_back(i2); r->local.push_back(k);
            r->xc.push_back(cx); r->yc.push_back(cy); r->th.push_back(E2.th[i2]);
        }
    }
    const int n = int(r->h2orig.size());
    r->n = n;
    if (n == 0) { r->form_margin = fm.m; return; }  // :254
    // clustering :274-279
    Clusterer cl(n, cfg, &fm);
    cl.x = r->xc; cl.y = r->yc; cl.th = r->th;
    ClusterOut co;
    cl.run(&co);
    r->xbar = co.xbar; r->ybar = co.ybar; r->thbar = co.thbar; r->label = co.label;
    for (int k = 0; k < n; k++) if (co.label[k] == k) r->n_clusters++;   // a label in use is the index of one of its members
    r->h2rep.resize(n);
    for (int k = 0; k < n; k++) r->h2rep[k] = r->h2orig[co.rep_local[k]];  // corrected(closest, 8)  EdgeClusterer.cpp:282
    r->rawat.resize(n);
    for (int k = 0; k < n; k++) r->rawat[k] = cand[k];  // HYPO2_idx_raw(local_idx) with the POST-correction index  :221
    {   // H2_Clusters :203-238 : key (H1, HYPO2_idx_raw(local)), value = members mapped the same way
        std::map<int, std::vector<int>> l2c;
        for (int k = 0; k < n; k++) l2c[co.label[k]].push_back(k);
        for (auto& kv : l2c) {
            std::vector<int> orig;
            for (int k : kv.second) orig.push_back(cand[k]);
            for (int oi : orig) r->clusters.push_back({oi, orig});
        }
    }
    r->processed = true;
    r->form_margin = fm.m;
    r->support.assign(size_t(n) * V, -2);
    if (keep) {
        r->wedge31.assign(size_t(V) * 2, kNaN); r->wedge32.assign(size_t(n) * V * 2, kNaN);
        r->tgt.assign(size_t(n) * V * 2, kNaN); r->margin.assign(size_t(n) * V, kInf); r->wflag.assign(size_t(n) * V, 0);
    }
    const double thr = std::cos(double(cfg.delta_theta_deg) / 180 * cfg.pi_truncated);  // getSupportedEdgels.cpp:35
    std::vector<int> stack_rows;  // supported_indices_stack(:,0)
    // LOOP 2 :310-398
    for (int vi = 0; vi < V; vi++) {
        const int v = o.vali_views[vi];
        const ViewEdges& E3 = o.views[v];
        ViewPrep local;
        const ViewPrep* vp;
        if (mode == 0) {
            // faithful: per (H1, view) copies and recomputation, EdgeSketch_Core.cpp:315-342
            ViewEdges copy = E3; (void)copy;
            build_view_prep(o, v, &local, true);
            vp = &local;
        } else {
            vp = &(*hoisted)[vi];
        }
        Margin m31;
        const M3 R31 = vp->g31.R21; const V3 T31 = vp->g31.T21;
        // wedge of the H1 edge in the validation view: only row 0 is used (:345)
        Wedge w31 = wedge_row(x1, y1, vp->g31, cfg.delta, cfg.img_rows, cfg.img_cols, &m31);
        if (mode == 0) {  // the reference evaluates it for all n identical rows (:334)
            for (int k = 1; k < n; k++) { Wedge t = wedge_row(x1, y1, vp->g31, cfg.delta, cfg.img_rows, cfg.img_cols, nullptr); (void)t; }
        }
        if (keep) { r->wedge31[2 * vi] = w31.lo; r->wedge31[2 * vi + 1] = w31.hi; }
        std::vector<int> idx31;
        if (!w31.flag) {
            if (mode == 0) {
                idx31 = scan_wedge_faithful(vp->ore31, w31.lo, w31.hi, &m31);                 // :348
                std::vector<int> e31 = scan_wedge_faithful(vp->ore31, w31.lo, w31.hi, nullptr);  // :349 (unused)
                (void)e31;
            } else {
                for (int i = 0; i < E3.n; i++) {
                    m31.rec(vp->ore31[i], w31.lo); m31.rec(vp->ore31[i], w31.hi);
                    if (vp->ore31[i] > w31.lo && vp->ore31[i] < w31.hi) idx31.push_back(i);
                }
            }
        } else {
            r->n_degenerate++;
        }
        // LOOP 3 :355-394
        for (int k = 0; k < n; k++) {
            Margin mv = m31;
            double t3x, t3y;
            reproject_tangent(x1, y1, th1, r->xbar[k], r->ybar[k], r->thbar[k], o.K1inv, o.K2inv, o.g21.R21, o.g21.T21,
                              R31, T31, &t3x, &t3y);  // :330
            Wedge w32 = wedge_row(r->xbar[k], r->ybar[k], vp->g32, cfg.delta, cfg.img_rows, cfg.img_cols, &mv);  // :340
            if (keep) {
                const size_t b = (size_t(k) * V + vi) * 2;
                r->wedge32[b] = w32.lo; r->wedge32[b + 1] = w32.hi; r->tgt[b] = t3x; r->tgt[b + 1] = t3y;
                r->wflag[size_t(k) * V + vi] = (w31.flag ? 8 : 0) | w32.flag;
            }
            int sup = -2;
            if (w32.flag) r->n_degenerate++;
            if (!w31.flag && !w32.flag) {
                std::vector<int> inl;
                if (mode == 0) {
                    std::vector<int> idx32 = scan_wedge_faithful(vp->ore32, w32.lo, w32.hi, nullptr);    // :362
                    std::vector<int> e32 = scan_wedge_faithful(vp->ore32, w32.lo, w32.hi, nullptr);      // :363 (unused)
                    (void)e32;
                    // is_Epipolar_Wedges_in_Parallel, EdgeSketch_Core.hpp:112-128 (evaluated after the scans, :366)
                    const double d0 = std::fabs(w31.lo - w32.lo), d1 = std::fabs(w31.lo - w32.hi),
                                 d2 = std::fabs(w31.hi - w32.lo), d3 = std::fabs(w31.hi - w32.hi);
                    const double dmax = std::max(std::max(d0, d1), std::max(d2, d3));
                    if (le(dmax, cfg.parallel_angle_deg, &mv)) { r->n_parallel++; goto row_done; }
                    std::set_intersection(idx31.begin(), idx31.end(), idx32.begin(), idx32.end(), std::back_inserter(inl));  // :369-374
                    for (int i : idx31) { mv.rec(vp->ore32[i], w32.lo); mv.rec(vp->ore32[i], w32.hi); }
                } else {
                    const double d0 = std::fabs(w31.lo - w32.lo), d1 = std::fabs(w31.lo - w32.hi),
                                 d2 = std::fabs(w31.hi - w32.lo), d3 = std::fabs(w31.hi - w32.hi);
                    const double dmax = std::max(std::max(d0, d1), std::max(d2, d3));
                    if (le(dmax, cfg.parallel_angle_deg, &mv)) { r->n_parallel++; goto row_done; }
                    for (int i : idx31) {
                        mv.rec(vp->ore32[i], w32.lo); mv.rec(vp->ore32[i], w32.hi);
                        if (vp->ore32[i] > w32.lo && vp->ore32[i] < w32.hi) inl.push_back(i);
                    }
                }
                r->n_inliers += int64_t(inl.size());
                sup = support_idx(t3x, t3y, vp->cs.data(), vp->sn.data(), inl.data(), int(inl.size()), thr, &mv);  // :381
            }
        row_done:
            r->support[size_t(k) * V + vi] = sup;
            if (keep) r->margin[size_t(k) * V + vi] = mv.m;
            if (sup != -2) { stack_rows.push_back(k); r->n_supported++; }  // :387-393
        }
    }
    // vote :402-443
    r->count.assign(n, 0);
    for (int k : stack_rows) r->count[k]++;
    for (int k = 0; k < n; k++)
        if (r->count[k] > 0 && r->count[k] >= cfg.max_support_views) r->valid.push_back({k, r->count[k]});
    if (r->valid.empty()) return;
    // Lowe ratio :480-508
    std::vector<std::pair<int, int>> vp = r->valid;  // (valid_idx -> here the row itself, count)
    std::sort(vp.begin(), vp.end(), [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.second > b.second; });
    const int highest = vp[0].second;
    const double threshold_value = highest * cfg.lowes_ratio;
    std::vector<int> sel;
    for (auto& pr : vp) if (pr.second > threshold_value) sel.push_back(pr.first);
    if (sel.empty()) sel.push_back(vp[0].first);
    // capacity N rows per H1 edge :515-521
    for (int row : sel)
        if (int(r->selected_order.size()) < cfg.n_views) r->selected_order.push_back(row);
    r->valid_selected.assign(r->valid.size(), 0);
    for (size_t q = 0; q < r->valid.size(); q++)
        for (int row : r->selected_order) if (row == r->valid[q].first) r->valid_selected[q] = 1;
}

}  // namespace

// ===================================================================================== C API
extern "C" {

void orc_default_config(orc_config* c) {
    std::memset(c, 0, sizeof(*c));
    c->delta = 0.4; c->delta_theta_deg = 15; c->parallel_angle_deg = 15;
    c->lowes_ratio = 0.6; c->epip_tangency_displ_thresh = 3; c->location_perturbation = 0.4;
    c->orient_perturbation = 0.174533; c->cluster_dist_thresh = 1; c->cluster_orient_thresh_deg = 20.0;
    c->cluster_gauss_sigma = 2.0; c->pi_truncated = 3.1415926;
    c->n_views = 50; c->img_rows = 800; c->img_cols = 800; c->max_support_views = 4; c->max_cluster_size = 10;
}

orc_ctx* orc_create(const orc_config* c) {
    orc_ctx* o = new orc_ctx();
    o->cfg = *c; o->n_views = c->n_views; o->views.resize(c->n_views);
    std::memset(&o->counts, 0, sizeof(o->counts));
    return o;
}
void orc_destroy(orc_ctx* o) { delete o; }

int orc_set_cameras(orc_ctx* o, int n, const double* R, const double* T, const double* K, int k_per_view) {
    if (n != o->n_views) return -1;
    o->R.resize(n); o->T.resize(n);
    for (int i = 0; i < n; i++) { std::memcpy(o->R[i].m, R + 9 * i, 72); std::memcpy(o->T[i].v, T + 3 * i, 24); }
    o->multi_K = k_per_view != 0;
    o->K.resize(o->multi_K ? n : 1);
    for (size_t i = 0; i < o->K.size(); i++) std::memcpy(o->K[i].m, K + 9 * i, 72);
    return 0;
}
int orc_set_view_edges(orc_ctx* o, int view, int n, const double* x, const double* y, const double* th, const double* s) {
    if (view < 0 || view >= o->n_views) return -1;
    ViewEdges& e = o->views[view];
    e.n = n; e.x.assign(x, x + n); e.y.assign(y, y + n); e.th.assign(th, th + n);
    if (s) e.s.assign(s, s + n); else e.s.assign(n, 1.0);
    return 0;
}

// Set_Hypothesis_Views_Camera + Set_Hypothesis_Views_Edgels, EdgeSketch_Core.cpp:115-166
int orc_set_hypothesis_pair(orc_ctx* o, int h1, int h2) {
    if (h1 < 0 || h2 < 0 || h1 >= o->n_views || h2 >= o->n_views || h1 == h2) return -1;
    o->h1 = h1; o->h2 = h2;
    o->K1 = o->Kof(h1); o->K2 = o->Kof(h2);
    o->K1inv = inverse3(o->K1); o->K2inv = inverse3(o->K2);
    o->g21 = pair_geometry(o->R[h1], o->T[h1], o->R[h2], o->T[h2], o->K1, o->K2);
    o->F21 = o->g21.F;
    o->vali_views.clear();
    for (int v = 0; v < o->n_views; v++) if (v != h1 && v != h2) o->vali_views.push_back(v);
    const ViewEdges& E1 = o->views[h1]; const ViewEdges& E2 = o->views[h2];
    o->ore_h2.resize(E2.n);
    for (int i = 0; i < E2.n; i++) o->ore_h2[i] = ore_angle(E2.x[i], E2.y[i], o->g21.e2pix);  // :161
    o->wedge_h1.resize(size_t(E1.n) * 2); o->wedge_h1_flag.resize(E1.n);
    for (int i = 0; i < E1.n; i++) {  // :163
        Wedge w = wedge_row(E1.x[i], E1.y[i], o->g21, o->cfg.delta, o->cfg.img_rows, o->cfg.img_cols, nullptr);
        o->wedge_h1[2 * i] = w.lo; o->wedge_h1[2 * i + 1] = w.hi; o->wedge_h1_flag[i] = w.flag;
    }
    return 0;
}

int orc_run_edge_sketch(orc_ctx* o, int h1_begin, int h1_end, int mode, int num_threads, int keep) {
    if (o->h1 < 0) return -1;
    const int E1n = o->views[o->h1].n;
    if (h1_begin < 0) h1_begin = 0;
    if (h1_end < 0 || h1_end > E1n) h1_end = E1n;
    const int V = int(o->vali_views.size());
    std::vector<ViewPrep> hoisted;
    if (mode != 0) {
        hoisted.resize(V);
        for (int vi = 0; vi < V; vi++) build_view_prep(*o, o->vali_views[vi], &hoisted[vi], true);
    }
    std::vector<H1Result> res(E1n);
#ifdef _OPENMP
    if (num_threads > 0) omp_set_num_threads(num_threads);
    const int chunk = std::max(1, num_threads);
#pragma omp parallel for schedule(static, chunk)   // EdgeSketch_Core.cpp:192-198
#endif
    for (int H1 = h1_begin; H1 < h1_end; H1++) process_h1(*o, H1, mode, keep != 0, &hoisted, &res[H1]);

    // ---- sequential assembly (outputs are independent of the thread count)
    orc_counts& c = o->counts; std::memset(&c, 0, sizeof(c));
    c.n_h1_total = E1n;
    o->m_of_h1.assign(E1n, 0); o->n_of_h1.assign(E1n, 0); o->off_of_h1.assign(E1n, -1); o->form_margin.assign(E1n, kInf);
    o->row_h1.clear(); o->row_h2orig.clear(); o->row_local.clear(); o->row_rawat.clear(); o->row_label.clear();
    o->row_h2rep.clear(); o->row_count.clear(); o->row_xc.clear(); o->row_yc.clear(); o->row_th.clear();
    o->row_xbar.clear(); o->row_ybar.clear(); o->row_thbar.clear(); o->support.clear();
    o->i_wedge31.clear(); o->i_wedge32.clear(); o->i_tgt.clear(); o->i_margin.clear(); o->i_wflag.clear();
    o->vr_h1.clear(); o->vr_row.clear(); o->vr_count.clear(); o->vr_sel.clear();
    o->p_h1.clear(); o->p_h2rep.clear(); o->p_row.clear(); o->p_count.clear();
    o->hypo2_clusters.clear();
    o->have_inter = keep != 0;
    int64_t off = 0;
    for (int H1 = 0; H1 < E1n; H1++) {
        H1Result& r = res[H1];
        o->m_of_h1[H1] = r.m; o->n_of_h1[H1] = r.processed ? r.n : (r.m > 0 ? r.n : 0); o->form_margin[H1] = r.form_margin;
        c.n_candidates += r.m;
        if (!r.processed) continue;
        o->off_of_h1[H1] = off;
        c.n_h1_processed++; c.n_hypotheses += r.n; c.n_validations += int64_t(r.n) * V;
        c.n_degenerate_wedges += r.n_degenerate; c.n_supported += r.n_supported; c.n_parallel += r.n_parallel; c.n_inliers += r.n_inliers; c.n_unique_hypotheses += r.n_clusters;
        for (int k = 0; k < r.n; k++) {
            o->row_h1.push_back(H1); o->row_h2orig.push_back(r.h2orig[k]); o->row_local.push_back(r.local[k]);
            o->row_rawat.push_back(r.rawat[k]); o->row_label.push_back(r.label[k]); o->row_h2rep.push_back(r.h2rep[k]);
            o->row_count.push_back(r.count.empty() ? 0 : r.count[k]);
            o->row_xc.push_back(r.xc[k]); o->row_yc.push_back(r.yc[k]); o->row_th.push_back(r.th[k]);
            o->row_xbar.push_back(r.xbar[k]); o->row_ybar.push_back(r.ybar[k]); o->row_thbar.push_back(r.thbar[k]);
        }
        o->support.insert(o->support.end(), r.support.begin(), r.support.end());
        if (keep) {
            o->i_wedge31.insert(o->i_wedge31.end(), r.wedge31.begin(), r.wedge31.end());
            o->i_wedge32.insert(o->i_wedge32.end(), r.wedge32.begin(), r.wedge32.end());
            o->i_tgt.insert(o->i_tgt.end(), r.tgt.begin(), r.tgt.end());
            o->i_margin.insert(o->i_margin.end(), r.margin.begin(), r.margin.end());
            o->i_wflag.insert(o->i_wflag.end(), r.wflag.begin(), r.wflag.end());
        }
        for (size_t q = 0; q < r.valid.size(); q++) {
            o->vr_h1.push_back(H1); o->vr_row.push_back(int32_t(off + r.valid[q].first));
            o->vr_count.push_back(r.valid[q].second); o->vr_sel.push_back(r.valid_selected[q]);
        }
        c.n_valid_rows += int64_t(r.valid.size());
        for (int row : r.selected_order) {
            o->p_h1.push_back(H1); o->p_h2rep.push_back(r.h2rep[row]); o->p_row.push_back(int32_t(off + row));
            o->p_count.push_back(r.count[row]);
        }
        c.n_pairs += int64_t(r.selected_order.size());
        for (auto& kv : r.clusters) o->hypo2_clusters[(int64_t(H1) << 32) | uint32_t(kv.first)] = kv.second;  // :559-566
        off += r.n;
    }
    return 0;
}

void orc_get_counts(orc_ctx* o, orc_counts* out) { *out = o->counts; }

int orc_get_h2_angles(orc_ctx* o, double* out) { std::memcpy(out, o->ore_h2.data(), o->ore_h2.size() * 8); return int(o->ore_h2.size()); }
int orc_get_h1_wedges(orc_ctx* o, double* out, int32_t* flag) {
    std::memcpy(out, o->wedge_h1.data(), o->wedge_h1.size() * 8);
    if (flag) std::memcpy(flag, o->wedge_h1_flag.data(), o->wedge_h1_flag.size() * 4);
    return int(o->wedge_h1_flag.size());
}
int orc_get_h1_table(orc_ctx* o, int32_t* m, int32_t* n, int64_t* row_off, double* fm) {
    const size_t E = o->m_of_h1.size();
    if (m) std::memcpy(m, o->m_of_h1.data(), E * 4);
    if (n) std::memcpy(n, o->n_of_h1.data(), E * 4);
    if (row_off) std::memcpy(row_off, o->off_of_h1.data(), E * 8);
    if (fm) std::memcpy(fm, o->form_margin.data(), E * 8);
    return int(E);
}
#define CPY(dst, src) do { if (dst) std::memcpy(dst, (src).data(), (src).size() * sizeof((src)[0])); } while (0)
int orc_get_rows(orc_ctx* o, int32_t* h1, int32_t* h2_orig, int32_t* local_idx, int32_t* raw_at_local, double* xc,
                 double* yc, double* theta_orig, double* xbar, double* ybar, double* thbar, int32_t* label,
                 int32_t* h2_rep, int32_t* count) {
    CPY(h1, o->row_h1); CPY(h2_orig, o->row_h2orig); CPY(local_idx, o->row_local); CPY(raw_at_local, o->row_rawat);
    CPY(xc, o->row_xc); CPY(yc, o->row_yc); CPY(theta_orig, o->row_th); CPY(xbar, o->row_xbar); CPY(ybar, o->row_ybar);
    CPY(thbar, o->row_thbar); CPY(label, o->row_label); CPY(h2_rep, o->row_h2rep); CPY(count, o->row_count);
    return int(o->row_h1.size());
}
int orc_get_support_matrix(orc_ctx* o, int32_t* out) { CPY(out, o->support); return int(o->support.size()); }
int orc_get_intermediates(orc_ctx* o, double* w31, double* w32, double* tgt, double* mg, int32_t* wflag) {
    if (!o->have_inter) return -1;
    CPY(w31, o->i_wedge31); CPY(w32, o->i_wedge32); CPY(tgt, o->i_tgt); CPY(mg, o->i_margin); CPY(wflag, o->i_wflag);
    return 0;
}
int orc_get_valid_rows(orc_ctx* o, int32_t* h1, int32_t* row, int32_t* count, int32_t* sel) {
    CPY(h1, o->vr_h1); CPY(row, o->vr_row); CPY(count, o->vr_count); CPY(sel, o->vr_sel);
    return int(o->vr_h1.size());
}
int orc_get_pairs(orc_ctx* o, int32_t* h1, int32_t* h2rep, int32_t* row, int32_t* count, int32_t* supports) {
    CPY(h1, o->p_h1); CPY(h2rep, o->p_h2rep); CPY(row, o->p_row); CPY(count, o->p_count);
    if (supports) {
        const int V = int(o->vali_views.size());
        for (size_t p = 0; p < o->p_row.size(); p++)
            std::memcpy(supports + p * V, o->support.data() + size_t(o->p_row[p]) * V, size_t(V) * 4);
    }
    return int(o->p_h1.size());
}

// Finalize_Edge_Pairs_and_Reconstruct_3D_Edges geometry, EdgeSketch_Core.cpp:625-676, + Stack_3D_Edges :742-748
int orc_finalize(orc_ctx* o, double* gamma_cam, double* gamma_world, double* tangent_world, double* h2c, int32_t* nanf) {
    const ViewEdges& E1 = o->views[o->h1]; const ViewEdges& E2 = o->views[o->h2];
    const size_t P = o->p_h1.size();
    for (size_t p = 0; p < P; p++) {
        const int a = o->p_h1[p], b = o->p_h2rep[p];
        double coeffs[3];
        epip_line_coeffs(o->F21, E1.x[a], E1.y[a], coeffs);
        double cx = kNaN, cy = kNaN;
        const bool ok = correct_one(coeffs, E2.x[b], E2.y[b], E2.th[b], o->cfg, &cx, &cy, nullptr);  // :636
        V3 G = {{0, 0, 0}}; int nf = 0;
        if (ok) G = triangulate2(E1.x[a], E1.y[a], cx, cy, o->g21.R21, o->g21.T21, o->K1inv);  // :656 (K_HYPO1 for both points)
        else nf = 2;
        if (std::isnan(G.v[0]) || std::isnan(G.v[1]) || std::isnan(G.v[2])) { G = V3{{0, 0, 0}}; nf |= 1; }  // :658-663
        V3 W = to_world(G, o->R[o->h1], o->T[o->h1]);  // :666, :745
        V3 tw = tangent3d_world(E1.x[a], E1.y[a], E1.th[a], E2.x[b], E2.y[b], E2.th[b], o->K1inv, o->K2inv, o->R[o->h1],
                                o->g21.R21);  // :675 (uncorrected original rows)
        for (int k = 0; k < 3; k++) {
            if (gamma_cam) gamma_cam[p * 3 + k] = G.v[k];
            if (gamma_world) gamma_world[p * 3 + k] = W.v[k];
            if (tangent_world) tangent_world[p * 3 + k] = tw.v[k];
        }
        if (h2c) { h2c[p * 2] = cx; h2c[p * 2 + 1] = cy; }
        if (nanf) nanf[p] = nf;
    }
    return int(P);
}

int orc_get_pair_cluster(orc_ctx* o, int p, int32_t* members, int cap) {  // EdgeSketch_Core.hpp:193-202
    if (p < 0 || size_t(p) >= o->p_h1.size()) return -1;
    auto it = o->hypo2_clusters.find((int64_t(o->p_h1[p]) << 32) | uint32_t(o->p_h2rep[p]));
    if (it == o->hypo2_clusters.end()) { if (cap > 0) members[0] = o->p_h2rep[p]; return 1; }
    const int n = int(it->second.size());
    for (int k = 0; k < n && k < cap; k++) members[k] = it->second[k];
    return n;
}

// EdgeSketch_Core.cpp:682-726 (which (image, edge) records a pass inserts) + :785-797 (one set per view)
int orc_accumulate_claims(orc_ctx* o, uint8_t** masks) {
    const ViewEdges& E1 = o->views[o->h1]; const ViewEdges& E2 = o->views[o->h2];
    const size_t P = o->p_h1.size();
    const int N = o->cfg.n_views, V = N - 2;
    int n_inserting = 0;
    for (size_t p = 0; p < P; p++) {
        const int a = o->p_h1[p], b = o->p_h2rep[p];
        double coeffs[3];
        epip_line_coeffs(o->F21, E1.x[a], E1.y[a], coeffs);
        double cx = kNaN, cy = kNaN;
        if (!correct_one(coeffs, E2.x[b], E2.y[b], E2.th[b], o->cfg, &cx, &cy, nullptr)) continue;   // :639-642
        const V3 G = triangulate2(E1.x[a], E1.y[a], cx, cy, o->g21.R21, o->g21.T21, o->K1inv);
        if (std::isnan(G.v[0]) || std::isnan(G.v[1]) || std::isnan(G.v[2])) continue;                 // :658-663
        n_inserting++;
        masks[o->h1][a] |= 1; masks[o->h2][b] |= 1;                                                   // :680-681
        auto it = o->hypo2_clusters.find((int64_t(a) << 32) | uint32_t(b));                           // hpp:193-202
        if (it != o->hypo2_clusters.end())
            for (int m : it->second)
                if (m >= 0 && m < int(E2.x.size()) && m != b) masks[o->h2][m] |= 1;                    // :684-692
        int col = 0;
        for (int v = 0; v < N; v++) {                                                                 // :698-724
            if (v == o->h1 || v == o->h2) continue;
            const int s = o->support[size_t(o->p_row[p]) * size_t(V) + size_t(col)];
            if (s != -2) masks[v][s] |= 1;
            col++;
        }
    }
    return n_inserting;
}

// EdgeSketch_Core.cpp:799-868
int orc_select_next_views(int n_views, const int64_t* claimed, const int64_t* total, const int32_t* history, int n_history,
                          double stop_ratio, int32_t* best1, int32_t* best2, double* avg_ratio) {
    std::vector<double> ratios;
    std::vector<std::pair<int, double>> frames;
    for (int i = 0; i < n_views; i++) {
        const double ratio = static_cast<double>(claimed[i]) / total[i];   // :806
        ratios.push_back(ratio);
        frames.push_back({i, ratio});
    }
    double sum = 0.0;
    for (double r : ratios) sum += r;                                      // std::accumulate(.., 0.0), :813
    const double avg = sum / ratios.size();
    std::sort(frames.begin(), frames.end(), [](const std::pair<int, double>& x, const std::pair<int, double>& y) { return x.second < y.second; });
    auto seen = [&](int v) { int c = 0; for (int k = 0; k < n_history; k++) c += (history[k] == v); return c; };
    auto pick = [&](size_t k, int keep) { return k < frames.size() ? frames[k].first : keep; };   // the reference indexes past the end (UB): keep
    int b1 = frames[0].first, b2 = frames[1].first, counter = 0;
    while (true) {                                                         // :828-854
        const int f1 = seen(b1), f2 = seen(b2);
        if (f1 == 0 && f2 == 0) break;
        if (f1 > 0 && f2 > 0) { b1 = pick(2 + counter, b1); b2 = pick(3 + counter, b2); }
        else if (f1 > 0) { b1 = pick(1 + counter, b1); b2 = pick(2 + counter, b2); }
        else if (f2 > 0) { b2 = pick(2 + counter, b2); }
        counter++;
        if (size_t(counter) > frames.size() - 2) break;
    }
    *best1 = b1; *best2 = b2; *avg_ratio = avg;
    return avg > stop_ratio ? 1 : 0;                                       // :862
}

void orc_pair_geometry(const double* R1, const double* T1, const double* R2, const double* T2, const double* K1,
                       const double* K2, double* R21, double* T21, double* F, double* e1pix, double* e2pix) {
    M3 r1, r2, k1, k2; V3 t1, t2;
    std::memcpy(r1.m, R1, 72); std::memcpy(r2.m, R2, 72); std::memcpy(k1.m, K1, 72); std::memcpy(k2.m, K2, 72);
    std::memcpy(t1.v, T1, 24); std::memcpy(t2.v, T2, 24);
    PairGeom g = pair_geometry(r1, t1, r2, t2, k1, k2);
    std::memcpy(R21, g.R21.m, 72); std::memcpy(T21, g.T21.v, 24); std::memcpy(F, g.F.m, 72);
    std::memcpy(e1pix, g.e1pix.v, 24); std::memcpy(e2pix, g.e2pix.v, 24);
}
void orc_wedge_row(double x, double y, const double* F, const double* e1pix, const double* e2pix, double delta, int rows,
                   int cols, double* lo, double* hi, int32_t* flag) {
    PairGeom g; std::memcpy(g.F.m, F, 72); std::memcpy(g.e1pix.v, e1pix, 24); std::memcpy(g.e2pix.v, e2pix, 24);
    Wedge w = wedge_row(x, y, g, delta, rows, cols, nullptr);
    *lo = w.lo; *hi = w.hi; *flag = w.flag;
}
void orc_triangulate(const double* p1, const double* p2, const double* R21, const double* T21, const double* K, double* out3) {
    M3 r, k; V3 t; std::memcpy(r.m, R21, 72); std::memcpy(k.m, K, 72); std::memcpy(t.v, T21, 24);
    V3 g = triangulate2(p1[0], p1[1], p2[0], p2[1], r, t, inverse3(k));
    std::memcpy(out3, g.v, 24);
}
void orc_tangent3d(const double* e1, const double* e2, const double* K1, const double* K2, const double* R1,
                   const double* T1, const double* R2, const double* T2, double* out3) {
    M3 r1, r2, k1, k2; V3 t1, t2;
    std::memcpy(r1.m, R1, 72); std::memcpy(r2.m, R2, 72); std::memcpy(k1.m, K1, 72); std::memcpy(k2.m, K2, 72);
    std::memcpy(t1.v, T1, 24); std::memcpy(t2.v, T2, 24);
    M3 R21 = rel_R21(r1, r2);
    V3 w = tangent3d_world(e1[0], e1[1], e1[2], e2[0], e2[1], e2[2], inverse3(k1), inverse3(k2), r1, R21);
    std::memcpy(out3, w.v, 24);
}

int orc_correct_one(const orc_config* cfg, const double* F21, double x1, double y1, double x2, double y2, double th2, double* cx, double* cy) {
    M3 F; std::memcpy(F.m, F21, 72);
    double c[3];
    epip_line_coeffs(F, x1, y1, c);
    return correct_one(c, x2, y2, th2, *cfg, cx, cy, nullptr) ? 1 : 0;
}
void orc_reproject_tangent(double x1, double y1, double th1, double x2, double y2, double th2, const double* K1, const double* K2,
                           const double* R21, const double* T21, const double* R31, const double* T31, double* t3) {
    M3 k1, k2, r21, r31; V3 t21, t31;
    std::memcpy(k1.m, K1, 72); std::memcpy(k2.m, K2, 72); std::memcpy(r21.m, R21, 72); std::memcpy(r31.m, R31, 72);
    std::memcpy(t21.v, T21, 24); std::memcpy(t31.v, T31, 24);
    reproject_tangent(x1, y1, th1, x2, y2, th2, inverse3(k1), inverse3(k2), r21, t21, r31, t31, &t3[0], &t3[1]);
}
double orc_ore_angle(double x, double y, double ex, double ey) { V3 e = {{ex, ey, 1.0}}; return ore_angle(x, y, e); }
void orc_inverse3(const double* A, double* out) { M3 a; std::memcpy(a.m, A, 72); M3 r = inverse3(a); std::memcpy(out, r.m, 72); }

// Stand-alone support test over pre-formed hypotheses (BASELINE.json config 5).  Follows
// EdgeSketch_Core.cpp:348-384 + getSupportedEdgels.cpp:32-47 with the per-view lists given.
int orc_support_test(int n_groups, const int64_t* group_off, int V, const int64_t* view_off, const double* ang31,
                     const double* ang32, const double* cs, const double* sn, const double* wedge31,
                     const double* wedge32, const double* tgt, double thr, double parallel_deg, int32_t* support,
                     int num_threads) {
    const int64_t H = group_off[n_groups];
#ifdef _OPENMP
    if (num_threads > 0) omp_set_num_threads(num_threads);
#pragma omp parallel for schedule(dynamic, 16) collapse(2)
#endif
    for (int v = 0; v < V; v++)
        for (int g = 0; g < n_groups; g++) {
            const int64_t b = view_off[v], e = view_off[v + 1];
            const double lo31 = wedge31[(size_t(v) * n_groups + g) * 2], hi31 = wedge31[(size_t(v) * n_groups + g) * 2 + 1];
            std::vector<int> idx31;
            for (int64_t i = b; i < e; i++) if (ang31[i] > lo31 && ang31[i] < hi31) idx31.push_back(int(i - b));
            for (int64_t r = group_off[g]; r < group_off[g + 1]; r++) {
                const double lo32 = wedge32[(size_t(v) * H + r) * 2], hi32 = wedge32[(size_t(v) * H + r) * 2 + 1];
                const double t3x = tgt[(size_t(v) * H + r) * 2], t3y = tgt[(size_t(v) * H + r) * 2 + 1];
                int sup = -2;
                const double d0 = std::fabs(lo31 - lo32), d1 = std::fabs(lo31 - hi32), d2 = std::fabs(hi31 - lo32),
                             d3 = std::fabs(hi31 - hi32);
                const double dmax = std::max(std::max(d0, d1), std::max(d2, d3));
                if (!(dmax <= parallel_deg) && !std::isnan(lo31) && !std::isnan(lo32)) {
                    std::vector<int> inl;
                    for (int i : idx31) if (ang32[b + i] > lo32 && ang32[b + i] < hi32) inl.push_back(i);
                    sup = support_idx(t3x, t3y, cs + b, sn + b, inl.data(), int(inl.size()), thr, nullptr);
                }
                support[size_t(v) * H + r] = sup;
            }
        }
    return 0;
}

}  // extern "C"
