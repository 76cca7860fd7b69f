// =====================================================================================
// cvx_math.cuh -- per-element FP64 geometry of the edge-sketch path (device functions).
//
// Two kinds of functions live here.  (1) The reference-order transcriptions (ore_angle, wedge_row, epipolar correction, Gaussian
// average, reprojection, triangulation, ...): plain IEEE double arithmetic in the reference's operation order, compiled with
// -fmad=false, so +,-,*,/,sqrt results are bit-identical to an un-fused CPU build and only atan/asin/sin/cos/tan/exp (CUDA
// libdevice vs glibc, <= 2 ulp) can differ.  (2) wedge_row_fast, the validation kernel's default: an algebraic rewrite of
// getOreListBar that is NOT in reference operation order (2 of its 11 libm calls remain, reciprocals shared); its bounds agree with
// wedge_row to < 2e-12 degrees, which the parity rule absorbs like the libm difference (tests/test_device_math_host.py,
// tests/test_gpu_parity.py::test_strict_libm_order_switch); cvx_config.strict_libm_order = 1 selects (1) in the kernel.
// Reference citations are file:line under Edge_Reconst/ of C-H-Chien/3D_Curvix.
//
// The functions are __host__ __device__ only so that tests/host_harness.cpp can unit-test this
// transcription on a CPU-only box; the product never calls them on the host.
// =====================================================================================
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CVX_HD __host__ __device__ __forceinline__
#else
#define CVX_HD inline
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace cvx {

// Per-pass constants of one hypothesis view pair (EdgeSketch_Core.cpp:115-143), computed on the host.
struct PassConst {
    double F21[9];        // epipolar line in view 2 of a view-1 pixel
    double e1_21[2];      // epipole in view 1 (pixels)            getOrientationList.cpp:47,49
    double e2_21[2];      // epipole in view 2 (pixels)            getOrientationList.cpp:48,50
    double R21[9], T21[3];
    double K1inv[9], K2inv[9];
    double R1[9], T1[3];  // absolute pose of view 1 (world transform of Stack_3D_Edges)
    // thresholds
    double delta, cos_thresh, parallel_deg;
    double loc_pert, tang_thresh, orient_pert;
    double cl_dist, cl_orient_rad, cl_wrap_deg, cl_sigma;
    double cl_sigma_inv;   // 1 / cl_sigma when cl_sigma is a power of two (x / sigma == x * inv exactly), else 0
    double cl_dist_sq;   // smallest double s with sqrt(s) >= cl_dist: "sqrt(s) < cl_dist" == "s < cl_dist_sq" exactly (sqrt is monotone)
    double lowes_ratio;
    int32_t rows, cols, max_support, max_cluster, n_vali, pad;
};

// Per-pass, per-validation-view constants.
struct ViewConst {
    double F31[9], e1_31[2], e2_31[2];   // (h1 -> v): F, epipole in h1, epipole in v
    double F32[9], e1_32[2], e2_32[2];   // (h2 -> v)
    double R31[9], T31[3];
    double pad;
};

CVX_HD double dot3(double a0, double a1, double a2, double b0, double b1, double b2) { return (a0 * b0 + a1 * b1) + a2 * b2; }
CVX_HD void mat_vec(const double* m, double x0, double x1, double x2, double* r) {
    r[0] = dot3(m[0], m[1], m[2], x0, x1, x2);
    r[1] = dot3(m[3], m[4], m[5], x0, x1, x2);
    r[2] = dot3(m[6], m[7], m[8], x0, x1, x2);
}
CVX_HD void mat_t_vec(const double* m, double x0, double x1, double x2, double* r) {  // m^T x
    r[0] = dot3(m[0], m[3], m[6], x0, x1, x2);
    r[1] = dot3(m[1], m[4], m[7], x0, x1, x2);
    r[2] = dot3(m[2], m[5], m[8], x0, x1, x2);
}
CVX_HD void cross3(const double* a, const double* b, double* c) {
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
CVX_HD double norm3(const double* a) { return sqrt((a[0] * a[0] + a[1] * a[1]) + a[2] * a[2]); }
// The libdevice FP64 transcendentals are ~100-200 instructions each; wedge_row alone calls atan seven times.  One shared
// out-of-line copy per function keeps the validation kernel's code footprint inside the instruction cache.
#if defined(__CUDACC__)
static __device__ __noinline__ double cvx_atan_dev(double x) { return atan(x); }
static __device__ __noinline__ double cvx_cos_dev(double x) { return cos(x); }
static __device__ __noinline__ void cvx_sincos_dev(double a, double* s, double* c) { sincos(a, s, c); }
#endif
CVX_HD double cvx_atan(double x) {
#if defined(__CUDA_ARCH__)
    return cvx_atan_dev(x);
#else
    return atan(x);
#endif
}
CVX_HD double cvx_cos(double x) {
#if defined(__CUDA_ARCH__)
    return cvx_cos_dev(x);
#else
    return cos(x);
#endif
}
CVX_HD void sincos_d(double a, double* s, double* c) {
#if defined(__CUDA_ARCH__)
    cvx_sincos_dev(a, s, c);
#else
    *s = sin(a); *c = cos(a);
#endif
}

// ---- epipolar angle of an edge about an epipole: getOrientationList.cpp:391-461 (getOreList / getOreListVali)
CVX_HD double ore_angle(double x, double y, double ex, double ey) {
    const double a = cvx_atan((y - ey) / (x - ex));
    double deg = (a * 180) / M_PI;
    if (deg < 0) deg = deg + 180;
    return deg;
}

// ---- epipolar wedge: getOrientationList.cpp:33-353 (getOreListBar), one row.
// Border crossing of the line A x + B y + C = 0 with the image rectangle; the two crossings are taken in the
// reference's slot order (left, right, top, bottom) and the one closer to the epipole in the target view wins.
CVX_HD bool wedge_border_point(double A, double B, double C, int rows, int cols, double e2x, double e2y, double* px, double* py) {
    const double fr = double(rows), fc = double(cols);
    const double yl = (-C) / B;                // x = 0          :153
    const double yr = (-(C + A * fc)) / B;     // x = cols       :154
    const double xt = (-C) / A;                // y = 0          :148
    const double xb = (-(C + B * fr)) / A;     // y = rows       :149
    int hits = 0;
    double sx0 = 0, sy0 = 0, sx1 = 0, sy1 = 0;
    if (yl >= 0 && yl <= fr) { sx0 = 0.0; sy0 = yl; hits = 1; }                                        // :203
    if (yr >= 0 && yr <= fr) { if (hits == 0) { sx0 = fc; sy0 = yr; } else { sx1 = fc; sy1 = yr; } hits++; }  // :212
    if (xt >= 0 && xt <= fc) { if (hits == 0) { sx0 = xt; sy0 = 0.0; } else if (hits == 1) { sx1 = xt; sy1 = 0.0; } hits++; }  // :221
    if (xb >= 0 && xb <= fc) { if (hits == 0) { sx0 = xb; sy0 = fr; } else if (hits == 1) { sx1 = xb; sy1 = fr; } hits++; }    // :229
    if (hits != 2) return false;   // the reference reads stale memory here; defined as "no wedge"
    const double ax = sx0 - e2x, ay = sy0 - e2y, bx = sx1 - e2x, by = sy1 - e2y;
    const double d0 = ax * ax + ay * ay, d1 = bx * bx + by * by;
    if (d0 < d1) { *px = sx0; *py = sy0; } else { *px = sx1; *py = sy1; }   // :237
    return true;
}
CVX_HD void wedge_border_slopes(double px, double py, double ore_deg, double e2x, double e2y, double delta, int cols, double* sp, double* sn) {
    const double dlt = fabs(cvx_cos(ore_deg / 180 * M_PI) * delta);   // :243
    double slope_p, slope_n;
    if (px == 0 || px == double(cols)) {                          // :246 vertical border: shift in y
        const double fdx = px - e2x;
        slope_p = (py - e2y - dlt) / fdx;
        slope_n = (py - e2y + dlt) / fdx;
    } else {                                                      // shift in x
        const double fdy = py - e2y;
        slope_p = fdy / (px - e2x - dlt);
        slope_n = fdy / (px - e2x + dlt);
    }
    *sp = cvx_atan(slope_p) / M_PI * 180;
    *sn = cvx_atan(slope_n) / M_PI * 180;
}
// returns flag (0 = ok; bit0/bit1: line 1/2 crosses != 2 borders; bit2: NaN) and writes lo/hi (NaN when flagged)
CVX_HD int wedge_row(double x, double y, const double* F, double e1x, double e1y, double e2x, double e2y, double delta, int rows,
                     int cols, double* lo, double* hi) {
    const double dy = y - e1y, dx = x - e1x;                      // :54-55
    const double d12 = sqrt((dy * dy + dx * dx) - delta * delta); // :58-60
    const double th = asin(delta / d12);                          // :62
    const double ang = cvx_atan(dy / dx);                             // :57,63
    double s1, c1, s2, c2;
    sincos_d(ang + th, &s1, &c1);                                 // :66,73-74
    sincos_d(ang - th, &s2, &c2);                                 // :67,75-76
    double dx1 = fabs(c1 * d12), dy1 = fabs(s1 * d12), dx2 = fabs(c2 * d12), dy2 = fabs(s2 * d12);
    if (dx < 0) { dx1 = -dx1; dx2 = -dx2; }                       // :87
    if (dy < 0) { dy1 = -dy1; dy2 = -dy2; }                       // :92
    const double p1x = dx1 + e1x, p1y = dy1 + e1y, p2x = dx2 + e1x, p2y = dy2 + e1y;   // :102-106
    const double A1 = (F[0] * p1x + F[1] * p1y) + F[2], B1 = (F[3] * p1x + F[4] * p1y) + F[5], C1 = (F[6] * p1x + F[7] * p1y) + F[8];  // :109
    const double A2 = (F[0] * p2x + F[1] * p2y) + F[2], B2 = (F[3] * p2x + F[4] * p2y) + F[5], C2 = (F[6] * p2x + F[7] * p2y) + F[8];  // :115
    double ore1 = cvx_atan(B1 / A1) / M_PI * 180;                     // :121
    double ore2 = cvx_atan(B2 / A2) / M_PI * 180;                     // :124
    if (ore1 < 0) ore1 = ore1 + 180;                              // :128
    if (ore2 < 0) ore2 = ore2 + 180;                              // :132
    const double ore_lo = (ore1 < ore2) ? ore1 : ore2, ore_hi = (ore1 < ore2) ? ore2 : ore1;   // :135-139
    double q1x, q1y, q2x, q2y;
    int flag = 0;
    if (!wedge_border_point(A1, B1, C1, rows, cols, e2x, e2y, &q1x, &q1y)) flag |= 1;
    if (!wedge_border_point(A2, B2, C2, rows, cols, e2x, e2y, &q2x, &q2y)) flag |= 2;
    const double qnan = nan("");
    if (flag) { *lo = qnan; *hi = qnan; return flag; }
    double s[4];
    wedge_border_slopes(q1x, q1y, ore_lo, e2x, e2y, delta, cols, &s[0], &s[1]);   // line 1 is shifted with the MIN angle (:243)
    wedge_border_slopes(q2x, q2y, ore_hi, e2x, e2y, delta, cols, &s[2], &s[3]);   // line 2 with the MAX angle (:309)
#pragma unroll
    for (int k = 0; k < 4; k++) if (s[k] < 0) s[k] = s[k] + 180;  // :337-348
    if (s[0] != s[0] || s[1] != s[1] || s[2] != s[2] || s[3] != s[3]) { *lo = qnan; *hi = qnan; return 4; }
    *lo = fmin(fmin(s[0], s[1]), fmin(s[2], s[3]));               // :349
    *hi = fmax(fmax(s[0], s[1]), fmax(s[2], s[3]));               // :350
    return 0;
}

// ---- the same wedge with the libm round trips removed (validation kernel: one call per unique hypothesis x view).
// getOreListBar goes angle -> (asin, atan) -> sincos -> point and slope -> atan -> cos; algebraically
//   cos(atan(t)) = 1/sqrt(1+t^2),  (cos,sin)(atan(dy/dx)) = (|dx|, sgn(dy/dx)|dy|)/r,  (sin,cos)(asin(u)) = (u, sqrt(1-u^2)),
// and the MIN/MAX ordering of the two epipolar-line angles follows from their slopes (atan is monotone).  What remains of libm
// are the four atan that produce the wedge bounds themselves, in the reference's expression.  The bounds agree with wedge_row to
// a few ulp (the same order as the glibc / libdevice difference of the libm chain, ~1e-13 degrees); the degenerate cases (NaN,
// != 2 border crossings) are the same expressions.  cvx_config.strict_libm_order = 1 selects wedge_row in the kernel instead.
CVX_HD bool ore_less(double t1, double t2) {   // ore(t) = atan(t) in [0,90) for t >= 0, atan(t) + 180 in (90,180) for t < 0
    const bool n1 = t1 < 0, n2 = t2 < 0;
    return (n1 == n2) ? (t1 < t2) : n2;
}
// reciprocal / reciprocal square root of the fast wedge (device: rcp.rn.f64 / rsqrt, a third of the instructions of a division)
CVX_HD double cvx_rcp(double x) {
#ifdef __CUDA_ARCH__
    return __drcp_rn(x);
#else
    return 1.0 / x;
#endif
}
CVX_HD double cvx_rsqrt(double x) {
#ifdef __CUDA_ARCH__
    return rsqrt(x);
#else
    return 1.0 / sqrt(x);
#endif
}
// wedge_border_point with the two reciprocals shared by its four border crossings (1 ulp from the divisions of the reference)
CVX_HD bool wedge_border_point_fast(double A, double B, double C, double iA, double iB, int rows, int cols, double e2x, double e2y, double* px,
                                    double* py) {
    const double fr = double(rows), fc = double(cols);
    const double yl = (-C) * iB;               // x = 0          :153
    const double yr = (-(C + A * fc)) * iB;    // x = cols       :154
    const double xt = (-C) * iA;               // y = 0          :148
    const double xb = (-(C + B * fr)) * iA;    // y = rows       :149
    int hits = 0;
    double sx0 = 0, sy0 = 0, sx1 = 0, sy1 = 0;
    if (yl >= 0 && yl <= fr) { sx0 = 0.0; sy0 = yl; hits = 1; }
    if (yr >= 0 && yr <= fr) { if (hits == 0) { sx0 = fc; sy0 = yr; } else { sx1 = fc; sy1 = yr; } hits++; }
    if (xt >= 0 && xt <= fc) { if (hits == 0) { sx0 = xt; sy0 = 0.0; } else if (hits == 1) { sx1 = xt; sy1 = 0.0; } hits++; }
    if (xb >= 0 && xb <= fc) { if (hits == 0) { sx0 = xb; sy0 = fr; } else if (hits == 1) { sx1 = xb; sy1 = fr; } hits++; }
    if (hits != 2) return false;
    const double ax = sx0 - e2x, ay = sy0 - e2y, bx = sx1 - e2x, by = sy1 - e2y;
    const double d0 = ax * ax + ay * ay, d1 = bx * bx + by * by;
    if (d0 < d1) { *px = sx0; *py = sy0; } else { *px = sx1; *py = sy1; }   // :237
    return true;
}
CVX_HD void wedge_border_slopes_fast(double px, double py, double dlt, double e2x, double e2y, int cols, double* sp, double* sn) {
    if (px == 0 || px == double(cols)) {                          // :246 vertical border: shift in y
        const double ifdx = cvx_rcp(px - e2x);
        *sp = (py - e2y - dlt) * ifdx;
        *sn = (py - e2y + dlt) * ifdx;
    } else {                                                      // shift in x
        const double fdy = py - e2y;
        *sp = fdy * cvx_rcp(px - e2x - dlt);
        *sn = fdy * cvx_rcp(px - e2x + dlt);
    }
}
CVX_HD double ore_of_slope(double t) {                            // :337-348: atan / M_PI * 180 (one multiplication here), + 180 if negative
    const double a = cvx_atan(t) * (180.0 / M_PI);
    return a < 0 ? a + 180 : a;
}
CVX_HD int wedge_row_fast(double x, double y, const double* F, double e1x, double e1y, double e2x, double e2y, double delta, int rows,
                          int cols, double* lo, double* hi) {
    const double dy = y - e1y, dx = x - e1x;                      // :54-55
    const double r2 = dy * dy + dx * dx;
    // With d12 = sqrt(r2 - delta^2) (:58-60), th = asin(delta / d12) (:62) and ang = atan(dy / dx) (:57), the tangent points are
    // d12 (cos, sin)(ang +- th) (:66-76).  d12 sin th = delta and d12 cos th = sqrt(d12^2 - delta^2) = sqrt(r2 - 2 delta^2) =: S,
    // so d12 itself drops out.  S is NaN exactly where the reference's sqrt / asin are (edge within sqrt(2) delta of the epipole).
    const double S = sqrt(r2 - 2.0 * (delta * delta));
    const double ir = cvx_rsqrt(r2);
    const double ca = fabs(dx) * ir;                              // cos / sin of atan(dy / dx)
    const double sa = ((dy < 0) != (dx < 0)) ? -(fabs(dy) * ir) : fabs(dy) * ir;
    double dx1 = fabs(ca * S - sa * delta), dy1 = fabs(sa * S + ca * delta);   // ang + th
    double dx2 = fabs(ca * S + sa * delta), dy2 = fabs(sa * S - ca * delta);   // ang - th
    if (dx < 0) { dx1 = -dx1; dx2 = -dx2; }                       // :87
    if (dy < 0) { dy1 = -dy1; dy2 = -dy2; }                       // :92
    const double p1x = dx1 + e1x, p1y = dy1 + e1y, p2x = dx2 + e1x, p2y = dy2 + e1y;   // :102-106
    const double A1 = (F[0] * p1x + F[1] * p1y) + F[2], B1 = (F[3] * p1x + F[4] * p1y) + F[5], C1 = (F[6] * p1x + F[7] * p1y) + F[8];  // :109
    const double A2 = (F[0] * p2x + F[1] * p2y) + F[2], B2 = (F[3] * p2x + F[4] * p2y) + F[5], C2 = (F[6] * p2x + F[7] * p2y) + F[8];  // :115
    const double iA1 = cvx_rcp(A1), iB1 = cvx_rcp(B1), iA2 = cvx_rcp(A2), iB2 = cvx_rcp(B2);
    const double t1 = B1 * iA1, t2 = B2 * iA2;                    // slopes whose atan the reference orders (:121-139)
    const double k1 = delta * cvx_rsqrt(1.0 + t1 * t1), k2 = delta * cvx_rsqrt(1.0 + t2 * t2);   // |cos(atan t)| * delta (:243)
    const bool first_lo = ore_less(t1, t2);
    const double dlt_lo = first_lo ? k1 : k2, dlt_hi = first_lo ? k2 : k1;
    double q1x, q1y, q2x, q2y;
    int flag = 0;
    if (!wedge_border_point_fast(A1, B1, C1, iA1, iB1, rows, cols, e2x, e2y, &q1x, &q1y)) flag |= 1;
    if (!wedge_border_point_fast(A2, B2, C2, iA2, iB2, rows, cols, e2x, e2y, &q2x, &q2y)) flag |= 2;
    const double qnan = nan("");
    if (flag) { *lo = qnan; *hi = qnan; return flag; }
    double m0, m1, m2, m3;                                        // the four slopes whose angles bound the wedge
    wedge_border_slopes_fast(q1x, q1y, dlt_lo, e2x, e2y, cols, &m0, &m1);   // line 1 is shifted with the MIN angle (:243)
    wedge_border_slopes_fast(q2x, q2y, dlt_hi, e2x, e2y, cols, &m2, &m3);   // line 2 with the MAX angle (:309)
    if (m0 != m0 || m1 != m1 || m2 != m2 || m3 != m3) { *lo = qnan; *hi = qnan; return 4; }
    // min / max of the four angles (:349-350) = angles of the extreme slopes in the ore order: two atan instead of four
    const double a = ore_less(m1, m0) ? m1 : m0, b = ore_less(m3, m2) ? m3 : m2;
    const double c = ore_less(m0, m1) ? m1 : m0, d = ore_less(m2, m3) ? m3 : m2;
    *lo = ore_of_slope(ore_less(b, a) ? b : a);
    *hi = ore_of_slope(ore_less(c, d) ? d : c);
    return 0;
}

// ---- epipolar correction of one view-2 candidate: PairEdgeHypo.cpp:225-293, util.cpp:68-106
CVX_HD void epip_line_coeffs(const double* F, double x1, double y1, double* c) {   // util.cpp:68-78
    const double l0 = (F[0] * x1 + F[1] * y1) + F[2], l1 = (F[3] * x1 + F[4] * y1) + F[5], l2 = (F[6] * x1 + F[7] * y1) + F[8];
    c[0] = -l0 / l1; c[1] = -1.0; c[2] = -l2 / l1;
}
CVX_HD double tangential_dist(const double* c, double x, double y, double theta, double* xi, double* yi) {   // util.cpp:96-106
    const double a_e = tan(theta), b_e = -1;
    const double c_e = -(a_e * x - y);
    const double den = c[0] * b_e - a_e * c[1];
    *xi = (c[1] * c_e - b_e * c[2]) / den;
    *yi = (c[2] * a_e - c_e * c[0]) / den;
    return sqrt((*xi - x) * (*xi - x) + (*yi - y) * (*yi - y));
}
CVX_HD bool correct_one(const double* c, double x, double y, double theta, double loc_pert, double tang_thresh, double orient_pert,
                        double* cx, double* cy) {
    const double a1 = c[0], b1 = c[1], c1 = c[2];
    const double q = (a1 * a1 + b1 * b1);
    const double ex = x - a1 * (a1 * x + b1 * y + c1) / q;        // util.cpp:85
    const double ey = y - b1 * (a1 * x + b1 * y + c1) / q;        // util.cpp:86
    const double nd = sqrt((x - ex) * (x - ex) + (y - ey) * (y - ey));
    if (nd < loc_pert) { *cx = ex; *cy = ey; return true; }       // PairEdgeHypo.cpp:243
    double xi, yi;
    double d = tangential_dist(c, x, y, theta, &xi, &yi);
    if (d < tang_thresh) { *cx = xi; *cy = yi; return true; }     // :255
    double st, ct;
    st = sin(theta); ct = cos(theta);
    const double p = a1 * ct + b1 * st;                            // :263
    const double dp = -a1 * st + b1 * ct;                          // :264
    if (p > 0 && dp < 0) theta -= orient_pert;                     // :267-270
    else if (p < 0 && dp < 0) theta -= orient_pert;
    else if (p > 0 && dp > 0) theta += orient_pert;
    else if (p < 0 && dp > 0) theta += orient_pert;
    d = tangential_dist(c, x, y, theta, &xi, &yi);                 // :274
    if (d < tang_thresh) { *cx = xi; *cy = yi; return true; }     // :276
    return false;
}

// ---- reprojection of a hypothesis tangent into a validation view: getReprojectedEdgel.cpp:32-43,132-176
CVX_HD void tgt_meters(double x, double y, double theta, const double* Kinv, const double* G, double* t) {
    double s, c;
    s = sin(theta); c = cos(theta);
    double tm[3];
    mat_vec(Kinv, c + x, s + y, 1.0, tm);
    t[0] = tm[0] - G[0]; t[1] = tm[1] - G[1]; t[2] = tm[2] - G[2];
}
// View-independent part (hoistable per hypothesis row): Gamma1, rho1 and the unit 3-D tangent in the view-1 frame.
struct RowGeom { double G1[3]; double rho1; double Tv1[3]; };
CVX_HD void row_geometry(double x1, double y1, double th1, double x2, double y2, double th2, const PassConst& pc, RowGeom* g) {
    mat_vec(pc.K1inv, x1, y1, 1.0, g->G1);
    double t1[3], G2[3], t2[3];
    tgt_meters(x1, y1, th1, pc.K1inv, g->G1, t1);
    mat_vec(pc.K2inv, x2, y2, 1.0, G2);
    double RG1[3];
    mat_vec(pc.R21, g->G1[0], g->G1[1], g->G1[2], RG1);
    g->rho1 = (pc.T21[0] - pc.T21[2] * G2[0]) / (RG1[2] * G2[0] - RG1[0]);   // :160
    tgt_meters(x2, y2, th2, pc.K2inv, G2, t2);
    double n1[3], c2[3], n2[3], c[3];
    cross3(t1, g->G1, n1);                                                    // :165
    cross3(t2, G2, c2);
    mat_t_vec(pc.R21, c2[0], c2[1], c2[2], n2);                               // :166
    cross3(n1, n2, c);
    const double cn = norm3(c);
    g->Tv1[0] = c[0] / cn; g->Tv1[1] = c[1] / cn; g->Tv1[2] = c[2] / cn;      // :167
}
CVX_HD void reproject_tangent(const RowGeom& g, const double* R31, const double* T31, double* t3x, double* t3y) {
    double R31G1[3];
    mat_vec(R31, g.G1[0], g.G1[1], g.G1[2], R31G1);
    const double rho3 = g.rho1 * R31G1[2] + T31[2];                           // :161
    // Gamma3 = 1/rho3 * ((R31 * rho1) * Gamma1 + T31)                        // :162
    const double r = g.rho1;
    double g3u[3];
    g3u[0] = dot3(R31[0] * r, R31[1] * r, R31[2] * r, g.G1[0], g.G1[1], g.G1[2]);
    g3u[1] = dot3(R31[3] * r, R31[4] * r, R31[5] * r, g.G1[0], g.G1[1], g.G1[2]);
    g3u[2] = dot3(R31[6] * r, R31[7] * r, R31[8] * r, g.G1[0], g.G1[1], g.G1[2]);
    const double ir3 = 1 / rho3;
    const double G3x = ir3 * (g3u[0] + T31[0]), G3y = ir3 * (g3u[1] + T31[1]), G3z = ir3 * (g3u[2] + T31[2]);
    double Tv3[3];
    mat_vec(R31, g.Tv1[0], g.Tv1[1], g.Tv1[2], Tv3);                          // :168
    double Tt[3] = {Tv3[0] - Tv3[2] * G3x, Tv3[1] - Tv3[2] * G3y, Tv3[2] - Tv3[2] * G3z};   // :169
    const double tn = norm3(Tt);
    *t3x = Tt[0] / tn; *t3y = Tt[1] / tn;                                     // :170
}

// ---- two-view linear triangulation: util.cpp:254-318 (N = 2).  The null direction of A^T A is taken from a
// cyclic Jacobi eigen-solve (stands in for JacobiSVD(ComputeFullV).matrixV().col(3)).
CVX_HD void smallest_eigvec4(double a[4][4], double* out) {
    double v[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < 4; p++) for (int q = p + 1; q < 4; q++) off += a[p][q] * a[p][q];
        const double dg = (a[0][0] * a[0][0] + a[1][1] * a[1][1]) + (a[2][2] * a[2][2] + a[3][3] * a[3][3]);
        if (off <= 1e-40 * dg) break;   // off-diagonal norm below 1e-20 of the diagonal: converged far below 1 ulp (JacobiSVD itself stops at 2 eps)
#if defined(__CUDA_ARCH__)
#pragma unroll   // static (p, q): the two 4x4 matrices stay in registers instead of local memory
#endif
        for (int p = 0; p < 3; p++)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
            for (int q = p + 1; q < 4; q++) {
                if (a[p][q] == 0.0) continue;
                const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; k++) { const double x = a[k][p], y = a[k][q]; a[k][p] = c * x - s * y; a[k][q] = s * x + c * y; }
                for (int k = 0; k < 4; k++) { const double x = a[p][k], y = a[q][k]; a[p][k] = c * x - s * y; a[q][k] = s * x + c * y; }
                for (int k = 0; k < 4; k++) { const double x = v[k][p], y = v[k][q]; v[k][p] = c * x - s * y; v[k][q] = s * x + c * y; }
            }
    }
    int best = 0;                                   // smallest diagonal entry, first minimum; static indexing only (registers)
    double bv = a[0][0];
    if (a[1][1] < bv) { bv = a[1][1]; best = 1; }
    if (a[2][2] < bv) { bv = a[2][2]; best = 2; }
    if (a[3][3] < bv) { bv = a[3][3]; best = 3; }
    for (int k = 0; k < 4; k++) out[k] = best == 0 ? v[k][0] : best == 1 ? v[k][1] : best == 2 ? v[k][2] : v[k][3];
}
CVX_HD void triangulate2(double x1, double y1, double x2, double y2, const double* R21, const double* T21, const double* Kinv, double* G) {
    double g1[3], g2[3];
    mat_vec(Kinv, x1, y1, 1.0, g1);
    mat_vec(Kinv, x2, y2, 1.0, g2);
    double A[4][4];
    A[0][0] = 0.0; A[0][1] = -1.0; A[0][2] = g1[1]; A[0][3] = 0.0;
    A[1][0] = 1.0; A[1][1] = 0.0; A[1][2] = -g1[0]; A[1][3] = 0.0;
    const double mx = g2[0], my = g2[1];
    A[2][0] = my * R21[6] - R21[3]; A[2][1] = my * R21[7] - R21[4]; A[2][2] = my * R21[8] - R21[5]; A[2][3] = my * T21[2] - T21[1];
    A[3][0] = R21[0] - mx * R21[6]; A[3][1] = R21[1] - mx * R21[7]; A[3][2] = R21[2] - mx * R21[8]; A[3][3] = T21[0] - mx * T21[2];
    double ATA[4][4];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            double s = A[0][i] * A[0][j];
            for (int k = 1; k < 4; k++) s += A[k][i] * A[k][j];
            ATA[i][j] = s;
        }
    double g[4];
    smallest_eigvec4(ATA, g);
    G[0] = g[0] / g[3]; G[1] = g[1] / g[3]; G[2] = g[2] / g[3];
}
// util.cpp:134-167: 3-D tangent from the two (uncorrected) 2-D edges, in world coordinates
CVX_HD void tangent3d_world(double x1, double y1, double th1, double x2, double y2, double th2, const PassConst& pc, double* tw) {
    double G1[3], G2[3], t1m[3], t2m[3];
    mat_vec(pc.K1inv, x1, y1, 1.0, G1);
    mat_vec(pc.K2inv, x2, y2, 1.0, G2);
    mat_vec(pc.K1inv, cos(th1), sin(th1), 0.0, t1m);
    mat_vec(pc.K2inv, cos(th2), sin(th2), 0.0, t2m);
    double n1[3], c2[3], n2[3], c[3];
    cross3(t1m, G1, n1);
    cross3(t2m, G2, c2);
    mat_t_vec(pc.R21, c2[0], c2[1], c2[2], n2);
    cross3(n1, n2, c);
    const double cn = norm3(c);
    mat_t_vec(pc.R1, c[0] / cn, c[1] / cn, c[2] / cn, tw);
}
CVX_HD void to_world(const double* G, const double* R, const double* T, double* W) {   // util.hpp:98-102
    mat_t_vec(R, G[0] - T[0], G[1] - T[1], G[2] - T[2], W);
}

// ---- clustering helpers: EdgeClusterer.cpp:39-50
CVX_HD double normalize_orientation(double o, double wrap_deg) {
    return (o * (180.0 / M_PI) < wrap_deg) ? o + M_PI : o;
}

}  // namespace cvx
