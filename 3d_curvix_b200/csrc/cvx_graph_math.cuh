// =====================================================================================
// cvx_graph_math.cuh -- per-element FP64 arithmetic of the 3-D edge graph stages (cvx_graph.cu): the closest vector from a point to a line, the
// Rodrigues rotation between two directions, Eigen 3.3's eulerAngles(0, 1, 2), the shortest-alignment rule and the quaternion route back to a
// rotation matrix, in the operation order of the reference (Edge_Reconst/util.cpp:123-131, 186-252) and of Eigen 3.3's un-vectorised double path.
// __host__ __device__ only so that tests/host_harness.cpp can check the transcription against the reference's own functions (oracle/_ref) on a
// CPU-only box; the product never runs it on the host.
// =====================================================================================
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CVXG_HD __host__ __device__ inline
#else
#define CVXG_HD inline
#endif

namespace cvxg {

constexpr double kPI = 3.1415926;                     // definitions.h:54 (the truncated PI of the reference)
constexpr double kMPI = 3.14159265358979323846;       // M_PI / EIGEN_PI

struct D3 { double x, y, z; };
CVXG_HD D3 ld3(const double* p, int64_t i) { return D3{p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }
CVXG_HD D3 sub3(D3 a, D3 b) { return D3{a.x - b.x, a.y - b.y, a.z - b.z}; }
CVXG_HD D3 add3(D3 a, D3 b) { return D3{a.x + b.x, a.y + b.y, a.z + b.z}; }
CVXG_HD D3 mul3(double s, D3 a) { return D3{s * a.x, s * a.y, s * a.z}; }
CVXG_HD double dot3(D3 a, D3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }   // Eigen 3.3 fixed-size dot: left to right
CVXG_HD double norm3(D3 a) { return sqrt(dot3(a, a)); }
CVXG_HD D3 cross3(D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
CVXG_HD D3 normalized3(D3 a) {     // Eigen::normalized(): v / sqrt(squaredNorm) when the norm is positive
    const double z = dot3(a, a);
    if (z > 0.0) { const double n = sqrt(z); return D3{a.x / n, a.y / n, a.z / n}; }
    return a;
}
// util.cpp:123-131: closest point of the line (P1, d1) to P2, minus P2
CVXG_HD D3 closest_vector_to_line(D3 P1, D3 d1, D3 P2) {
    const double a = dot3(sub3(P2, P1), d1);
    return sub3(add3(P1, mul3(a, d1)), P2);
}
struct M33 { double m[3][3]; };
CVXG_HD D3 mat_vec3(const M33& A, D3 v) {
    return D3{(A.m[0][0] * v.x + A.m[0][1] * v.y) + A.m[0][2] * v.z, (A.m[1][0] * v.x + A.m[1][1] * v.y) + A.m[1][2] * v.z,
              (A.m[2][0] * v.x + A.m[2][1] * v.y) + A.m[2][2] * v.z};
}
// util.cpp:226-239 getRodriguesRotationMatrix: I + [v1 x v2]x + 1/(1+c) [v1 x v2]x^2 of the normalised inputs
CVXG_HD M33 rodrigues(D3 v1, D3 v2) {
    const double n1 = norm3(v1), n2 = norm3(v2);
    v1 = D3{v1.x / n1, v1.y / n1, v1.z / n1};
    v2 = D3{v2.x / n2, v2.y / n2, v2.z / n2};
    const D3 w = cross3(v1, v2);
    const double c = dot3(v1, v2);
    const double coeff = 1.0 / (1.0 + c);
    const double vx[3][3] = {{0.0, -w.z, w.y}, {w.z, 0.0, -w.x}, {-w.y, w.x, 0.0}};   // util.cpp:33-36 getSkewSymmetric
    M33 R;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            const double p = ((coeff * vx[i][0]) * vx[0][j] + (coeff * vx[i][1]) * vx[1][j]) + (coeff * vx[i][2]) * vx[2][j];
            R.m[i][j] = ((i == j ? 1.0 : 0.0) + vx[i][j]) + p;
        }
    return R;
}
// Eigen 3.3 MatrixBase::eulerAngles(0, 1, 2), then degrees (util.cpp:219-224 getAlignEulerAnglesDegrees)
CVXG_HD D3 align_euler_deg(D3 v1, D3 v2) {
    const M33 R = rodrigues(v1, v2);
    double r0 = atan2(R.m[1][2], R.m[2][2]);
    const double c2 = sqrt(R.m[0][0] * R.m[0][0] + R.m[0][1] * R.m[0][1]);
    double r1;
    if (r0 > 0.0) {
        if (r0 > 0.0) r0 -= kMPI; else r0 += kMPI;
        r1 = atan2(-R.m[0][2], -c2);
    } else {
        r1 = atan2(-R.m[0][2], c2);
    }
    const double s1 = sin(r0), c1 = cos(r0);
    const double r2 = atan2(s1 * R.m[2][0] - c1 * R.m[1][0], c1 * R.m[1][1] - s1 * R.m[2][1]);
    const double k = 180.0 / kMPI;
    return D3{(-r0) * k, (-r1) * k, (-r2) * k};
}
CVXG_HD bool large_deg(D3 e) { return e.x < -90 || e.x > 90 || e.y < -90 || e.y > 90 || e.z < -90 || e.z > 90; }   // util.cpp:215-217
// util.cpp:186-213 getShortestAlignEulerAnglesDegrees
CVXG_HD D3 shortest_align_euler_deg(D3 v1, D3 v2) {
    const D3 e12 = align_euler_deg(v1, v2), e21 = align_euler_deg(v2, v1);
    if (large_deg(e12) && large_deg(e21)) {
        const D3 v2r = D3{-v2.x, -v2.y, -v2.z};
        const D3 e12r = align_euler_deg(v1, v2r);
        if (large_deg(e12r)) { const D3 e = align_euler_deg(v2r, v1); return D3{e.x * -1, e.y * -1, e.z * -1}; }
        return e12r;
    }
    if (large_deg(e12)) return D3{e21.x * -1, e21.y * -1, e21.z * -1};
    return e12;
}
// util.cpp:241-252 euler_to_rotation_matrix: (yaw about Z) * (pitch about Y) * (roll about X) as quaternions, then to a matrix
struct Quat { double w, x, y, z; };
CVXG_HD Quat qmul(Quat a, Quat b) {
    return Quat{a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
                a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x};
}
CVXG_HD M33 euler_to_rotation(double roll, double pitch, double yaw) {
    const double hr = 0.5 * roll, hp = 0.5 * pitch, hy = 0.5 * yaw;
    const double sr = sin(hr), sp = sin(hp), sy = sin(hy);
    const Quat qr{cos(hr), sr * 1.0, sr * 0.0, sr * 0.0}, qp{cos(hp), sp * 0.0, sp * 1.0, sp * 0.0}, qy{cos(hy), sy * 0.0, sy * 0.0, sy * 1.0};
    const Quat q = qmul(qmul(qy, qp), qr);
    const double tx = 2.0 * q.x, ty = 2.0 * q.y, tz = 2.0 * q.z;
    const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w, txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    M33 R;
    R.m[0][0] = 1.0 - (tyy + tzz); R.m[0][1] = txy - twz; R.m[0][2] = txz + twy;
    R.m[1][0] = txy + twz; R.m[1][1] = 1.0 - (txx + tzz); R.m[1][2] = tyz - twx;
    R.m[2][0] = txz - twy; R.m[2][1] = tyz + twx; R.m[2][2] = 1.0 - (txx + tyy);
    return R;
}

}  // namespace cvxg
