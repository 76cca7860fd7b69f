// Eigen stand-in, part 4: inverse, determinant, Jacobi SVD, Euler angles, AngleAxis / Quaternion, stream output.
// Each routine restates the algorithm Eigen 3.3 documents for it.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include "ops.hpp"

namespace Eigen {

namespace internal {
// cofactor (i,j) of a 3x3, cyclic index form
template <class M> inline typename M::Scalar cofactor3(const M& m, int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return m.coeff(i1, j1) * m.coeff(i2, j2) - m.coeff(i1, j2) * m.coeff(i2, j1);
}
}  // namespace internal

// Matrix3d::inverse(): the first column of cofactors gives det = sum(cofactor_col0 .* col0); result = cofactor^T * (1/det).
// 2x2: adjugate * (1/det).  Other sizes (not on the edge-sketch path): Gauss-Jordan with partial pivoting.
template <class D> inline typename DenseBase<D>::PlainObject DenseBase<D>::inverse() const {
    const D& m = derived();
    EIGEN_SHIM_CHECK(m.rows() == m.cols(), "inverse of a non-square matrix");
    const Index n = m.rows();
    PlainObject r;
    r.resize(n, n);
    if (n == 3) {
        const Scalar c0 = internal::cofactor3(m, 0, 0), c1 = internal::cofactor3(m, 1, 0), c2 = internal::cofactor3(m, 2, 0);
        const Scalar det = internal::ordered_sum<3, Scalar>(3, [&](Index k) { return (k == 0 ? c0 : (k == 1 ? c1 : c2)) * m.coeff(k, 0); });
        const Scalar invdet = Scalar(1) / det;
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) r.coeffRef(j, i) = internal::cofactor3(m, i, j) * invdet;
    } else if (n == 2) {
        const Scalar invdet = Scalar(1) / (m.coeff(0, 0) * m.coeff(1, 1) - m.coeff(1, 0) * m.coeff(0, 1));
        r.coeffRef(0, 0) = m.coeff(1, 1) * invdet;
        r.coeffRef(1, 0) = -m.coeff(1, 0) * invdet;
        r.coeffRef(0, 1) = -m.coeff(0, 1) * invdet;
        r.coeffRef(1, 1) = m.coeff(0, 0) * invdet;
    } else if (n == 1) {
        r.coeffRef(0, 0) = Scalar(1) / m.coeff(0, 0);
    } else {
        Matrix<Scalar, Dynamic, Dynamic> a(m);
        r.setIdentity();
        for (Index c = 0; c < n; c++) {
            Index p = c;
            for (Index i = c + 1; i < n; i++) if (std::abs(a(i, c)) > std::abs(a(p, c))) p = i;
            if (p != c) for (Index j = 0; j < n; j++) { std::swap(a(p, j), a(c, j)); std::swap(r.coeffRef(p, j), r.coeffRef(c, j)); }
            const Scalar d = a(c, c);
            for (Index j = 0; j < n; j++) { a(c, j) = a(c, j) / d; r.coeffRef(c, j) = r.coeff(c, j) / d; }
            for (Index i = 0; i < n; i++) if (i != c) {
                const Scalar f = a(i, c);
                for (Index j = 0; j < n; j++) { a(i, j) = a(i, j) - f * a(c, j); r.coeffRef(i, j) = r.coeff(i, j) - f * r.coeff(c, j); }
            }
        }
    }
    return r;
}
template <class D> inline typename DenseBase<D>::Scalar DenseBase<D>::determinant() const {
    const D& m = derived();
    const Index n = m.rows();
    if (n == 1) return m.coeff(0, 0);
    if (n == 2) return m.coeff(0, 0) * m.coeff(1, 1) - m.coeff(1, 0) * m.coeff(0, 1);
    if (n == 3) return m.coeff(0, 0) * internal::cofactor3(m, 0, 0) + m.coeff(1, 0) * internal::cofactor3(m, 1, 0) + m.coeff(2, 0) * internal::cofactor3(m, 2, 0);
    Matrix<Scalar, Dynamic, Dynamic> a(m);
    Scalar det = 1;
    for (Index c = 0; c < n; c++) {
        Index p = c;
        for (Index i = c + 1; i < n; i++) if (std::abs(a(i, c)) > std::abs(a(p, c))) p = i;
        if (a(p, c) == Scalar(0)) return Scalar(0);
        if (p != c) { for (Index j = 0; j < n; j++) std::swap(a(p, j), a(c, j)); det = -det; }
        det = det * a(c, c);
        for (Index i = c + 1; i < n; i++) { const Scalar f = a(i, c) / a(c, c); for (Index j = c; j < n; j++) a(i, j) = a(i, j) - f * a(c, j); }
    }
    return det;
}

// =====================================================================================
// Two-sided Jacobi SVD of a square matrix (what JacobiSVD does once the matrix is square): sweep over the
// 2x2 sub-blocks, make each real-symmetric with a rotation from the left, diagonalise it with a Jacobi
// rotation, accumulate the rotations into U and V, stop when every off-diagonal pair is below
// 2*epsilon * (largest diagonal entry seen); then make the singular values positive and sort them descending.
// =====================================================================================
template <class M> class JacobiSVD {
public:
    typedef typename M::Scalar Scalar;
    typedef Matrix<Scalar, Dynamic, Dynamic> Mat;
    typedef Matrix<Scalar, Dynamic, 1> Vec;
    JacobiSVD() {}
    template <class D> explicit JacobiSVD(const DenseBase<D>& a, unsigned int = 0) { compute(a); }
    template <class D> JacobiSVD& compute(const DenseBase<D>& a_, unsigned int = 0) {
        Mat a(a_.derived());
        // a tall matrix is reduced to square by a thin QR (modified Gram-Schmidt suffices for this shim)
        EIGEN_SHIM_CHECK(a.rows() >= a.cols(), "JacobiSVD shim: rows >= cols only");
        const Index n = a.cols();
        Mat Q;
        if (a.rows() != n) {
            Q = Mat::Zero(a.rows(), n);
            Mat Rm = Mat::Zero(n, n);
            Mat w(a);
            for (Index j = 0; j < n; j++) {
                Scalar nr = w.col(j).norm();
                Rm(j, j) = nr;
                for (Index i = 0; i < a.rows(); i++) Q(i, j) = nr > Scalar(0) ? w(i, j) / nr : Scalar(0);
                for (Index k = j + 1; k < n; k++) {
                    Scalar d = 0;
                    for (Index i = 0; i < a.rows(); i++) d += Q(i, j) * w(i, k);
                    Rm(j, k) = d;
                    for (Index i = 0; i < a.rows(); i++) w(i, k) -= d * Q(i, j);
                }
            }
            a = Rm;
        }
        Scalar scale = 0;
        for (Index k = 0; k < a.size(); k++) scale = std::max(scale, Scalar(std::abs(a(k))));
        if (scale == Scalar(0)) scale = Scalar(1);
        Mat W = a / scale;
        Mat U = Mat::Identity(n, n), V = Mat::Identity(n, n);
        const Scalar precision = Scalar(2) * std::numeric_limits<Scalar>::epsilon();
        const Scalar considerAsZero = (std::numeric_limits<Scalar>::min)();
        Scalar maxDiag = 0;
        for (Index i = 0; i < n; i++) maxDiag = std::max(maxDiag, Scalar(std::abs(W(i, i))));
        bool finished = false;
        int sweeps = 0;
        while (!finished && sweeps++ < 1000) {
            finished = true;
            for (Index p = 1; p < n; p++) {
                for (Index q = 0; q < p; q++) {
                    const Scalar threshold = std::max(considerAsZero, precision * maxDiag);
                    if (std::abs(W(p, q)) > threshold || std::abs(W(q, p)) > threshold) {
                        finished = false;
                        // 2x2 block m = [W(p,p) W(p,q); W(q,p) W(q,q)]
                        Scalar m00 = W(p, p), m01 = W(p, q), m10 = W(q, p), m11 = W(q, q);
                        // rot1 makes it symmetric
                        Scalar c1, s1;
                        const Scalar t = m00 + m11, d = m10 - m01;
                        if (std::abs(d) < considerAsZero) { s1 = 0; c1 = 1; }
                        else { const Scalar u = t / d, tmp = std::sqrt(Scalar(1) + u * u); s1 = Scalar(1) / tmp; c1 = u / tmp; }
                        // m <- rot1 applied on the left: row0' = c*row0 + s*row1, row1' = -s*row0 + c*row1
                        const Scalar n00 = c1 * m00 + s1 * m10, n01 = c1 * m01 + s1 * m11, n11 = -s1 * m01 + c1 * m11;
                        // Jacobi rotation diagonalising the symmetric [n00 n01; n01 n11]
                        Scalar cr, sr;
                        const Scalar deno = Scalar(2) * std::abs(n01);
                        if (deno < considerAsZero) { cr = 1; sr = 0; }
                        else {
                            const Scalar tau = (n00 - n11) / deno, w = std::sqrt(tau * tau + Scalar(1));
                            const Scalar tt = tau > Scalar(0) ? Scalar(1) / (tau + w) : Scalar(1) / (tau - w);
                            const Scalar sign_t = tt > Scalar(0) ? Scalar(1) : Scalar(-1);
                            const Scalar nn = Scalar(1) / std::sqrt(tt * tt + Scalar(1));
                            sr = -sign_t * (n01 / std::abs(n01)) * std::abs(tt) * nn;
                            cr = nn;
                        }
                        // left rotation = rot1 * right^T  (composition of plane rotations (c,s))
                        const Scalar cl = c1 * cr + s1 * sr, sl = s1 * cr - c1 * sr;
                        // W <- L * W on rows p,q ; U <- U * L^T on columns p,q
                        for (Index j = 0; j < n; j++) {
                            const Scalar xp = W(p, j), xq = W(q, j);
                            W(p, j) = cl * xp + sl * xq;
                            W(q, j) = -sl * xp + cl * xq;
                        }
                        for (Index i = 0; i < n; i++) {
                            const Scalar xp = U(i, p), xq = U(i, q);
                            U(i, p) = cl * xp + sl * xq;
                            U(i, q) = -sl * xp + cl * xq;
                        }
                        // W <- W * R on columns p,q ; V <- V * R
                        for (Index i = 0; i < n; i++) {
                            const Scalar xp = W(i, p), xq = W(i, q);
                            W(i, p) = cr * xp - sr * xq;
                            W(i, q) = sr * xp + cr * xq;
                        }
                        for (Index i = 0; i < n; i++) {
                            const Scalar xp = V(i, p), xq = V(i, q);
                            V(i, p) = cr * xp - sr * xq;
                            V(i, q) = sr * xp + cr * xq;
                        }
                        maxDiag = std::max(maxDiag, std::max(Scalar(std::abs(W(p, p))), Scalar(std::abs(W(q, q)))));
                    }
                }
            }
        }
        m_sv.resize(n);
        for (Index i = 0; i < n; i++) {
            const Scalar a_ii = W(i, i);
            m_sv(i) = std::abs(a_ii);
            if (a_ii < Scalar(0)) for (Index r = 0; r < n; r++) U(r, i) = -U(r, i);
        }
        for (Index i = 0; i < n; i++) m_sv(i) = m_sv(i) * scale;
        for (Index i = 0; i < n; i++) {   // selection sort, descending, swapping the vectors along
            Index pos = i;
            for (Index k = i + 1; k < n; k++) if (m_sv(k) > m_sv(pos)) pos = k;
            if (m_sv(pos) == Scalar(0)) break;
            if (pos != i) {
                std::swap(m_sv(i), m_sv(pos));
                for (Index r = 0; r < n; r++) { std::swap(U(r, i), U(r, pos)); std::swap(V(r, i), V(r, pos)); }
            }
        }
        m_V = V;
        m_U = (Q.size() > 0) ? Mat(Q * U) : U;
        return *this;
    }
    const Mat& matrixU() const { return m_U; }
    const Mat& matrixV() const { return m_V; }
    const Vec& singularValues() const { return m_sv; }

private:
    Mat m_U, m_V;
    Vec m_sv;
};
template <class D> inline JacobiSVD<Matrix<typename DenseBase<D>::Scalar, Dynamic, Dynamic>> DenseBase<D>::jacobiSvd(unsigned int options) const {
    return JacobiSVD<Matrix<Scalar, Dynamic, Dynamic>>(derived(), options);
}

// Euler angles (a0,a1,a2) of a rotation matrix, Eigen 3.3 convention (first angle in [0, pi])
template <class D> inline Matrix<typename DenseBase<D>::Scalar, 3, 1> DenseBase<D>::eulerAngles(Index a0, Index a1, Index a2) const {
    const D& m = derived();
    Matrix<Scalar, 3, 1> res;
    const Index odd = ((a0 + 1) % 3 == a1) ? 0 : 1;
    const Index i = a0, j = (a0 + 1 + odd) % 3, k = (a0 + 2 - odd) % 3;
    const Scalar pi = Scalar(EIGEN_PI);
    if (a0 == a2) {
        res[0] = std::atan2(m.coeff(j, i), m.coeff(k, i));
        if ((odd && res[0] < Scalar(0)) || ((!odd) && res[0] > Scalar(0))) {
            if (res[0] > Scalar(0)) res[0] -= pi; else res[0] += pi;
            const Scalar s2 = std::sqrt(m.coeff(j, i) * m.coeff(j, i) + m.coeff(k, i) * m.coeff(k, i));
            res[1] = -std::atan2(s2, m.coeff(i, i));
        } else {
            const Scalar s2 = std::sqrt(m.coeff(j, i) * m.coeff(j, i) + m.coeff(k, i) * m.coeff(k, i));
            res[1] = std::atan2(s2, m.coeff(i, i));
        }
        const Scalar s1 = std::sin(res[0]), c1 = std::cos(res[0]);
        res[2] = std::atan2(c1 * m.coeff(j, k) - s1 * m.coeff(k, k), c1 * m.coeff(j, j) - s1 * m.coeff(k, j));
    } else {
        res[0] = std::atan2(m.coeff(j, k), m.coeff(k, k));
        const Scalar c2 = std::sqrt(m.coeff(i, i) * m.coeff(i, i) + m.coeff(i, j) * m.coeff(i, j));
        if ((odd && res[0] < Scalar(0)) || ((!odd) && res[0] > Scalar(0))) {
            if (res[0] > Scalar(0)) res[0] -= pi; else res[0] += pi;
            res[1] = std::atan2(-m.coeff(i, k), -c2);
        } else {
            res[1] = std::atan2(-m.coeff(i, k), c2);
        }
        const Scalar s1 = std::sin(res[0]), c1 = std::cos(res[0]);
        res[2] = std::atan2(s1 * m.coeff(k, i) - c1 * m.coeff(j, i), c1 * m.coeff(j, j) - s1 * m.coeff(k, j));
    }
    if (!odd) res = -res;
    return res;
}

// =====================================================================================
// AngleAxis / Quaternion (used only by util.cpp:euler_to_rotation_matrix, off the edge-sketch path)
// =====================================================================================
template <class S> class Quaternion;
template <class S> class AngleAxis {
public:
    typedef Matrix<S, 3, 1> Vector3;
    AngleAxis() : m_angle(0) { m_axis.setZero(); }
    template <class D> AngleAxis(const S& angle, const DenseBase<D>& axis) : m_angle(angle), m_axis(axis) {}
    S angle() const { return m_angle; }
    const Vector3& axis() const { return m_axis; }
    Matrix<S, 3, 3> toRotationMatrix() const { return Quaternion<S>(*this).toRotationMatrix(); }
private:
    S m_angle;
    Vector3 m_axis;
};
template <class S> class Quaternion {
public:
    Quaternion() : m_w(1), m_x(0), m_y(0), m_z(0) {}
    Quaternion(const S& w, const S& x, const S& y, const S& z) : m_w(w), m_x(x), m_y(y), m_z(z) {}
    Quaternion(const AngleAxis<S>& aa) {
        const S ha = S(0.5) * aa.angle();
        const S s = std::sin(ha);
        m_w = std::cos(ha); m_x = s * aa.axis().x(); m_y = s * aa.axis().y(); m_z = s * aa.axis().z();
    }
    S w() const { return m_w; }
    S x() const { return m_x; }
    S y() const { return m_y; }
    S z() const { return m_z; }
    Quaternion operator*(const Quaternion& b) const {
        const Quaternion& a = *this;
        return Quaternion(a.m_w * b.m_w - a.m_x * b.m_x - a.m_y * b.m_y - a.m_z * b.m_z,
                          a.m_w * b.m_x + a.m_x * b.m_w + a.m_y * b.m_z - a.m_z * b.m_y,
                          a.m_w * b.m_y + a.m_y * b.m_w + a.m_z * b.m_x - a.m_x * b.m_z,
                          a.m_w * b.m_z + a.m_z * b.m_w + a.m_x * b.m_y - a.m_y * b.m_x);
    }
    Matrix<S, 3, 3> toRotationMatrix() const {
        Matrix<S, 3, 3> r;
        const S tx = S(2) * m_x, ty = S(2) * m_y, tz = S(2) * m_z;
        const S twx = tx * m_w, twy = ty * m_w, twz = tz * m_w, txx = tx * m_x, txy = ty * m_x, txz = tz * m_x;
        const S tyy = ty * m_y, tyz = tz * m_y, tzz = tz * m_z;
        r(0, 0) = S(1) - (tyy + tzz); r(0, 1) = txy - twz; r(0, 2) = txz + twy;
        r(1, 0) = txy + twz; r(1, 1) = S(1) - (txx + tzz); r(1, 2) = tyz - twx;
        r(2, 0) = txz - twy; r(2, 1) = tyz + twx; r(2, 2) = S(1) - (txx + tyy);
        return r;
    }
private:
    S m_w, m_x, m_y, m_z;
};
template <class S> inline Quaternion<S> operator*(const AngleAxis<S>& a, const AngleAxis<S>& b) { return Quaternion<S>(a) * Quaternion<S>(b); }
template <class S> inline Quaternion<S> operator*(const Quaternion<S>& a, const AngleAxis<S>& b) { return a * Quaternion<S>(b); }
template <class S> inline Quaternion<S> operator*(const AngleAxis<S>& a, const Quaternion<S>& b) { return Quaternion<S>(a) * b; }
typedef AngleAxis<double> AngleAxisd;
typedef AngleAxis<float> AngleAxisf;
typedef Quaternion<double> Quaterniond;
typedef Quaternion<float> Quaternionf;

// =====================================================================================
// Stream output in Eigen's default IOFormat: the stream's own precision, one space between coefficients,
// "\n" between rows, every coefficient right-aligned to the widest one of the whole matrix, no trailing newline.
// =====================================================================================
template <class D> inline std::ostream& operator<<(std::ostream& s, const DenseBase<D>& m_) {
    const D& m = m_.derived();
    if (m.size() == 0) return s;
    Index width = 0;
    for (Index j = 0; j < m.cols(); j++)
        for (Index i = 0; i < m.rows(); i++) {
            std::stringstream sstr;
            sstr.copyfmt(s);
            sstr << m.coeff(i, j);
            width = std::max<Index>(width, Index(sstr.str().length()));
        }
    for (Index i = 0; i < m.rows(); i++) {
        if (i) s << "\n";
        if (width) s.width(width);
        s << m.coeff(i, 0);
        for (Index j = 1; j < m.cols(); j++) {
            s << " ";
            if (width) s.width(width);
            s << m.coeff(i, j);
        }
    }
    return s;
}

}  // namespace Eigen
