// Eigen stand-in, part 3: arithmetic.  Every operator evaluates eagerly, element by element, in the order the
// source expression states (Eigen's expression templates evaluate the same scalar expression per coefficient);
// reductions follow internal::ordered_sum.  TEST INFRASTRUCTURE ONLY -- see oracle/ref_shim/README.md.
#pragma once
#include "plain.hpp"

namespace Eigen {

namespace internal {
template <class A, class B> struct cwise_result {
    typedef typename traits<A>::Scalar Scalar;
    typedef typename plain<Scalar, pick(traits<A>::Rows, traits<B>::Rows), pick(traits<A>::Cols, traits<B>::Cols),
                           (traits<A>::IsArray != 0)>::type type;
};
template <class A, class B> struct product_result {
    typedef typename traits<A>::Scalar Scalar;
    typedef Matrix<Scalar, traits<A>::Rows, traits<B>::Cols> type;
};
template <class A, class B, bool BothArrays = (traits<A>::IsArray && traits<B>::IsArray)> struct mul_result { typedef typename product_result<A, B>::type type; };
template <class A, class B> struct mul_result<A, B, true> { typedef typename cwise_result<A, B>::type type; };

template <class R, class A, class B, class F> inline R cwise_binary(const DenseBase<A>& a_, const DenseBase<B>& b_, const F& f) {
    const A& a = a_.derived();
    const B& b = b_.derived();
    R r;
    if (a.rows() == b.rows() && a.cols() == b.cols()) {
        r.resize(a.rows(), a.cols());
        for (Index j = 0; j < a.cols(); j++)
            for (Index i = 0; i < a.rows(); i++) r.coeffRef(i, j) = f(a.coeff(i, j), b.coeff(i, j));
    } else {   // vectors of equal length in different orientation never occur in valid Eigen code; sizes must agree
        EIGEN_SHIM_CHECK(false, "element-wise operation on mismatched shapes");
    }
    return r;
}
template <class A, class F> inline typename DenseBase<A>::PlainObject cwise_unary(const DenseBase<A>& a_, const F& f) {
    const A& a = a_.derived();
    typename DenseBase<A>::PlainObject r;
    r.resize(a.rows(), a.cols());
    for (Index j = 0; j < a.cols(); j++)
        for (Index i = 0; i < a.rows(); i++) r.coeffRef(i, j) = f(a.coeff(i, j));
    return r;
}
}  // namespace internal

// ---- sums and differences
template <class A, class B> inline typename internal::cwise_result<A, B>::type operator+(const DenseBase<A>& a, const DenseBase<B>& b) {
    typedef typename internal::traits<A>::Scalar S;
    return internal::cwise_binary<typename internal::cwise_result<A, B>::type>(a, b, [](S x, S y) { return x + y; });
}
template <class A, class B> inline typename internal::cwise_result<A, B>::type operator-(const DenseBase<A>& a, const DenseBase<B>& b) {
    typedef typename internal::traits<A>::Scalar S;
    return internal::cwise_binary<typename internal::cwise_result<A, B>::type>(a, b, [](S x, S y) { return x - y; });
}
template <class A> inline typename DenseBase<A>::PlainObject operator-(const DenseBase<A>& a) {
    typedef typename internal::traits<A>::Scalar S;
    return internal::cwise_unary(a, [](S x) { return -x; });
}

// ---- scalar on either side (the scalar is first converted to the object's scalar type, as Eigen does)
#define EIGEN_SHIM_ARITH(T) typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0
template <class A, class T, EIGEN_SHIM_ARITH(T)> inline typename DenseBase<A>::PlainObject operator*(const DenseBase<A>& a, const T& s) {
    typedef typename internal::traits<A>::Scalar S;
    const S ss = S(s);
    return internal::cwise_unary(a, [ss](S x) { return x * ss; });
}
template <class A, class T, EIGEN_SHIM_ARITH(T)> inline typename DenseBase<A>::PlainObject operator*(const T& s, const DenseBase<A>& a) {
    typedef typename internal::traits<A>::Scalar S;
    const S ss = S(s);
    return internal::cwise_unary(a, [ss](S x) { return ss * x; });
}
template <class A, class T, EIGEN_SHIM_ARITH(T)> inline typename DenseBase<A>::PlainObject operator/(const DenseBase<A>& a, const T& s) {
    typedef typename internal::traits<A>::Scalar S;
    const S ss = S(s);
    return internal::cwise_unary(a, [ss](S x) { return x / ss; });
}
// array-only in Eigen: scalar / array, array +- scalar
template <class A, class T, EIGEN_SHIM_ARITH(T)> inline typename DenseBase<A>::PlainObject operator/(const T& s, const DenseBase<A>& a) {
    typedef typename internal::traits<A>::Scalar S;
    const S ss = S(s);
    return internal::cwise_unary(a, [ss](S x) { return ss / x; });
}
template <class A, class T, EIGEN_SHIM_ARITH(T)> inline typename DenseBase<A>::PlainObject operator+(const DenseBase<A>& a, const T& s) {
    typedef typename internal::traits<A>::Scalar S;
    const S ss = S(s);
    return internal::cwise_unary(a, [ss](S x) { return x + ss; });
}
template <class A, class T, EIGEN_SHIM_ARITH(T)> inline typename DenseBase<A>::PlainObject operator-(const DenseBase<A>& a, const T& s) {
    typedef typename internal::traits<A>::Scalar S;
    const S ss = S(s);
    return internal::cwise_unary(a, [ss](S x) { return x - ss; });
}

// ---- products: matrix product for matrices, element-wise for arrays
template <class A, class B> inline typename internal::mul_result<A, B>::type operator*(const DenseBase<A>& a_, const DenseBase<B>& b_) {
    typedef typename internal::traits<A>::Scalar S;
    static_assert(int(internal::traits<A>::IsArray) == int(internal::traits<B>::IsArray), "array * matrix is not allowed");
    if constexpr (internal::traits<A>::IsArray) {
        return internal::cwise_binary<typename internal::mul_result<A, B>::type>(a_, b_, [](S x, S y) { return x * y; });
    } else {
        const A& a = a_.derived();
        const B& b = b_.derived();
        EIGEN_SHIM_CHECK(a.cols() == b.rows(), "matrix product inner dimension");
        typename internal::mul_result<A, B>::type r;
        r.resize(a.rows(), b.cols());
        constexpr int K = internal::pick(internal::traits<A>::Cols, internal::traits<B>::Rows);
        const Index n = a.cols();
        for (Index j = 0; j < b.cols(); j++)
            for (Index i = 0; i < a.rows(); i++)
                r.coeffRef(i, j) = internal::ordered_sum<K, S>(n, [&](Index k) { return a.coeff(i, k) * b.coeff(k, j); });
        return r;
    }
}
template <class A, class B> inline typename internal::cwise_result<A, B>::type operator/(const DenseBase<A>& a, const DenseBase<B>& b) {
    typedef typename internal::traits<A>::Scalar S;
    return internal::cwise_binary<typename internal::cwise_result<A, B>::type>(a, b, [](S x, S y) { return x / y; });
}

// ---- comparisons (arrays): the result is an int array usable with select()
#define EIGEN_SHIM_CMP(OP)                                                                                                         \
    template <class A, class T, EIGEN_SHIM_ARITH(T)>                                                                               \
    inline Array<int, internal::traits<A>::Rows, internal::traits<A>::Cols> operator OP(const DenseBase<A>& a_, const T& s) {      \
        const A& a = a_.derived();                                                                                                 \
        typedef typename internal::traits<A>::Scalar S;                                                                            \
        const S ss = S(s);                                                                                                         \
        Array<int, internal::traits<A>::Rows, internal::traits<A>::Cols> r;                                                        \
        r.resize(a.rows(), a.cols());                                                                                              \
        for (Index j = 0; j < a.cols(); j++)                                                                                       \
            for (Index i = 0; i < a.rows(); i++) r.coeffRef(i, j) = (a.coeff(i, j) OP ss) ? 1 : 0;                                 \
        return r;                                                                                                                  \
    }
EIGEN_SHIM_CMP(<)
EIGEN_SHIM_CMP(<=)
EIGEN_SHIM_CMP(>)
EIGEN_SHIM_CMP(>=)
#undef EIGEN_SHIM_CMP

// matrix == matrix is a bool in Eigen (all coefficients equal)
template <class A, class B> inline bool operator==(const DenseBase<A>& a_, const DenseBase<B>& b_) {
    const A& a = a_.derived();
    const B& b = b_.derived();
    if (a.rows() != b.rows() || a.cols() != b.cols()) return false;
    for (Index j = 0; j < a.cols(); j++)
        for (Index i = 0; i < a.rows(); i++)
            if (!(a.coeff(i, j) == b.coeff(i, j))) return false;
    return true;
}
template <class A, class B> inline bool operator!=(const DenseBase<A>& a, const DenseBase<B>& b) { return !(a == b); }

// =====================================================================================
// DenseBase members that need the plain types
// =====================================================================================
template <class D> inline auto DenseBase<D>::transpose() const {
    typename internal::plain<Scalar, ColsAtCompileTime, RowsAtCompileTime, (IsArrayKind != 0)>::type r;
    const D& a = derived();
    r.resize(a.cols(), a.rows());
    for (Index j = 0; j < a.cols(); j++)
        for (Index i = 0; i < a.rows(); i++) r.coeffRef(j, i) = a.coeff(i, j);
    return r;
}
template <class D> inline auto DenseBase<D>::array() const { return Array<Scalar, RowsAtCompileTime, ColsAtCompileTime>(derived()); }
template <class D> inline auto DenseBase<D>::matrix() const { return Matrix<Scalar, RowsAtCompileTime, ColsAtCompileTime>(derived()); }

template <class D> inline typename DenseBase<D>::Scalar DenseBase<D>::sum() const {
    const D& a = derived();
    return internal::ordered_sum<SizeAtCompileTime, Scalar>(a.rows() * a.cols(), [&](Index k) { return a.coeff(k); });
}
template <class D> inline typename DenseBase<D>::Scalar DenseBase<D>::squaredNorm() const {
    const D& a = derived();
    return internal::ordered_sum<SizeAtCompileTime, Scalar>(a.rows() * a.cols(), [&](Index k) { const Scalar v = a.coeff(k); return v * v; });
}
template <class D> inline typename DenseBase<D>::PlainObject DenseBase<D>::normalized() const {
    const Scalar n = norm();   // Eigen 3.3: derived() / norm()
    return internal::cwise_unary(*this, [n](Scalar x) { return x / n; });
}
template <class D> template <class O> inline typename DenseBase<D>::Scalar DenseBase<D>::dot(const DenseBase<O>& o_) const {
    const D& a = derived();
    const O& b = o_.derived();
    EIGEN_SHIM_CHECK(a.size() == b.size(), "dot of different lengths");
    constexpr int K = internal::pick(SizeAtCompileTime, DenseBase<O>::SizeAtCompileTime);
    return internal::ordered_sum<K, Scalar>(a.size(), [&](Index k) { return a.coeff(k) * b.coeff(k); });
}
template <class D> template <class O> inline auto DenseBase<D>::cross(const DenseBase<O>& o_) const {
    const D& a = derived();
    const O& b = o_.derived();
    EIGEN_SHIM_CHECK(a.size() == 3 && b.size() == 3, "cross needs 3-vectors");
    Matrix<Scalar, 3, 1> r;
    r.coeffRef(0) = a.coeff(1) * b.coeff(2) - a.coeff(2) * b.coeff(1);
    r.coeffRef(1) = a.coeff(2) * b.coeff(0) - a.coeff(0) * b.coeff(2);
    r.coeffRef(2) = a.coeff(0) * b.coeff(1) - a.coeff(1) * b.coeff(0);
    return r;
}
template <class D> inline typename DenseBase<D>::Scalar DenseBase<D>::minCoeff() const {
    const D& a = derived();
    EIGEN_SHIM_CHECK(a.size() > 0, "minCoeff of an empty object");
    Scalar m = a.coeff(0);
    for (Index k = 1; k < a.size(); k++) { const Scalar v = a.coeff(k); m = (v < m) ? v : m; }   // numext::mini
    return m;
}
template <class D> inline typename DenseBase<D>::Scalar DenseBase<D>::maxCoeff() const {
    const D& a = derived();
    EIGEN_SHIM_CHECK(a.size() > 0, "maxCoeff of an empty object");
    Scalar m = a.coeff(0);
    for (Index k = 1; k < a.size(); k++) { const Scalar v = a.coeff(k); m = (m < v) ? v : m; }   // numext::maxi
    return m;
}
template <class D> inline typename DenseBase<D>::PlainObject DenseBase<D>::cwiseAbs() const { return internal::cwise_unary(*this, [](Scalar x) { return Scalar(std::abs(x)); }); }
template <class D> template <class O> inline typename DenseBase<D>::PlainObject DenseBase<D>::cwiseProduct(const DenseBase<O>& o) const {
    return internal::cwise_binary<PlainObject>(*this, o, [](Scalar x, Scalar y) { return x * y; });
}
template <class D> inline bool DenseBase<D>::hasNaN() const {
    const D& a = derived();
    for (Index k = 0; k < a.size(); k++) { const Scalar v = a.coeff(k); if (!(v == v)) return true; }
    return false;
}
template <class D> inline bool DenseBase<D>::allFinite() const {
    const D& a = derived();
    for (Index k = 0; k < a.size(); k++) { const Scalar v = a.coeff(k) - a.coeff(k); if (!(v == v)) return false; }
    return true;
}
// Eigen: ||a - b||^2 <= prec^2 * min(||a||^2, ||b||^2)
template <class D> template <class O> inline bool DenseBase<D>::isApprox(const DenseBase<O>& o, double prec) const {
    const Scalar d2 = (derived() - o.derived()).squaredNorm();
    return d2 <= Scalar(prec) * Scalar(prec) * std::min(squaredNorm(), o.squaredNorm());
}
template <class D> inline typename DenseBase<D>::Scalar DenseBase<D>::trace() const {
    const D& a = derived();
    return internal::ordered_sum<Dynamic, Scalar>(std::min(a.rows(), a.cols()), [&](Index k) { return a.coeff(k, k); });
}

#define EIGEN_SHIM_UNARY(NAME, EXPR) \
    template <class D> inline typename DenseBase<D>::PlainObject DenseBase<D>::NAME() const { return internal::cwise_unary(*this, [](Scalar x) { return Scalar(EXPR); }); }
EIGEN_SHIM_UNARY(sqrt, std::sqrt(x))
EIGEN_SHIM_UNARY(abs, std::abs(x))
EIGEN_SHIM_UNARY(atan, std::atan(x))
EIGEN_SHIM_UNARY(asin, std::asin(x))
EIGEN_SHIM_UNARY(acos, std::acos(x))
EIGEN_SHIM_UNARY(cos, std::cos(x))
EIGEN_SHIM_UNARY(sin, std::sin(x))
EIGEN_SHIM_UNARY(tan, std::tan(x))
EIGEN_SHIM_UNARY(exp, std::exp(x))
EIGEN_SHIM_UNARY(square, x * x)
#undef EIGEN_SHIM_UNARY

template <class D> template <class A, class B> inline auto DenseBase<D>::select(const DenseBase<A>& a_, const DenseBase<B>& b_) const {
    const D& c = derived();
    const A& a = a_.derived();
    const B& b = b_.derived();
    typename DenseBase<A>::PlainObject r;
    r.resize(c.rows(), c.cols());
    for (Index j = 0; j < c.cols(); j++)
        for (Index i = 0; i < c.rows(); i++) r.coeffRef(i, j) = c.coeff(i, j) ? a.coeff(i, j) : b.coeff(i, j);
    return r;
}

}  // namespace Eigen
