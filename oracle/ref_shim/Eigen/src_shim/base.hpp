// =====================================================================================
// Eigen stand-in, part 1: storage, traits, the CRTP base every dense object derives from.
// TEST INFRASTRUCTURE ONLY -- see oracle/ref_shim/README.md.  This is NOT Eigen and holds no Eigen
// source: it restates, for the small API subset the 3D Curvix reference uses, the semantics Eigen 3.3
// documents (column-major dense storage, element-wise arithmetic in source order, Matrix3d::inverse()
// as cofactors * (1/det), a two-sided Jacobi SVD).  Everything evaluates eagerly into a plain object.
// =====================================================================================
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <limits>
#include <sstream>
#include <type_traits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>   // Eigen/Core pulls omp.h in when OpenMP is on; EdgeSketch_Core.cpp relies on that
#endif

// 0: every reduction (matrix product inner sums, norm, dot, sum) adds left to right.
// 1: sensitivity probe -- compile-time-sized reductions use the balanced tree a0 + (a1 + a2) that an
//    unrolled scalar reduction produces; run-time-sized ones stay left to right.
#ifndef EIGEN_SHIM_SUM_ORDER
#define EIGEN_SHIM_SUM_ORDER 0
#endif

#ifdef EIGEN_SHIM_BOUNDS_CHECK
#define EIGEN_SHIM_CHECK(cond, what)                                                              \
    do {                                                                                          \
        if (!(cond)) {                                                                            \
            std::fprintf(stderr, "[eigen-shim] %s failed: %s (%s:%d)\n", what, #cond, __FILE__, __LINE__); \
            std::abort();                                                                         \
        }                                                                                         \
    } while (0)
#else
#define EIGEN_SHIM_CHECK(cond, what) ((void)0)
#endif

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW
#define EIGEN_PI 3.141592653589793238462643383279502884197169399375105820974944592307816406

namespace Eigen {

typedef std::ptrdiff_t Index;
enum { Dynamic = -1 };
enum { ColMajor = 0, RowMajor = 1, AutoAlign = 0, DontAlign = 2 };
enum { Unaligned = 0, Aligned = 16 };
enum { ComputeFullU = 0x04, ComputeThinU = 0x08, ComputeFullV = 0x10, ComputeThinV = 0x20 };

template <class S, int R, int C, int Opt = 0, int MR = R, int MC = C> class Matrix;
template <class S, int R, int C, int Opt = 0, int MR = R, int MC = C> class Array;
template <class S, int R, int C> class Block;
template <class P, int MapOpt = 0> class Map;
template <class M> class JacobiSVD;
template <class D> class DenseBase;
template <class D> class CommaInitializer;

namespace internal {

// value put into every cell that real Eigen would leave uninitialised (default construction of a
// fixed-size object, resize, the new cells of conservativeResize): reads of it surface as NaN.
template <class S> inline S sentinel() {
    if constexpr (std::is_floating_point<S>::value) return std::numeric_limits<S>::quiet_NaN();
    else return std::numeric_limits<S>::min();
}
constexpr int pick(int a, int b) { return a != Dynamic ? a : b; }
constexpr int mul_dim(int a, int b) { return (a == Dynamic || b == Dynamic) ? Dynamic : a * b; }

template <class D> struct traits;
template <class S, int R, int C, int O, int MR, int MC> struct traits<Matrix<S, R, C, O, MR, MC>> {
    typedef S Scalar;
    enum { Rows = R, Cols = C, IsArray = 0 };
};
template <class S, int R, int C, int O, int MR, int MC> struct traits<Array<S, R, C, O, MR, MC>> {
    typedef S Scalar;
    enum { Rows = R, Cols = C, IsArray = 1 };
};
template <class S, int R, int C> struct traits<Block<S, R, C>> {
    typedef S Scalar;
    enum { Rows = R, Cols = C, IsArray = 0 };
};
template <class P, int O> struct traits<Map<P, O>> {
    typedef typename traits<P>::Scalar Scalar;
    enum { Rows = traits<P>::Rows, Cols = traits<P>::Cols, IsArray = traits<P>::IsArray };
};

template <class S, int R, int C, bool IsArr> struct plain { typedef Matrix<S, R, C> type; };
template <class S, int R, int C> struct plain<S, R, C, true> { typedef Array<S, R, C> type; };

// ---- storage: inline for fixed sizes, heap for anything with a run-time dimension
template <class S, int R, int C, bool Dyn = (R == Dynamic || C == Dynamic)> struct Storage {
    S d[R * C > 0 ? R * C : 1];
    Storage() { for (int i = 0; i < R * C; i++) d[i] = sentinel<S>(); }
    Index rows() const { return R; }
    Index cols() const { return C; }
    S* data() { return d; }
    const S* data() const { return d; }
    void resize(Index r, Index c) { EIGEN_SHIM_CHECK(r == R && c == C, "fixed-size resize"); (void)r; (void)c; }
    void conservativeResize(Index r, Index c) { resize(r, c); }
};
template <class S, int R, int C> struct Storage<S, R, C, true> {
    std::vector<S> v;
    Index r, c;
    Storage() : r(R == Dynamic ? 0 : R), c(C == Dynamic ? 0 : C) {}
    Storage(const Storage&) = default;
    Storage& operator=(const Storage&) = default;
    Storage(Storage&& o) noexcept : v(std::move(o.v)), r(o.r), c(o.c) { o.reset(); }
    Storage& operator=(Storage&& o) noexcept { v = std::move(o.v); r = o.r; c = o.c; o.reset(); return *this; }
    void reset() { v.clear(); r = (R == Dynamic ? 0 : R); c = (C == Dynamic ? 0 : C); if (R != Dynamic && C != Dynamic) v.assign(size_t(r * c), sentinel<S>()); else if (R != Dynamic) c = 0; else if (C != Dynamic) r = 0; }
    Index rows() const { return r; }
    Index cols() const { return c; }
    S* data() { return v.data(); }
    const S* data() const { return v.data(); }
    // Eigen's resize keeps the buffer when the size is unchanged and otherwise reallocates (contents lost)
    void resize(Index nr, Index nc) {
        EIGEN_SHIM_CHECK((R == Dynamic || nr == R) && (C == Dynamic || nc == C) && nr >= 0 && nc >= 0, "resize");
        if (nr * nc != r * c) v.assign(size_t(nr * nc), sentinel<S>());
        r = nr; c = nc;
    }
    // keeps the overlapping top-left block, new cells are uninitialised in Eigen (sentinel here)
    void conservativeResize(Index nr, Index nc) {
        EIGEN_SHIM_CHECK((R == Dynamic || nr == R) && (C == Dynamic || nc == C) && nr >= 0 && nc >= 0, "conservativeResize");
        if (nr == r && nc == c) return;
        if (nr == r) { v.resize(size_t(nr * nc), sentinel<S>()); c = nc; return; }   // column-major: columns append
        std::vector<S> nv(size_t(nr * nc), sentinel<S>());
        const Index mr = std::min(nr, r), mc = std::min(nc, c);
        for (Index j = 0; j < mc; j++)
            for (Index i = 0; i < mr; i++) nv[size_t(i + j * nr)] = v[size_t(i + j * r)];
        v.swap(nv);
        r = nr; c = nc;
    }
};

// ---- reductions in a defined order
template <int Start, int Len, class F> inline auto tree_sum(const F& term) {
    if constexpr (Len == 1) return term(Index(Start));
    else return tree_sum<Start, Len / 2>(term) + tree_sum<Start + Len / 2, Len - Len / 2>(term);
}
// sum of term(0..n-1); K = the compile-time length or Dynamic
template <int K, class S, class F> inline S ordered_sum(Index n, const F& term) {
#if EIGEN_SHIM_SUM_ORDER == 1
    if constexpr (K != Dynamic && K > 0) { (void)n; return tree_sum<0, K>(term); }
#endif
    if (n <= 0) return S(0);
    S acc = term(Index(0));
    for (Index k = 1; k < n; k++) acc = acc + term(k);
    return acc;
}

}  // namespace internal

// =====================================================================================
// DenseBase: everything that can be read.  A derived class D provides rows(), cols(), data(),
// outerStride() (column-major: element (i,j) lives at data()[i + j*outerStride()]).
// =====================================================================================
template <class D> class DenseBase {
public:
    typedef typename internal::traits<D>::Scalar Scalar;
    enum {
        RowsAtCompileTime = internal::traits<D>::Rows,
        ColsAtCompileTime = internal::traits<D>::Cols,
        SizeAtCompileTime = internal::mul_dim(internal::traits<D>::Rows, internal::traits<D>::Cols),
        IsArrayKind = internal::traits<D>::IsArray,
        IsVectorAtCompileTime = (internal::traits<D>::Rows == 1 || internal::traits<D>::Cols == 1)
    };
    typedef typename internal::plain<Scalar, RowsAtCompileTime, ColsAtCompileTime, (IsArrayKind != 0)>::type PlainObject;

    const D& derived() const { return *static_cast<const D*>(this); }
    D& derived() { return *static_cast<D*>(this); }

    Index size() const { return derived().rows() * derived().cols(); }
    Index innerSize() const { return derived().rows(); }
    Index outerSize() const { return derived().cols(); }

    Scalar coeff(Index i, Index j) const {
        EIGEN_SHIM_CHECK(i >= 0 && i < derived().rows() && j >= 0 && j < derived().cols(), "coeff(i,j)");
        return derived().data()[i + j * derived().outerStride()];
    }
    Scalar coeff(Index k) const {
        EIGEN_SHIM_CHECK(k >= 0 && k < size(), "coeff(k)");
        const Index r = derived().rows();
        return r == 1 ? coeff(0, k) : coeff(k % r, k / r);
    }
    Scalar& coeffRef(Index i, Index j) const {   // views do not carry const-ness (shim simplification)
        EIGEN_SHIM_CHECK(i >= 0 && i < derived().rows() && j >= 0 && j < derived().cols(), "coeffRef(i,j)");
        return const_cast<Scalar*>(derived().data())[i + j * derived().outerStride()];
    }
    Scalar& coeffRef(Index k) const {
        EIGEN_SHIM_CHECK(k >= 0 && k < size(), "coeffRef(k)");
        const Index r = derived().rows();
        return r == 1 ? coeffRef(0, k) : coeffRef(k % r, k / r);
    }
    Scalar operator()(Index i, Index j) const { return coeff(i, j); }
    Scalar operator()(Index k) const { return coeff(k); }
    Scalar operator[](Index k) const { return coeff(k); }
    Scalar x() const { return coeff(0); }
    Scalar y() const { return coeff(1); }
    Scalar z() const { return coeff(2); }
    Scalar w() const { return coeff(3); }
    // double(expr) on a 1x1 product
    explicit operator Scalar() const { EIGEN_SHIM_CHECK(size() == 1, "scalar conversion of a non-1x1"); return coeff(0, 0); }

    PlainObject eval() const { return PlainObject(derived()); }

    // ---- views.  A Block points into this object's memory; it must not outlive it.
    typedef Block<Scalar, 1, ColsAtCompileTime> RowXpr;
    typedef Block<Scalar, RowsAtCompileTime, 1> ColXpr;
    typedef Block<Scalar, Dynamic, Dynamic> BlockXpr;
    RowXpr row(Index i) const { return RowXpr(&coeffRefUnchecked(i, 0), 1, derived().cols(), derived().outerStride(), i, derived().rows()); }
    ColXpr col(Index j) const { return ColXpr(&coeffRefUnchecked(0, j), derived().rows(), 1, derived().outerStride(), j, derived().cols()); }
    BlockXpr block(Index i, Index j, Index nr, Index nc) const {
        EIGEN_SHIM_CHECK(i >= 0 && j >= 0 && nr >= 0 && nc >= 0 && i + nr <= derived().rows() && j + nc <= derived().cols(), "block");
        return BlockXpr(&coeffRefUnchecked(i, j), nr, nc, derived().outerStride());
    }
    template <int NR, int NC> Block<Scalar, NR, NC> block(Index i, Index j) const {
        EIGEN_SHIM_CHECK(i >= 0 && j >= 0 && i + NR <= derived().rows() && j + NC <= derived().cols(), "block<,>");
        return Block<Scalar, NR, NC>(&coeffRefUnchecked(i, j), NR, NC, derived().outerStride());
    }
    BlockXpr topRows(Index n) const { return block(0, 0, n, derived().cols()); }
    BlockXpr bottomRows(Index n) const { return block(derived().rows() - n, 0, n, derived().cols()); }
    BlockXpr leftCols(Index n) const { return block(0, 0, derived().rows(), n); }
    BlockXpr rightCols(Index n) const { return block(0, derived().cols() - n, derived().rows(), n); }
    // vector segments (row or column vectors)
    auto segment(Index start, Index n) const {
        if constexpr (ColsAtCompileTime == 1) return Block<Scalar, Dynamic, 1>(&coeffRefUnchecked(start, 0), n, 1, derived().outerStride(), start + n, derived().rows() + 1);
        else return Block<Scalar, 1, Dynamic>(&coeffRefUnchecked(0, start), 1, n, derived().outerStride(), start + n, derived().cols() + 1);
    }
    template <int N> auto segment(Index start) const {
        if constexpr (ColsAtCompileTime == 1) return Block<Scalar, N, 1>(&coeffRefUnchecked(start, 0), N, 1, derived().outerStride(), start + N, derived().rows() + 1);
        else return Block<Scalar, 1, N>(&coeffRefUnchecked(0, start), 1, N, derived().outerStride(), start + N, derived().cols() + 1);
    }
    auto head(Index n) const { return segment(0, n); }
    auto tail(Index n) const { return segment(size() - n, n); }
    template <int N> auto head() const { return this->template segment<N>(0); }
    template <int N> auto tail() const { return this->template segment<N>(size() - N); }

    // ---- defined after the plain types (ops.hpp / linalg.hpp)
    auto transpose() const;
    auto array() const;
    auto matrix() const;
    Scalar sum() const;
    Scalar mean() const { return sum() / Scalar(size()); }
    Scalar squaredNorm() const;
    Scalar norm() const { return std::sqrt(squaredNorm()); }
    PlainObject normalized() const;
    template <class O> Scalar dot(const DenseBase<O>& o) const;
    template <class O> auto cross(const DenseBase<O>& o) const;
    Scalar minCoeff() const;
    Scalar maxCoeff() const;
    PlainObject cwiseAbs() const;
    template <class O> PlainObject cwiseProduct(const DenseBase<O>& o) const;
    bool hasNaN() const;
    bool allFinite() const;
    template <class O> bool isApprox(const DenseBase<O>& o, double prec = 1e-12) const;
    PlainObject inverse() const;
    Scalar determinant() const;
    Scalar trace() const;
    JacobiSVD<Matrix<Scalar, Dynamic, Dynamic>> jacobiSvd(unsigned int options = 0) const;
    Matrix<Scalar, 3, 1> eulerAngles(Index a0, Index a1, Index a2) const;
    // element-wise functions (Eigen offers them on arrays; harmless on matrices here)
    PlainObject sqrt() const;
    PlainObject abs() const;
    PlainObject atan() const;
    PlainObject asin() const;
    PlainObject acos() const;
    PlainObject cos() const;
    PlainObject sin() const;
    PlainObject tan() const;
    PlainObject exp() const;
    PlainObject square() const;
    // (cond_array).select(a, b)
    template <class A, class B> auto select(const DenseBase<A>& a, const DenseBase<B>& b) const;

private:
    Scalar& coeffRefUnchecked(Index i, Index j) const {
        return const_cast<Scalar*>(derived().data())[i + j * derived().outerStride()];
    }
};

}  // namespace Eigen
