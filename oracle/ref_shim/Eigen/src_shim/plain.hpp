// Eigen stand-in, part 2: writable objects -- Matrix, Array, Block (view), Map (view), comma initialiser.
// TEST INFRASTRUCTURE ONLY -- see oracle/ref_shim/README.md.
#pragma once
#include "base.hpp"

namespace Eigen {

namespace internal {
// dst = src for any two dense objects; vectors may be assigned across orientation (Eigen allows
// row-vector = column-vector).  The right-hand side is always an evaluated object, so there is no aliasing.
template <class Dst, class Src> inline void assign(Dst& dst, const DenseBase<Src>& src_) {
    const Src& src = src_.derived();
    const Index sr = src.rows(), sc = src.cols();
    dst.resizeLike(sr, sc);
    typedef typename traits<Dst>::Scalar S;
    if (dst.rows() == sr && dst.cols() == sc) {
        for (Index j = 0; j < sc; j++)
            for (Index i = 0; i < sr; i++) dst.coeffRef(i, j) = S(src.coeff(i, j));
    } else {
        // Vector across orientation.  A fixed-size vector assigned from a LONGER run-time-sized vector takes its leading
        // coefficients: that is what an NDEBUG build of Eigen does (the size assertion is compiled out and the loop runs over
        // the destination), and the reference relies on it: "Eigen::Vector2d pt_H1 = Edges_HYPO1_final.row(0)" with a 1x4
        // row (EdgeSketch_Core.cpp:648-649, :719).
        EIGEN_SHIM_CHECK((dst.rows() == 1 || dst.cols() == 1) && (sr == 1 || sc == 1) && dst.size() <= src.size(),
                         "assignment of mismatched shapes");
        const Index n = dst.size() < src.size() ? dst.size() : src.size();
        for (Index k = 0; k < n; k++) dst.coeffRef(k) = S(src.coeff(k));
    }
}
}  // namespace internal

// =====================================================================================
// WritableBase: assignment, compound assignment, setters, comma initialiser
// =====================================================================================
template <class D> class WritableBase : public DenseBase<D> {
public:
    typedef DenseBase<D> Base;
    typedef typename Base::Scalar Scalar;
    using Base::derived;
    using Base::operator();
    using Base::operator[];
    using Base::x;
    using Base::y;
    using Base::z;
    using Base::w;

    Scalar& operator()(Index i, Index j) { return this->coeffRef(i, j); }
    Scalar& operator()(Index k) { return this->coeffRef(k); }
    Scalar& operator[](Index k) { return this->coeffRef(k); }
    Scalar& x() { return this->coeffRef(0); }
    Scalar& y() { return this->coeffRef(1); }
    Scalar& z() { return this->coeffRef(2); }
    Scalar& w() { return this->coeffRef(3); }

    template <class O> D& assignFrom(const DenseBase<O>& o) { internal::assign(derived(), o); return derived(); }

    template <class O> D& operator+=(const DenseBase<O>& o) {
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = this->coeff(i, j) + lin_or_ij(o, i, j);
        return derived();
    }
    template <class O> D& operator-=(const DenseBase<O>& o) {
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = this->coeff(i, j) - lin_or_ij(o, i, j);
        return derived();
    }
    template <class T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0> D& operator*=(const T& s) {
        const Scalar ss = Scalar(s);
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = this->coeff(i, j) * ss;
        return derived();
    }
    template <class T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0> D& operator/=(const T& s) {
        const Scalar ss = Scalar(s);
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = this->coeff(i, j) / ss;
        return derived();
    }

    D& setConstant(const Scalar& v) {
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = v;
        return derived();
    }
    D& fill(const Scalar& v) { return setConstant(v); }
    D& setZero() { return setConstant(Scalar(0)); }
    D& setOnes() { return setConstant(Scalar(1)); }
    D& setIdentity() {
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = Scalar(i == j ? 1 : 0);
        return derived();
    }
    void normalize() {
        const Scalar n = this->norm();   // Eigen 3.3: *this /= norm()
        for (Index j = 0; j < derived().cols(); j++)
            for (Index i = 0; i < derived().rows(); i++) this->coeffRef(i, j) = this->coeff(i, j) / n;
    }

    template <class T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0>
    CommaInitializer<D> operator<<(const T& s) { return CommaInitializer<D>(derived(), Scalar(s)); }
    template <class O> CommaInitializer<D> operator<<(const DenseBase<O>& o) { return CommaInitializer<D>(derived(), o); }

private:
    template <class O> static typename O::Scalar lin_or_ij(const DenseBase<O>& o, Index i, Index j) {
        const O& d = o.derived();
        if (d.rows() == 0 || d.cols() == 0) return typename O::Scalar(0);
        return (i < d.rows() && j < d.cols()) ? d.coeff(i, j) : d.coeff(i + j);   // vector across orientation
    }
};

// Eigen's comma initialiser: scalars fill the current row left to right; a block starts a new band of
// rows when the previous band is full.  finished() hands the target back.
template <class D> class CommaInitializer {
public:
    typedef typename internal::traits<D>::Scalar Scalar;
    CommaInitializer(D& x, const Scalar& s) : m_x(x), m_row(0), m_col(1), m_band(1) { m_x.coeffRef(0, 0) = s; }
    template <class O> CommaInitializer(D& x, const DenseBase<O>& o) : m_x(x), m_row(0), m_col(0), m_band(o.derived().rows()) { put(o); }
    template <class T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0>
    CommaInitializer& operator,(const T& s) {
        if (m_col == m_x.cols()) { m_row += m_band; m_col = 0; m_band = 1; }
        m_x.coeffRef(m_row, m_col++) = Scalar(s);
        return *this;
    }
    template <class O> CommaInitializer& operator,(const DenseBase<O>& o) {
        if (m_col == m_x.cols()) { m_row += m_band; m_col = 0; m_band = o.derived().rows(); }
        put(o);
        return *this;
    }
    D& finished() { return m_x; }

private:
    template <class O> void put(const DenseBase<O>& o_) {
        const O& o = o_.derived();
        for (Index j = 0; j < o.cols(); j++)
            for (Index i = 0; i < o.rows(); i++) m_x.coeffRef(m_row + i, m_col + j) = Scalar(o.coeff(i, j));
        m_col += o.cols();
    }
    D& m_x;
    Index m_row, m_col, m_band;
};

// =====================================================================================
// Plain objects.  PlainBase holds what Matrix and Array share.
// =====================================================================================
template <class D, class S, int R, int C> class PlainBase : public WritableBase<D> {
public:
    typedef WritableBase<D> WBase;
    typedef S Scalar;

    Index rows() const { return m_s.rows(); }
    Index cols() const { return m_s.cols(); }
    Index outerStride() const { return m_s.rows(); }
    S* data() { return m_s.data(); }
    const S* data() const { return m_s.data(); }

    void resize(Index r, Index c) { m_s.resize(r, c); }
    void resize(Index n) {
        if constexpr (R == 1) m_s.resize(1, n); else m_s.resize(n, C == Dynamic ? 1 : C);
    }
    void conservativeResize(Index r, Index c) { m_s.conservativeResize(r, c); }
    void conservativeResize(Index n) {
        if constexpr (R == 1) m_s.conservativeResize(1, n); else m_s.conservativeResize(n, C == Dynamic ? 1 : C);
    }
    // shape this object so that an (sr x sc) source can be assigned (vectors may change orientation)
    void resizeLike(Index sr, Index sc) {
        if (rows() == sr && cols() == sc) return;
        if ((R == Dynamic || R == sr) && (C == Dynamic || C == sc)) { m_s.resize(sr, sc); return; }
        if ((R == Dynamic || R == sc) && (C == Dynamic || C == sr) && (sr == 1 || sc == 1)) { m_s.resize(sc, sr); return; }
    }

    static D Constant(Index r, Index c, const S& v) { D m; m.resize(r, c); m.setConstant(v); return m; }
    static D Constant(Index n, const S& v) { D m; m.resize(n); m.setConstant(v); return m; }
    static D Constant(const S& v) { D m; m.setConstant(v); return m; }
    static D Zero(Index r, Index c) { return Constant(r, c, S(0)); }
    static D Zero(Index n) { return Constant(n, S(0)); }
    static D Zero() { return Constant(S(0)); }
    static D Ones(Index r, Index c) { return Constant(r, c, S(1)); }
    static D Ones(Index n) { return Constant(n, S(1)); }
    static D Ones() { return Constant(S(1)); }
    static D Identity() { D m; m.setIdentity(); return m; }
    static D Identity(Index r, Index c) { D m; m.resize(r, c); m.setIdentity(); return m; }
    static D Unit(Index k) { D m; m.setZero(); m.coeffRef(k) = S(1); return m; }
    static D UnitX() { return Unit(0); }
    static D UnitY() { return Unit(1); }
    static D UnitZ() { return Unit(2); }
    static D UnitW() { return Unit(3); }

protected:
    PlainBase() {}
    // the 1..4-argument constructors of Eigen: sizes for run-time-sized objects, coefficients for
    // fixed-size vectors of matching length
    template <class T0> void init1(const T0& a) {
        if constexpr (R != Dynamic && C != Dynamic) { static_assert(R * C == 1, "1-argument constructor on a fixed-size object"); m_s.d[0] = S(a); }
        else resize(Index(a));
    }
    template <class T0, class T1> void init2(const T0& a, const T1& b) {
        if constexpr (R != Dynamic && C != Dynamic) { static_assert(R * C == 2, "2 coefficients need a 2-vector"); m_s.d[0] = S(a); m_s.d[1] = S(b); }
        else m_s.resize(Index(a), Index(b));
    }
    internal::Storage<S, R, C> m_s;
};

#define EIGEN_SHIM_PLAIN_CTORS(Class)                                                                          \
    Class() {}                                                                                                 \
    Class(const Class&) = default;                                                                             \
    Class(Class&&) = default;                                                                                  \
    Class& operator=(const Class&) = default;                                                                  \
    Class& operator=(Class&&) = default;                                                                       \
    template <class O> Class(const DenseBase<O>& o) { internal::assign(*this, o); }                            \
    template <class O> Class& operator=(const DenseBase<O>& o) { internal::assign(*this, o); return *this; }   \
    template <class T0, typename std::enable_if<std::is_arithmetic<T0>::value, int>::type = 0>                 \
    explicit Class(const T0& a) { this->init1(a); }                                                            \
    template <class T0, class T1, typename std::enable_if<std::is_arithmetic<T0>::value && std::is_arithmetic<T1>::value, int>::type = 0> \
    Class(const T0& a, const T1& b) { this->init2(a, b); }                                                     \
    template <class T0, class T1, class T2, typename std::enable_if<std::is_arithmetic<T0>::value && std::is_arithmetic<T1>::value && std::is_arithmetic<T2>::value, int>::type = 0> \
    Class(const T0& a, const T1& b, const T2& c) {                                                             \
        static_assert(R != Dynamic && C != Dynamic && R * C == 3, "3 coefficients need a fixed 3-vector");     \
        this->m_s.d[0] = S(a); this->m_s.d[1] = S(b); this->m_s.d[2] = S(c);                                   \
    }                                                                                                          \
    template <class T0, class T1, class T2, class T3, typename std::enable_if<std::is_arithmetic<T0>::value && std::is_arithmetic<T3>::value, int>::type = 0> \
    Class(const T0& a, const T1& b, const T2& c, const T3& d) {                                                \
        static_assert(R != Dynamic && C != Dynamic && R * C == 4, "4 coefficients need a fixed 4-vector");     \
        this->m_s.d[0] = S(a); this->m_s.d[1] = S(b); this->m_s.d[2] = S(c); this->m_s.d[3] = S(d);            \
    }

template <class S, int R, int C, int Opt, int MR, int MC>
class Matrix : public PlainBase<Matrix<S, R, C, Opt, MR, MC>, S, R, C> {
    static_assert((Opt & RowMajor) == 0, "the shim stores column-major only");
public:
    typedef PlainBase<Matrix, S, R, C> PBase;
    EIGEN_SHIM_PLAIN_CTORS(Matrix)
};

template <class S, int R, int C, int Opt, int MR, int MC>
class Array : public PlainBase<Array<S, R, C, Opt, MR, MC>, S, R, C> {
public:
    typedef PlainBase<Array, S, R, C> PBase;
    EIGEN_SHIM_PLAIN_CTORS(Array)
};

// =====================================================================================
// Views
// =====================================================================================
template <class S, int R, int C> class Block : public WritableBase<Block<S, R, C>> {
public:
    typedef S Scalar;
    Block(S* p, Index r, Index c, Index stride, Index chk_i = 0, Index chk_n = 1) : m_p(p), m_r(r), m_c(c), m_stride(stride) {
        EIGEN_SHIM_CHECK(chk_i >= 0 && chk_i < chk_n, "row/col/segment index");
        (void)chk_i; (void)chk_n;
    }
    Block(const Block&) = default;
    Index rows() const { return m_r; }
    Index cols() const { return m_c; }
    Index outerStride() const { return m_stride; }
    S* data() const { return m_p; }
    void resizeLike(Index, Index) {}
    // assignment writes through the view
    Block& operator=(const Block& o) { internal::assign(*this, o); return *this; }
    template <class O> Block& operator=(const DenseBase<O>& o) { internal::assign(*this, o); return *this; }
    template <class O> const Block& operator=(const DenseBase<O>& o) const { Block b(*this); internal::assign(b, o); return *this; }

private:
    S* m_p;
    Index m_r, m_c, m_stride;
};

template <class P, int MapOpt> class Map : public WritableBase<Map<P, MapOpt>> {
public:
    typedef typename internal::traits<P>::Scalar Scalar;
    enum { R = internal::traits<P>::Rows, C = internal::traits<P>::Cols };
    Map(const Scalar* p, Index n) : m_p(const_cast<Scalar*>(p)), m_r(int(R) == 1 ? 1 : n), m_c(int(R) == 1 ? n : (int(C) == int(Dynamic) ? 1 : Index(C))) {}
    Map(const Scalar* p, Index r, Index c) : m_p(const_cast<Scalar*>(p)), m_r(r), m_c(c) {}
    explicit Map(const Scalar* p) : m_p(const_cast<Scalar*>(p)), m_r(R), m_c(C) { static_assert(R != Dynamic && C != Dynamic, "sizes needed"); }
    Map(const Map&) = default;
    Index rows() const { return m_r; }
    Index cols() const { return m_c; }
    Index outerStride() const { return m_r; }
    Scalar* data() const { return m_p; }
    void resizeLike(Index, Index) {}
    Map& operator=(const Map& o) { internal::assign(*this, o); return *this; }
    template <class O> Map& operator=(const DenseBase<O>& o) { internal::assign(*this, o); return *this; }

private:
    Scalar* m_p;
    Index m_r, m_c;
};

// ---- the usual typedefs
#define EIGEN_SHIM_TYPEDEFS(Type, Suffix)                          \
    typedef Matrix<Type, 2, 2> Matrix2##Suffix;                    \
    typedef Matrix<Type, 3, 3> Matrix3##Suffix;                    \
    typedef Matrix<Type, 4, 4> Matrix4##Suffix;                    \
    typedef Matrix<Type, Dynamic, Dynamic> MatrixX##Suffix;        \
    typedef Matrix<Type, 2, 1> Vector2##Suffix;                    \
    typedef Matrix<Type, 3, 1> Vector3##Suffix;                    \
    typedef Matrix<Type, 4, 1> Vector4##Suffix;                    \
    typedef Matrix<Type, Dynamic, 1> VectorX##Suffix;              \
    typedef Matrix<Type, 1, 2> RowVector2##Suffix;                 \
    typedef Matrix<Type, 1, 3> RowVector3##Suffix;                 \
    typedef Matrix<Type, 1, 4> RowVector4##Suffix;                 \
    typedef Matrix<Type, 1, Dynamic> RowVectorX##Suffix;           \
    typedef Array<Type, Dynamic, Dynamic> ArrayXX##Suffix;         \
    typedef Array<Type, Dynamic, 1> ArrayX##Suffix;
EIGEN_SHIM_TYPEDEFS(double, d)
EIGEN_SHIM_TYPEDEFS(float, f)
EIGEN_SHIM_TYPEDEFS(int, i)
#undef EIGEN_SHIM_TYPEDEFS

}  // namespace Eigen
