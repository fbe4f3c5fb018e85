// TEST INFRASTRUCTURE — stand-in for <RcppArmadillo.h> so that the reference's OWN translation units
// (/root/reference/src/*.cpp, compiled where they lie, never copied) build without R, Rcpp or Armadillo.
// It provides just the subset of the Armadillo / Rcpp / Rmath interfaces those files use, with eager
// evaluation and plain loops. Third-party arithmetic that the reference delegates to libraries is restated:
//   arma::chol(.,"lower")        LAPACK dpotrf there -> unblocked left-looking column Cholesky here
//   arma::solve(trimatl/trimatu) LAPACK dtrtrs there -> forward / back substitution here
//   mat * vec, mat * mat         BLAS dgemv/dgemm there -> column-sweep (axpy) loops here
//   R::pnorm                     Rmath nmath/pnorm.c there -> Cody (1969) rational Chebyshev approximation here
//   R::rnorm                     R's RNG there -> normals handed in by the caller (refshim_set_normals)
// Only oracle/Makefile's `ref` target includes this header. Nothing under gpirt_b200/ does.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <limits>
#include <memory>
#include <numeric>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#ifndef M_2PI
#define M_2PI 6.283185307179586476925286766559  // Rmath.h
#endif

namespace arma {

#if defined(ARMA_64BIT_WORD)
typedef unsigned long long uword;
#else
typedef unsigned int uword;
#endif

namespace fill {
struct zeros_t {};
struct ones_t {};
struct eye_t {};
struct value_t { double v; };
static const zeros_t zeros{};
static const ones_t ones{};
static const eye_t eye{};
inline value_t value(double v) { return value_t{v}; }
}  // namespace fill

template <class T> class Col;
template <class T> class Row;
template <class T> class Mat;
template <class T> struct subview_col;
template <class T> struct subview_row;

// assignable single element (beta.slice(h).col(j).row(p) = x)
template <class T>
struct elem_ref {
    T* p;
    elem_ref& operator=(T v) { *p = v; return *this; }
    operator T() const { return *p; }
};

// contiguous piece of a column
template <class T>
struct subview_col {
    T* p;
    uword n_elem, n_rows;
    static const uword n_cols = 1;
    subview_col(T* p_, uword n) : p(p_), n_elem(n), n_rows(n) {}
    T& operator()(uword i) const { return p[i]; }
    T& operator[](uword i) const { return p[i]; }
    subview_col subvec(uword a, uword b) const { return subview_col(p + a, b - a + 1); }
    subview_col head(uword k) const { return subview_col(p, k); }
    elem_ref<T> row(uword i) const { return elem_ref<T>{p + i}; }
    Row<T> t() const;
    const subview_col& operator=(const subview_col& o) const {
        if (o.n_elem != n_elem) throw std::logic_error("subview_col: size mismatch");
        std::memmove(p, o.p, n_elem * sizeof(T));
        return *this;
    }
    const subview_col& operator=(const Col<T>& o) const;
    const subview_col& operator=(const Mat<T>& o) const;
};

// strided piece of a row
template <class T>
struct subview_row {
    T* p;
    uword stride, n_elem, n_cols;
    static const uword n_rows = 1;
    subview_row(T* p_, uword stride_, uword n) : p(p_), stride(stride_), n_elem(n), n_cols(n) {}
    T& operator()(uword i) const { return p[i * stride]; }
    T& operator[](uword i) const { return p[i * stride]; }
    subview_row subvec(uword a, uword b) const { return subview_row(p + a * stride, stride, b - a + 1); }
    Col<T> t() const;
    const subview_row& operator=(const subview_row& o) const {
        if (o.n_elem != n_elem) throw std::logic_error("subview_row: size mismatch");
        for (uword i = 0; i < n_elem; ++i) p[i * stride] = o.p[i * o.stride];
        return *this;
    }
    const subview_row& operator=(const Row<T>& o) const;
};

template <class T>
struct diagview {
    T* p;
    uword stride, n;
    void operator+=(T v) const { for (uword i = 0; i < n; ++i) p[i * stride] += v; }
};

template <class T>
class Col {
public:
    std::vector<T> store;
    uword n_elem = 0, n_rows = 0;
    static const uword n_cols = 1;
    void set_size(uword n) { store.resize(n); n_elem = n_rows = n; }
    Col() {}
    explicit Col(uword n) { set_size(n); }
    Col(uword n, fill::zeros_t) { set_size(n); std::fill(store.begin(), store.end(), T(0)); }
    Col(uword n, fill::ones_t) { set_size(n); std::fill(store.begin(), store.end(), T(1)); }
    Col(uword n, fill::value_t v) { set_size(n); std::fill(store.begin(), store.end(), T(v.v)); }
    Col(const subview_col<T>& s) { set_size(s.n_elem); std::copy(s.p, s.p + s.n_elem, store.begin()); }
    Col(const Mat<T>& m);
    Col& operator=(const subview_col<T>& s) { set_size(s.n_elem); std::copy(s.p, s.p + s.n_elem, store.begin()); return *this; }
    Col& operator=(const Mat<T>& m);
    T* memptr() { return store.data(); }
    const T* memptr() const { return store.data(); }
    T& operator()(uword i) { return store[i]; }
    const T& operator()(uword i) const { return store[i]; }
    T& operator[](uword i) { return store[i]; }
    const T& operator[](uword i) const { return store[i]; }
    subview_col<T> head(uword k) const { return subview_col<T>(const_cast<T*>(store.data()), k); }
    subview_col<T> subvec(uword a, uword b) const { return subview_col<T>(const_cast<T*>(store.data()) + a, b - a + 1); }
    Col elem(const Col<uword>& idx) const {
        Col out(idx.n_elem);
        for (uword k = 0; k < idx.n_elem; ++k) out[k] = store[idx[k]];
        return out;
    }
    Row<T> t() const;
    T min() const { return *std::min_element(store.begin(), store.end()); }
    T max() const { return *std::max_element(store.begin(), store.end()); }
    void zeros() { std::fill(store.begin(), store.end(), T(0)); }
};

template <class T>
class Row {
public:
    std::vector<T> store;
    uword n_elem = 0, n_cols = 0;
    static const uword n_rows = 1;
    Row() {}
    explicit Row(uword n) : store(n), n_elem(n), n_cols(n) {}
    T& operator()(uword i) { return store[i]; }
    const T& operator()(uword i) const { return store[i]; }
    Col<T> t() const { Col<T> c(n_elem); std::copy(store.begin(), store.end(), c.store.begin()); return c; }
};

template <class T>
Row<T> subview_col<T>::t() const { Row<T> r(n_elem); std::copy(p, p + n_elem, r.store.begin()); return r; }
template <class T>
Col<T> subview_row<T>::t() const { Col<T> c(n_elem); for (uword i = 0; i < n_elem; ++i) c[i] = p[i * stride]; return c; }
template <class T>
Row<T> Col<T>::t() const { Row<T> r(n_elem); std::copy(store.begin(), store.end(), r.store.begin()); return r; }
template <class T>
const subview_col<T>& subview_col<T>::operator=(const Col<T>& o) const {
    if (o.n_elem != n_elem) throw std::logic_error("subview_col = vec: size mismatch");
    std::copy(o.store.begin(), o.store.end(), p);
    return *this;
}
template <class T>
const subview_row<T>& subview_row<T>::operator=(const Row<T>& o) const {
    if (o.n_elem != n_elem) throw std::logic_error("subview_row = rowvec: size mismatch");
    for (uword i = 0; i < n_elem; ++i) p[i * stride] = o.store[i];
    return *this;
}

// ---- Mat: column-major; either owns its storage or views a slice of a cube ----------------------
template <class T>
class Mat {
public:
    std::vector<T> store;
    T* mem = nullptr;
    bool view = false;
    uword n_rows = 0, n_cols = 0, n_elem = 0;

    void set_size(uword r, uword c) {
        if (view) {
            if (r != n_rows || c != n_cols) throw std::logic_error("Mat view: size mismatch");
            return;
        }
        store.resize(size_t(r) * c);
        mem = store.data();
        n_rows = r; n_cols = c; n_elem = r * c;
    }
    Mat() {}
    Mat(uword r, uword c) { set_size(r, c); }
    Mat(uword r, uword c, fill::zeros_t) { set_size(r, c); zeros(); }
    Mat(uword r, uword c, fill::ones_t) { set_size(r, c); std::fill(mem, mem + n_elem, T(1)); }
    Mat(uword r, uword c, fill::eye_t) { set_size(r, c); zeros(); for (uword i = 0; i < std::min(r, c); ++i) at(i, i) = T(1); }
    static Mat make_view(T* p, uword r, uword c) { Mat m; m.mem = p; m.view = true; m.n_rows = r; m.n_cols = c; m.n_elem = r * c; return m; }
    Mat(const Mat& o) { set_size(o.n_rows, o.n_cols); std::copy(o.mem, o.mem + o.n_elem, mem); }
    Mat(Mat&& o) noexcept {
        if (o.view) { view = true; mem = o.mem; n_rows = o.n_rows; n_cols = o.n_cols; n_elem = o.n_elem; }  // only used while a cube builds its slice table
        else { store = std::move(o.store); mem = store.data(); n_rows = o.n_rows; n_cols = o.n_cols; n_elem = o.n_elem; o.n_rows = o.n_cols = o.n_elem = 0; o.mem = nullptr; }
    }
    Mat(const subview_col<T>& s) { set_size(s.n_elem, 1); std::copy(s.p, s.p + s.n_elem, mem); }
    Mat(const Col<T>& c) { set_size(c.n_elem, 1); std::copy(c.store.begin(), c.store.end(), mem); }
    Mat& operator=(const Mat& o) {
        if (this == &o) return *this;
        set_size(o.n_rows, o.n_cols);
        std::copy(o.mem, o.mem + o.n_elem, mem);
        return *this;
    }
    Mat& operator=(Mat&& o) {
        if (view || o.view) return (*this = static_cast<const Mat&>(o));
        store = std::move(o.store); mem = store.data(); n_rows = o.n_rows; n_cols = o.n_cols; n_elem = o.n_elem;
        o.n_rows = o.n_cols = o.n_elem = 0; o.mem = nullptr;
        return *this;
    }
    T& at(uword i, uword j) { return mem[i + size_t(j) * n_rows]; }
    const T& at(uword i, uword j) const { return mem[i + size_t(j) * n_rows]; }
    // Armadillo checks bounds in operator() unless ARMA_NO_DEBUG is defined (RcppArmadillo does not define it); .at() is unchecked
    void check_bounds(uword i, uword j) const { if (i >= n_rows || j >= n_cols) throw std::logic_error("Mat::operator(): index out of bounds"); }
    T& operator()(uword i, uword j) { check_bounds(i, j); return at(i, j); }
    const T& operator()(uword i, uword j) const { check_bounds(i, j); return at(i, j); }
    T& operator()(uword i) { return mem[i]; }
    const T& operator()(uword i) const { return mem[i]; }
    T& operator[](uword i) { return mem[i]; }
    const T& operator[](uword i) const { return mem[i]; }
    T* memptr() { return mem; }
    const T* memptr() const { return mem; }
    subview_col<T> col(uword j) const { return subview_col<T>(const_cast<T*>(mem) + size_t(j) * n_rows, n_rows); }
    subview_row<T> row(uword i) const { return subview_row<T>(const_cast<T*>(mem) + i, n_rows, n_cols); }
    Mat rows(const Col<uword>& idx) const {
        Mat out(idx.n_elem, n_cols);
        for (uword j = 0; j < n_cols; ++j)
            for (uword k = 0; k < idx.n_elem; ++k) out.at(k, j) = at(idx[k], j);
        return out;
    }
    diagview<T> diag() { return diagview<T>{mem, n_rows + 1, std::min(n_rows, n_cols)}; }
    Mat t() const {
        Mat out(n_cols, n_rows);
        for (uword j = 0; j < n_cols; ++j)
            for (uword i = 0; i < n_rows; ++i) out.at(j, i) = at(i, j);
        return out;
    }
    void zeros() { std::fill(mem, mem + n_elem, T(0)); }
    void clamp(T lo, T hi) { for (uword i = 0; i < n_elem; ++i) mem[i] = mem[i] < lo ? lo : (mem[i] > hi ? hi : mem[i]); }
    T min() const { return *std::min_element(mem, mem + n_elem); }
    T max() const { return *std::max_element(mem, mem + n_elem); }
    void swap(Mat& o) {
        if (view || o.view) throw std::logic_error("Mat::swap on a view");
        store.swap(o.store); std::swap(n_rows, o.n_rows); std::swap(n_cols, o.n_cols); std::swap(n_elem, o.n_elem);
        mem = store.data(); o.mem = o.store.data();
    }
    Mat& operator-=(const Mat& o) {
        if (o.n_rows != n_rows || o.n_cols != n_cols) throw std::logic_error("Mat -=: size mismatch");
        for (uword i = 0; i < n_elem; ++i) mem[i] -= o.mem[i];
        return *this;
    }
};

template <class T>
Col<T>::Col(const Mat<T>& m) {
    if (m.n_cols != 1 && m.n_elem != 0) throw std::logic_error("vec(mat): not a column");
    set_size(m.n_elem); std::copy(m.mem, m.mem + m.n_elem, store.begin());
}
template <class T>
Col<T>& Col<T>::operator=(const Mat<T>& m) {
    if (m.n_cols != 1 && m.n_elem != 0) throw std::logic_error("vec = mat: not a column");
    set_size(m.n_elem); std::copy(m.mem, m.mem + m.n_elem, store.begin());
    return *this;
}
template <class T>
const subview_col<T>& subview_col<T>::operator=(const Mat<T>& o) const {
    if (o.n_elem != n_elem) throw std::logic_error("subview_col = mat: size mismatch");
    std::copy(o.mem, o.mem + o.n_elem, p);
    return *this;
}

typedef Col<double> vec;
typedef Col<uword> uvec;
typedef Row<double> rowvec;
typedef Mat<double> mat;
typedef Mat<uword> umat;

// ---- cube -----------------------------------------------------------------------------------------
class cube;
struct subview_cube_col {  // X.col(c) of a cube: n_rows x 1 x n_slices, assignable from an n_rows x n_slices matrix
    cube& c;
    uword j;
    void operator=(const mat& m) const;
};

class cube {
public:
    std::vector<double> store;
    std::vector<mat> slices;
    uword n_rows = 0, n_cols = 0, n_slices = 0, n_elem = 0;
    void rebuild() {
        slices.clear();
        slices.reserve(n_slices);
        for (uword s = 0; s < n_slices; ++s) slices.push_back(mat::make_view(store.data() + size_t(s) * n_rows * n_cols, n_rows, n_cols));
    }
    void set_size(uword r, uword c, uword s) {
        n_rows = r; n_cols = c; n_slices = s; n_elem = r * c * s;
        store.resize(size_t(n_elem));
        rebuild();
    }
    cube() {}
    cube(uword r, uword c, uword s) { set_size(r, c, s); }
    cube(uword r, uword c, uword s, fill::zeros_t) { set_size(r, c, s); zeros(); }
    cube(const cube& o) { set_size(o.n_rows, o.n_cols, o.n_slices); store = o.store; rebuild(); }
    cube(cube&& o) noexcept : store(std::move(o.store)), n_rows(o.n_rows), n_cols(o.n_cols), n_slices(o.n_slices), n_elem(o.n_elem) {
        rebuild(); o.n_rows = o.n_cols = o.n_slices = o.n_elem = 0; o.slices.clear();
    }
    cube& operator=(const cube& o) {
        if (this == &o) return *this;
        n_rows = o.n_rows; n_cols = o.n_cols; n_slices = o.n_slices; n_elem = o.n_elem; store = o.store; rebuild();
        return *this;
    }
    cube& operator=(cube&& o) {
        store = std::move(o.store); n_rows = o.n_rows; n_cols = o.n_cols; n_slices = o.n_slices; n_elem = o.n_elem; rebuild();
        o.n_rows = o.n_cols = o.n_slices = o.n_elem = 0; o.slices.clear();
        return *this;
    }
    mat& slice(uword s) { return slices[s]; }
    const mat& slice(uword s) const { return slices[s]; }
    double& operator()(uword i, uword j, uword s) { return store[i + size_t(j) * n_rows + size_t(s) * n_rows * n_cols]; }
    const double& operator()(uword i, uword j, uword s) const { return store[i + size_t(j) * n_rows + size_t(s) * n_rows * n_cols]; }
    subview_cube_col col(uword j) { return subview_cube_col{*this, j}; }
    mat row(uword i) const {  // 1 x n_cols x n_slices read as an n_cols x n_slices matrix
        mat out(n_cols, n_slices);
        for (uword s = 0; s < n_slices; ++s)
            for (uword j = 0; j < n_cols; ++j) out(j, s) = (*this)(i, j, s);
        return out;
    }
    void zeros() { std::fill(store.begin(), store.end(), 0.0); }
    void swap(cube& o) {
        store.swap(o.store); std::swap(n_rows, o.n_rows); std::swap(n_cols, o.n_cols); std::swap(n_slices, o.n_slices); std::swap(n_elem, o.n_elem);
        rebuild(); o.rebuild();
    }
    double* memptr() { return store.data(); }
    const double* memptr() const { return store.data(); }
};
inline void subview_cube_col::operator=(const mat& m) const {
    if (m.n_rows != c.n_rows || m.n_cols != c.n_slices) throw std::logic_error("cube.col() = mat: size mismatch");
    for (uword s = 0; s < c.n_slices; ++s)
        for (uword i = 0; i < c.n_rows; ++i) c(i, j, s) = m(i, s);
}

template <class T>
class field {
public:
    std::vector<T> store;
    uword n_rows = 0, n_cols = 0, n_elem = 0;
    field() {}
    explicit field(uword n) : store(n), n_rows(n), n_cols(1), n_elem(n) {}
    field(uword r, uword c) : store(size_t(r) * c), n_rows(r), n_cols(c), n_elem(r * c) {}
    T& operator()(uword i) { return store[i]; }
    const T& operator()(uword i) const { return store[i]; }
    T& operator()(uword i, uword j) { return store[i + size_t(j) * n_rows]; }
    const T& operator()(uword i, uword j) const { return store[i + size_t(j) * n_rows]; }
};

// ---- generators -----------------------------------------------------------------------------------
template <class V> struct ones_impl;
template <> struct ones_impl<vec> { static vec make(uword n) { return vec(n, fill::ones); } };
template <> struct ones_impl<mat> { static mat make(uword r, uword c) { return mat(r, c, fill::ones); } };
template <class V> inline V ones(uword n) { return ones_impl<V>::make(n); }
template <class V> inline V ones(uword r, uword c) { return ones_impl<V>::make(r, c); }

// regspace(start, delta, end): N = 1 + floor((end - start) / delta), x[i] = start + i * delta
template <class V> struct regspace_impl;
template <> struct regspace_impl<vec> {
    static vec make(double a, double d, double b) {
        const uword N = uword(1) + uword(std::floor((b - a) / d));
        vec x(N);
        for (uword i = 0; i < N; ++i) x[i] = a + double(i) * d;
        return x;
    }
};
template <> struct regspace_impl<uvec> {
    static uvec make(uword a, uword b) {
        if (b < a || b == uword(-1)) return uvec();
        uvec x(b - a + 1);
        for (uword i = 0; i < x.n_elem; ++i) x[i] = a + i;
        return x;
    }
};
template <class V> inline V regspace(double a, double d, double b) { return regspace_impl<V>::make(a, d, b); }
template <class V> inline V regspace(uword a, uword b) { return regspace_impl<V>::make(a, b); }

// linspace(start, end, N): x[i] = start + i * (end - start) / (N - 1), last element set to end
inline vec linspace_impl(double a, double b, uword N) {
    vec x(N);
    if (N == 1) { x[0] = b; return x; }
    const double delta = (b - a) / double(N - 1);
    for (uword i = 0; i + 1 < N; ++i) x[i] = a + double(i) * delta;
    x[N - 1] = b;
    return x;
}
template <class V = vec> inline V linspace(double a, double b, uword N) { return linspace_impl(a, b, N); }

// ---- element-wise and products --------------------------------------------------------------------
inline vec operator+(const vec& a, const vec& b) {
    if (a.n_elem != b.n_elem) throw std::logic_error("vec + vec: size mismatch");
    vec r(a.n_elem);
    for (uword i = 0; i < a.n_elem; ++i) r[i] = a[i] + b[i];
    return r;
}
inline vec operator+(const vec& a, const subview_col<double>& b) {
    if (a.n_elem != b.n_elem) throw std::logic_error("vec + subview: size mismatch");
    vec r(a.n_elem);
    for (uword i = 0; i < a.n_elem; ++i) r[i] = a[i] + b[i];
    return r;
}
inline vec operator+(const subview_col<double>& a, const subview_col<double>& b) { return vec(a) + b; }
inline mat operator+(const mat& a, double s) { mat r(a.n_rows, a.n_cols); for (uword i = 0; i < a.n_elem; ++i) r[i] = a[i] + s; return r; }
inline mat operator/(const mat& a, double s) { mat r(a.n_rows, a.n_cols); for (uword i = 0; i < a.n_elem; ++i) r[i] = a[i] / s; return r; }
inline mat pow(const mat& a, int e) { mat r(a.n_rows, a.n_cols); for (uword i = 0; i < a.n_elem; ++i) r[i] = std::pow(a[i], e); return r; }
inline vec pow(const vec& a, int e) { vec r(a.n_elem); for (uword i = 0; i < a.n_elem; ++i) r[i] = std::pow(a[i], e); return r; }

// y = A x in column-sweep (reference-BLAS dgemv 'N') order
inline void gemv_raw(const mat& A, const double* x, double* y) {
    for (uword i = 0; i < A.n_rows; ++i) y[i] = 0.0;
    for (uword j = 0; j < A.n_cols; ++j) {
        const double xj = x[j];
        const double* cj = A.mem + size_t(j) * A.n_rows;
        for (uword i = 0; i < A.n_rows; ++i) y[i] += cj[i] * xj;
    }
}
inline vec operator*(const mat& A, const vec& x) {
    if (A.n_cols != x.n_elem) throw std::logic_error("mat * vec: size mismatch");
    vec y(A.n_rows);
    gemv_raw(A, x.memptr(), y.memptr());
    return y;
}
inline vec operator*(const mat& A, const subview_col<double>& x) {
    if (A.n_cols != x.n_elem) throw std::logic_error("mat * subview_col: size mismatch");
    vec y(A.n_rows);
    gemv_raw(A, x.p, y.memptr());
    return y;
}
inline mat operator*(const mat& A, const mat& B) {
    if (A.n_cols != B.n_rows) throw std::logic_error("mat * mat: size mismatch");
    mat Cm(A.n_rows, B.n_cols);
#pragma omp parallel for schedule(static) if (A.n_rows * B.n_cols * A.n_cols > 4000000)
    for (uword j = 0; j < B.n_cols; ++j) gemv_raw(A, B.mem + size_t(j) * B.n_rows, Cm.mem + size_t(j) * A.n_rows);
    return Cm;
}

// ---- factorisation and triangular solves ----------------------------------------------------------
inline mat chol(const mat& S, const char* layout) {
    if (std::string(layout) != "lower") throw std::logic_error("refshim chol: only \"lower\"");
    if (S.n_rows != S.n_cols) throw std::logic_error("chol(): given matrix must be square sized");
    const uword n = S.n_rows;
    mat A(S);
    for (uword j = 0; j < n; ++j) {
        double* cj = A.mem + size_t(j) * n;
        for (uword k = 0; k < j; ++k) {
            const double* ck = A.mem + size_t(k) * n;
            const double ljk = ck[j];
            if (ljk == 0.0) continue;
            for (uword i = j; i < n; ++i) cj[i] -= ck[i] * ljk;
        }
        double d = cj[j];
        if (!(d > 0.0)) throw std::runtime_error("chol(): decomposition failed");
        d = std::sqrt(d);
        cj[j] = d;
        for (uword i = j + 1; i < n; ++i) cj[i] /= d;
        for (uword i = 0; i < j; ++i) cj[i] = 0.0;
    }
    return A;
}
struct tri_lower { const mat& m; };
struct tri_upper { mat m; };
inline tri_lower trimatl(const mat& m) { return tri_lower{m}; }
inline tri_upper trimatu(const mat& m) { return tri_upper{m}; }
inline mat solve(const tri_lower& Lw, const mat& B) {
    const mat& L = Lw.m;
    const uword n = L.n_rows;
    if (B.n_rows != n) throw std::logic_error("solve(): size mismatch");
    mat X(B);
#pragma omp parallel for schedule(static) if (X.n_cols > 8)
    for (uword c = 0; c < X.n_cols; ++c) {
        double* x = X.mem + size_t(c) * n;
        for (uword j = 0; j < n; ++j) {
            const double xj = x[j] / L.at(j, j);
            x[j] = xj;
            const double* cj = L.mem + size_t(j) * n;
            for (uword i = j + 1; i < n; ++i) x[i] -= cj[i] * xj;
        }
    }
    return X;
}
inline mat solve(const tri_upper& Uw, const mat& B) {
    const mat& U = Uw.m;
    const uword n = U.n_rows;
    if (B.n_rows != n) throw std::logic_error("solve(): size mismatch");
    mat X(B);
#pragma omp parallel for schedule(static) if (X.n_cols > 8)
    for (uword c = 0; c < X.n_cols; ++c) {
        double* x = X.mem + size_t(c) * n;
        for (uword jj = n; jj-- > 0;) {
            double s = x[jj];
            for (uword i = jj + 1; i < n; ++i) s -= U.at(jj, i) * x[i];
            x[jj] = s / U.at(jj, jj);
        }
    }
    return X;
}

// ---- queries ----------------------------------------------------------------------------------------
inline uvec find_finite(const subview_col<double>& s) {
    uvec out;
    for (uword i = 0; i < s.n_elem; ++i) if (std::isfinite(s[i])) out.store.push_back(i);
    out.n_elem = out.n_rows = out.store.size();
    return out;
}
inline uvec find_finite(const subview_row<double>& s) {
    uvec out;
    for (uword i = 0; i < s.n_elem; ++i) if (std::isfinite(s[i])) out.store.push_back(i);
    out.n_elem = out.n_rows = out.store.size();
    return out;
}
inline uvec vectorise(const umat& m) { uvec v(m.n_elem); std::copy(m.mem, m.mem + m.n_elem, v.store.begin()); return v; }
inline double mean(const uvec& v) { double s = 0.0; for (uword i = 0; i < v.n_elem; ++i) s += double(v[i]); return v.n_elem ? s / double(v.n_elem) : 0.0; }

// interp1(X, Y, XI, YI, "linear"): piecewise-linear interpolation on (X, Y) sorted by X; outside the range -> NaN
inline void interp1(const vec& X, const vec& Y, const vec& XI, vec& YI, const char* /*method*/) {
    std::vector<uword> ord(X.n_elem);
    std::iota(ord.begin(), ord.end(), uword(0));
    std::stable_sort(ord.begin(), ord.end(), [&](uword a, uword b) { return X[a] < X[b]; });
    std::vector<double> xs, ys;
    for (uword k : ord) {
        if (!xs.empty() && X[k] == xs.back()) continue;  // unique abscissae
        xs.push_back(X[k]); ys.push_back(Y[k]);
    }
    YI.set_size(XI.n_elem);
    for (uword q = 0; q < XI.n_elem; ++q) {
        const double x = XI[q];
        if (xs.empty() || x < xs.front() || x > xs.back()) { YI[q] = std::numeric_limits<double>::quiet_NaN(); continue; }
        size_t hi = std::lower_bound(xs.begin(), xs.end(), x) - xs.begin();
        if (xs[hi] == x) { YI[q] = ys[hi]; continue; }
        const size_t lo = hi - 1;
        const double w = (x - xs[lo]) / (xs[hi] - xs[lo]);
        YI[q] = (1.0 - w) * ys[lo] + w * ys[hi];
    }
}
inline void interp1(const vec& X, const subview_col<double>& Y, const vec& XI, vec& YI, const char* method) { interp1(X, vec(Y), XI, YI, method); }

}  // namespace arma

// ---- Rmath ------------------------------------------------------------------------------------------
extern "C" double refshim_next_normal(void);  // defined in ref_capi.cpp: the caller-supplied N(0,1) stream
extern "C" double refshim_next_unif(void);    // defined by whoever uses R::unif_rand (tests/rshim/driver.cpp)
namespace R {
// pnorm(x, mu, sigma, lower_tail, log_p): Cody (1969) as in nmath/pnorm.c (pnorm_both); only the
// lower_tail = 1, log_p = 0 combination the reference uses is restated.
inline double pnorm(double x, double mu, double sigma, int lower_tail, int log_p) {
    if (!lower_tail || log_p) throw std::logic_error("refshim pnorm: only lower tail, non-log");
    if (std::isnan(x) || std::isnan(mu) || std::isnan(sigma)) return x + mu + sigma;
    if (!std::isfinite(x) && mu == x) return std::numeric_limits<double>::quiet_NaN();
    if (sigma <= 0) {
        if (sigma < 0) return std::numeric_limits<double>::quiet_NaN();
        return (x < mu) ? 0.0 : 1.0;
    }
    const double pq = (x - mu) / sigma;
    if (!std::isfinite(pq)) return (x < mu) ? 0.0 : 1.0;
    x = pq;
    static const double a[5] = {2.2352520354606839287, 161.02823106855587881, 1067.6894854603709582, 18154.981253343561249,
                                0.065682337918207449113};
    static const double b[4] = {47.20258190468824187, 976.09855173777669322, 10260.932208618978205, 45507.789335026729956};
    static const double c[9] = {0.39894151208813466764, 8.8831497943883759412, 93.506656132177855979, 597.27027639480026226,
                                2494.5375852903726711,  6848.1904505362823326, 11602.651437647350124, 9842.7148383839780218,
                                1.0765576773720192317e-8};
    static const double d[8] = {22.266688044328115691, 235.38790178262499861, 1519.377599407554805,  6485.558298266760755,
                                18615.571640885098091, 34900.952721145977266, 38912.003286093271411, 19685.429676859990727};
    static const double p[6] = {0.21589853405795699,     0.1274011611602473639, 0.022235277870649807,
                                0.001421619193227893466, 2.9112874951168792e-5, 0.02307344176494017303};
    static const double q[5] = {1.28426009614491121, 0.468238212480865118, 0.0659881378689285515, 0.00378239633202758244,
                                7.29751555083966205e-5};
    const double M_SQRT_32 = 5.656854249492380195206754896838, M_1_SQRT_2PI = 0.398942280401432677939946059934;
    const double eps = std::numeric_limits<double>::epsilon() * 0.5;
    double xden, xnum, temp, del, xsq, cum;
    const double y = std::fabs(x);
    if (y <= 0.67448975) {
        if (y > eps) {
            xsq = x * x;
            xnum = a[4] * xsq;
            xden = xsq;
            for (int i = 0; i < 3; ++i) { xnum = (xnum + a[i]) * xsq; xden = (xden + b[i]) * xsq; }
        } else {
            xnum = xden = 0.0;
        }
        temp = x * (xnum + a[3]) / (xden + b[3]);
        return 0.5 + temp;
    }
    if (y <= M_SQRT_32) {
        xnum = c[8] * y;
        xden = y;
        for (int i = 0; i < 7; ++i) { xnum = (xnum + c[i]) * y; xden = (xden + d[i]) * y; }
        temp = (xnum + c[7]) / (xden + d[7]);
        xsq = std::trunc(y * 16) / 16;
        del = (y - xsq) * (y + xsq);
        cum = std::exp(-xsq * xsq * 0.5) * std::exp(-del * 0.5) * temp;
        return x > 0. ? 1.0 - cum : cum;
    }
    if (-37.5193 < x && x < 8.2924) {
        xsq = 1.0 / (x * x);
        xnum = p[5] * xsq;
        xden = xsq;
        for (int i = 0; i < 4; ++i) { xnum = (xnum + p[i]) * xsq; xden = (xden + q[i]) * xsq; }
        temp = xsq * (xnum + p[4]) / (xden + q[4]);
        temp = (M_1_SQRT_2PI - temp) / y;
        xsq = std::trunc(x * 16) / 16;
        del = (x - xsq) * (x + xsq);
        cum = std::exp(-xsq * xsq * 0.5) * std::exp(-del * 0.5) * temp;
        return x > 0. ? 1.0 - cum : cum;
    }
    return x > 0 ? 1.0 : 0.0;
}
// rnorm(mu, sigma) = mu + sigma * norm_rand() (nmath/rnorm.c); norm_rand() comes from the caller here
inline double rnorm(double mu, double sigma) {
    if (std::isnan(mu) || !std::isfinite(sigma) || sigma < 0.) return std::numeric_limits<double>::quiet_NaN();
    if (sigma == 0. || !std::isfinite(mu)) return mu;
    return mu + sigma * refshim_next_normal();
}
inline double unif_rand() { return refshim_next_unif(); }
}  // namespace R

// ---- Rcpp -------------------------------------------------------------------------------------------
inline void Rprintf(const char*, ...) {}
struct refshim_nil_t {};
static const refshim_nil_t R_NilValue{};

namespace Rcpp {
struct RObject {
    enum Kind { NIL, CUBE, VEC, LIST } kind = NIL;
    std::shared_ptr<arma::cube> cube;
    std::shared_ptr<arma::vec> vec;
    std::shared_ptr<std::vector<RObject>> items;
    std::shared_ptr<std::vector<std::string>> names;
};
inline RObject wrap(const arma::cube& c) { RObject o; o.kind = RObject::CUBE; o.cube = std::make_shared<arma::cube>(c); return o; }
inline RObject wrap(const arma::vec& v) { RObject o; o.kind = RObject::VEC; o.vec = std::make_shared<arma::vec>(v); return o; }
struct List;
inline RObject wrap(const RObject& o) { return o; }
struct NamedValue { std::string name; RObject value; };
template <class T> inline NamedValue Named(const char* name, const T& v) { return NamedValue{name, wrap(v)}; }
struct List : RObject {
    List() { kind = LIST; items = std::make_shared<std::vector<RObject>>(); names = std::make_shared<std::vector<std::string>>(); }
    explicit List(int n) : List() { items->resize(n); names->resize(n); }
    List(const refshim_nil_t&) { kind = NIL; }
    RObject& operator[](int i) { return (*items)[i]; }
    const RObject& operator[](const std::string& key) const {
        for (size_t i = 0; i < names->size(); ++i) if ((*names)[i] == key) return (*items)[i];
        throw std::out_of_range("List: no element named " + key);
    }
    template <class... A> static List create(const A&... a) {
        List l;
        const NamedValue arr[] = {a...};
        for (const NamedValue& nv : arr) { l.items->push_back(nv.value); l.names->push_back(nv.name); }
        return l;
    }
};
inline void checkUserInterrupt() {}
inline void stop(const std::string& msg) { throw std::runtime_error(msg); }  // Rcpp::stop -> R error through END_RCPP
struct NullStream { template <class T> NullStream& operator<<(const T&) { return *this; } };
static NullStream Rcout, Rcerr;
struct NumericVector {};
struct Function { template <class... A> void operator()(A...) const {} };
struct EnvSlot {
    operator Function() const { return Function(); }
    operator NumericVector() const { return NumericVector(); }
    EnvSlot& operator=(const NumericVector&) { return *this; }
};
struct Environment {
    Environment(const char*) {}
    EnvSlot operator[](const char*) { return EnvSlot(); }
};
}  // namespace Rcpp
