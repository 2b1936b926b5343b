// oracle/ref_stubs/f90rt.h -- TEST INFRASTRUCTURE.  Run-time support for the C++ that oracle/f90_to_cpp.py produces from the
// reference's Fortran 90 sources (fortran/raytracer.f95, util.f95, bmodel_dipole.f95, constants.f95): rank-1 and rank-2
// arrays with Fortran's element-wise expression semantics, 1-based indexing, sections, and the array intrinsics the
// sources use.  What this header decides (a Fortran compiler would): element-wise operations evaluate each element with
// the scalar operation in source order; SUM, DOT_PRODUCT, MAXVAL, MINVAL run over the elements in index order starting
// from the first (gfortran inlines them so without -ffast-math); the right-hand side of an array assignment is evaluated
// completely before the left-hand side is stored; COMPLEX arithmetic follows -fcx-fortran-rules (compile with that flag).
#pragma once
#include <math.h>
#include <complex>
#include <stdio.h>
#include <stdlib.h>
#include <map>
#include <memory>
#include <string>
#include <type_traits>
#include <vector>

typedef std::complex<double> cd;

template <class T> struct Vec;
template <class T> struct is_vec : std::false_type {};
template <class T> struct is_vec<Vec<T>> : std::true_type {};

template <class T> struct Vec {
  std::vector<T> d;
  bool allocated = false;
  Vec() {}
  explicit Vec(int n) : d((size_t)(n > 0 ? n : 0), T()), allocated(true) {}
  Vec(std::initializer_list<T> l) : d(l), allocated(true) {}
  int size() const { return (int)d.size(); }
  T& operator()(int i) { return d[(size_t)(i - 1)]; }
  const T& operator()(int i) const { return d[(size_t)(i - 1)]; }
  // whole-array assignment: conforming shapes; an unallocated / differently sized target takes the shape (allocatable LHS)
  template <class U> Vec& operator=(const Vec<U>& o) {
    std::vector<T> t(o.d.size());
    for (size_t i = 0; i < o.d.size(); ++i) t[i] = (T)o.d[i];
    d.swap(t); allocated = true; return *this;
  }
  Vec& operator=(const Vec& o) { if (this != &o) { d = o.d; allocated = true; } return *this; }
  template <class U, class = typename std::enable_if<!is_vec<U>::value>::type> Vec& operator=(const U& s) { for (auto& e : d) e = (T)s; return *this; }
  struct Sec {
    Vec& v; int lo, hi;
    template <class U> Sec& operator=(const Vec<U>& o) { for (int i = lo; i <= hi; ++i) v(i) = (T)o(i - lo + 1); return *this; }
    Sec& operator=(const Sec& o) { Vec t = o; return *this = t; }
    template <class U, class = typename std::enable_if<!is_vec<U>::value>::type> Sec& operator=(const U& s) { for (int i = lo; i <= hi; ++i) v(i) = (T)s; return *this; }
    operator Vec() const { Vec r(hi - lo + 1); for (int i = lo; i <= hi; ++i) r(i - lo + 1) = v(i); return r; }
  };
  Sec sec(int lo, int hi) { return Sec{*this, lo, hi}; }                                                             // as the target of an assignment
  Vec rsec(int lo, int hi) const { Vec r(hi - lo + 1); for (int i = lo; i <= hi; ++i) r(i - lo + 1) = (*this)(i); return r; }  // as a value
};

template <class T> struct Mat {  // column-major, 1-based
  std::vector<T> d;
  int n1 = 0, n2 = 0;
  bool allocated = false;
  Mat() {}
  Mat(int a, int b) : d((size_t)a * b, T()), n1(a), n2(b), allocated(true) {}
  T& operator()(int i, int j) { return d[(size_t)(i - 1) + (size_t)(j - 1) * n1]; }
  const T& operator()(int i, int j) const { return d[(size_t)(i - 1) + (size_t)(j - 1) * n1]; }
  struct Col {
    Mat& m; int j;
    template <class U> Col& operator=(const Vec<U>& o) { for (int i = 1; i <= m.n1; ++i) m(i, j) = (T)o(i); return *this; }
    template <class U, class = typename std::enable_if<!is_vec<U>::value>::type> Col& operator=(const U& s) { for (int i = 1; i <= m.n1; ++i) m(i, j) = (T)s; return *this; }
    operator Vec<T>() const { Vec<T> r(m.n1); for (int i = 1; i <= m.n1; ++i) r(i) = m(i, j); return r; }
  };
  Col col(int j) { return Col{*this, j}; }
  Vec<T> rcol(int j) const { Vec<T> r(n1); for (int i = 1; i <= n1; ++i) r(i) = (*this)(i, j); return r; }
  Vec<T> row(int i) const { Vec<T> r(n2); for (int j = 1; j <= n2; ++j) r(j) = (*this)(i, j); return r; }
  struct Block {
    Mat& m; int a, b, c, e;
    Block& operator=(const Mat& o) { for (int j = c; j <= e; ++j) for (int i = a; i <= b; ++i) m(i, j) = o(i - a + 1, j - c + 1); return *this; }
  };
  Block block(int a, int b, int c, int e) { return Block{*this, a, b, c, e}; }
};

// rank-3 view (an assumed-shape dummy, or the section a(ind,:,:,:) of a rank-4 array) and rank-4 view over storage in
// Fortran order that the harness owns
template <class T> struct Arr3 {
  T* p = nullptr; long s1 = 0, s2 = 0, s3 = 0; int n1 = 0, n2 = 0, n3 = 0;
  T& operator()(int i, int j, int k) const { return p[(i - 1) * s1 + (j - 1) * s2 + (k - 1) * s3]; }
  // the section (a1:b1, a2:b2, a3:b3); a scalar subscript i is passed as i:i and keeps its dimension with extent 1
  Arr3 sec3(int a1, int b1, int a2, int b2, int a3, int b3) const {
    Arr3 r = *this; r.p = &(*this)(a1, a2, a3); r.n1 = b1 - a1 + 1; r.n2 = b2 - a2 + 1; r.n3 = b3 - a3 + 1; return r; }
};
// value of a rank-3 array expression (elements are independent of one another: the order of evaluation is free; the right-hand
// side is complete before the assignment stores, as the language requires)
struct Val3 {
  std::vector<double> d; int n1 = 0, n2 = 0, n3 = 0;
  Val3(int a, int b, int c) : d((size_t)a * b * c), n1(a), n2(b), n3(c) {}
  double& operator()(int i, int j, int k) { return d[(size_t)(i - 1) + (size_t)n1 * ((j - 1) + (size_t)n2 * (k - 1))]; }
  double operator()(int i, int j, int k) const { return d[(size_t)(i - 1) + (size_t)n1 * ((j - 1) + (size_t)n2 * (k - 1))]; }
};
inline Val3 operator-(const Arr3<double>& a, const Arr3<double>& b) {
  Val3 r(a.n1, a.n2, a.n3);
  for (int k = 1; k <= a.n3; ++k) for (int j = 1; j <= a.n2; ++j) for (int i = 1; i <= a.n1; ++i) r(i, j, k) = a(i, j, k) - b(i, j, k);
  return r; }
inline Val3 operator/(Val3 a, double s) { for (double& v : a.d) v = v / s; return a; }
inline void F_assign3(const Arr3<double>& dst, const Val3& v) {
  for (int k = 1; k <= dst.n3; ++k) for (int j = 1; j <= dst.n2; ++j) for (int i = 1; i <= dst.n1; ++i) dst(i, j, k) = v(i, j, k); }
inline void F_assign3(const Arr3<double>& dst, double s) {
  for (int k = 1; k <= dst.n3; ++k) for (int j = 1; j <= dst.n2; ++j) for (int i = 1; i <= dst.n1; ++i) dst(i, j, k) = s; }
template <class T> struct Arr4 {
  T* p = nullptr; int n1 = 0, n2 = 0, n3 = 0, n4 = 0;
  std::shared_ptr<std::vector<T>> own;  // set by ALLOCATE; empty when the harness lends its storage
  Arr4() {}
  Arr4(int a, int b, int c, int d) : n1(a), n2(b), n3(c), n4(d), own(new std::vector<T>((size_t)a * b * c * d)) { p = own->data(); }
  T& operator()(int i, int j, int k, int l) const { return p[(size_t)(i - 1) + (size_t)n1 * ((j - 1) + (size_t)n2 * ((k - 1) + (size_t)n3 * (l - 1)))]; }
  Arr3<T> slice1(int i) const { Arr3<T> r; r.p = p + (i - 1); r.s1 = n1; r.s2 = (long)n1 * n2; r.s3 = (long)n1 * n2 * n3; r.n1 = n2; r.n2 = n3; r.n3 = n4; return r; }
};

// ---- element-wise arithmetic
template <class A, class B> using common_t = typename std::common_type<A, B>::type;
#define F90_BINOP(op)                                                                                                      \
  template <class A, class B> Vec<common_t<A, B>> operator op(const Vec<A>& a, const Vec<B>& b) {                           \
    Vec<common_t<A, B>> r(a.size()); for (int i = 1; i <= a.size(); ++i) r(i) = (common_t<A, B>)a(i) op (common_t<A, B>)b(i); return r; } \
  template <class A, class B, class = typename std::enable_if<!is_vec<B>::value>::type> Vec<common_t<A, B>> operator op(const Vec<A>& a, const B& b) { \
    Vec<common_t<A, B>> r(a.size()); for (int i = 1; i <= a.size(); ++i) r(i) = (common_t<A, B>)a(i) op (common_t<A, B>)b; return r; }    \
  template <class A, class B, class = typename std::enable_if<!is_vec<A>::value>::type> Vec<common_t<A, B>> operator op(const A& a, const Vec<B>& b) { \
    Vec<common_t<A, B>> r(b.size()); for (int i = 1; i <= b.size(); ++i) r(i) = (common_t<A, B>)a op (common_t<A, B>)b(i); return r; }
F90_BINOP(+) F90_BINOP(-) F90_BINOP(*) F90_BINOP(/)
template <class A> Vec<A> operator-(const Vec<A>& a) { Vec<A> r(a.size()); for (int i = 1; i <= a.size(); ++i) r(i) = -a(i); return r; }

// REAL op COMPLEX: Fortran converts the real operand to COMPLEX (x, +0.0) and performs the COMPLEX operation in full --
// gcc's complex lowering does not shortcut an operand whose imaginary part is the constant 0.0 while signed zeros are
// honoured (tree-complex.c, some_nonzerop) -- so the signs of zero imaginary parts, and with them the branch of a later
// SQRT, are those of the full operation (libstdc++'s mixed operators scale or negate instead).
inline cd operator+(double a, const cd& z) { return cd(a, 0.0) + z; }
inline cd operator+(const cd& z, double a) { return z + cd(a, 0.0); }
inline cd operator-(double a, const cd& z) { return cd(a, 0.0) - z; }
inline cd operator-(const cd& z, double a) { return z - cd(a, 0.0); }
inline cd operator*(double a, const cd& z) { return cd(a, 0.0) * z; }
inline cd operator*(const cd& z, double a) { return z * cd(a, 0.0); }
inline cd operator/(double a, const cd& z) { return cd(a, 0.0) / z; }
inline cd operator/(const cd& z, double a) { return z / cd(a, 0.0); }
inline cd operator*(const cd& a, int b) { return a * cd((double)b, 0.0); }
inline cd operator*(int a, const cd& b) { return cd((double)a, 0.0) * b; }

// ---- intrinsics
namespace F90 {
inline double sqrt(double x) { return ::sqrt(x); }
inline cd sqrt(const cd& x) { return std::sqrt(x); }
inline double abs(double x) { return ::fabs(x); }
inline double cos(double x) { return ::cos(x); }
inline double sin(double x) { return ::sin(x); }
inline double acos(double x) { return ::acos(x); }
inline double atan2(double y, double x) { return ::atan2(y, x); }
inline double max(double a, double b) { return (b > a || a != a) ? b : a; }  // MAX(a, b) as gfortran <= 8 expands it (a decision, not a fact)
inline double real(const cd& z) { return z.real(); }
inline double real(double x) { return x; }
inline double real(float x) { return (double)x; }
// REAL(x) without KIND: default REAL of a real argument, the real part (same kind) of a COMPLEX one
inline float real_default(double x) { return (float)x; }
inline float real_default(float x) { return x; }
inline double real_default(const cd& z) { return z.real(); }
inline int mod(int a, int p) { return a % p; }
inline double imag(const cd& z) { return z.imag(); }
inline cd cmplx_dp(double a, double b) { return cd(a, b); }
inline cd cmplx_sp(double a, double b) { return cd((double)(float)a, (double)(float)b); }  // CMPLX without KIND is default (single) complex
template <class T> Vec<T> sqrt(const Vec<T>& v) { Vec<T> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = F90::sqrt(v(i)); return r; }
inline Vec<double> abs(const Vec<double>& v) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = ::fabs(v(i)); return r; }
inline Vec<double> real(const Vec<cd>& v) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = v(i).real(); return r; }
inline Vec<double> real(const Vec<double>& v) { return v; }
inline Vec<double> real_default(const Vec<cd>& v) { return real(v); }
inline Vec<float> real_default(const Vec<double>& v) { Vec<float> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = (float)v(i); return r; }
inline Vec<double> imag(const Vec<cd>& v) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = v(i).imag(); return r; }
template <class T> T sum(const Vec<T>& v) { T s = T(); for (int i = 1; i <= v.size(); ++i) s = s + v(i); return s; }
inline double dot_product(const Vec<double>& a, const Vec<double>& b) { double s = 0.0; for (int i = 1; i <= a.size(); ++i) s = s + a(i) * b(i); return s; }
inline cd dot_product(const Vec<cd>& a, const Vec<cd>& b) { cd s(0.0, 0.0); for (int i = 1; i <= a.size(); ++i) s = s + std::conj(a(i)) * b(i); return s; }
inline double maxval(const Vec<double>& v) { double m = v(1); for (int i = 2; i <= v.size(); ++i) if (v(i) > m) m = v(i); return m; }
inline double minval(const Vec<double>& v) { double m = v(1); for (int i = 2; i <= v.size(); ++i) if (v(i) < m) m = v(i); return m; }
inline Vec<double> cos(const Vec<double>& v) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = ::cos(v(i)); return r; }
inline Vec<double> sin(const Vec<double>& v) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = ::sin(v(i)); return r; }
inline Mat<double> reshape(const Vec<double>& v, int a, int b) { Mat<double> m(a, b); for (int i = 0; i < a * b; ++i) m.d[(size_t)i] = v.d[(size_t)i]; return m; }
inline Mat<double> transpose(const Mat<double>& m) { Mat<double> t(m.n2, m.n1); for (int i = 1; i <= m.n1; ++i) for (int j = 1; j <= m.n2; ++j) t(j, i) = m(i, j); return t; }
// MATMUL(matrix, vector) as gfortran inlines it: c = 0; do j; do i; c(i) = c(i) + a(i,j) * b(j)
inline Vec<double> matmul(const Mat<double>& a, const Vec<double>& b) {
  Vec<double> c(a.n1);
  for (int j = 1; j <= a.n2; ++j) for (int i = 1; i <= a.n1; ++i) c(i) = c(i) + a(i, j) * b(j);
  return c;
}
inline double exp(double x) { return ::exp(x); }
inline Vec<double> exp(const Vec<double>& v) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = ::exp(v(i)); return r; }
inline Vec<int> iota(int lo, int hi) { Vec<int> r(hi - lo + 1 > 0 ? hi - lo + 1 : 0); for (int i = lo; i <= hi; ++i) r(i - lo + 1) = i; return r; }
template <class T> int size(const Arr3<T>& a, int dim) { return dim == 1 ? a.n1 : (dim == 2 ? a.n2 : a.n3); }
// MAXLOC((/ (ind, ind = 1, n) /), MASK = 0.0 <= (xi - x)): the largest index whose node is not above xi; 0 if there is none
inline int maxloc_iota_mask(int n, double xi, const Vec<double>& x) { int best = 0; for (int i = 1; i <= n; ++i) if (0.0 <= (xi - x(i))) best = i; return best; }
template <class T> int size(const Vec<T>& v) { return v.size(); }
template <class T> int size(const Mat<T>& m, int dim) { return dim == 1 ? m.n1 : m.n2; }
template <class T> int size(const Vec<T>& v, int dim) { (void)dim; return v.size(); }
template <class T> bool allocated(const Vec<T>& v) { return v.allocated; }
}  // namespace F90

// array constructor (/ a, b, c /) of scalars and arrays
inline void F_cat_into(Vec<double>&) {}
template <class... R> void F_cat_into(Vec<double>& out, double a, const R&... r);
template <class... R> void F_cat_into(Vec<double>& out, const Vec<double>& a, const R&... r) { for (int i = 1; i <= a.size(); ++i) out.d.push_back(a(i)); F_cat_into(out, r...); }
template <class... R> void F_cat_into(Vec<double>& out, double a, const R&... r) { out.d.push_back(a); F_cat_into(out, r...); }
template <class... A> Vec<double> F_cat(const A&... a) { Vec<double> out; out.allocated = true; F_cat_into(out, a...); return out; }

template <int N, class T> T F_powc(const T& x) {
  constexpr int tbl[9] = {0, 1, 1, 2, 2, 3, 3, 4, 4};
  static_assert(N >= 1 && N <= 8, "exponent");
  if constexpr (N == 1) return x;
  else return F_powc<N - tbl[N]>(x) * F_powc<tbl[N]>(x);
}
inline double F_pow(double x, double y) { return ::pow(x, y); }
// x**n, integer n at run time: libgfortran's pow_r8_i4 / libgcc's __powidf2 (square and multiply from the low bit)
inline double F_pow(double x, int m) {
  unsigned n = m < 0 ? -(unsigned)m : (unsigned)m;
  double y = (n % 2) ? x : 1;
  while (n >>= 1) { x = x * x; if (n % 2) y *= x; }
  return m < 0 ? 1 / y : y;
}
inline Vec<double> F_pow(const Vec<double>& v, double y) { Vec<double> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = ::pow(v(i), y); return r; }
inline double F_toreal(double x) { return x; }
inline double F_toreal(const cd& z) { return z.real(); }  // COMPLEX assigned to REAL: the real part
inline Vec<cd> F_tocplx(const Vec<cd>& v) { return v; }
inline Vec<cd> F_tocplx(const Vec<double>& v) { Vec<cd> r(v.size()); for (int i = 1; i <= v.size(); ++i) r(i) = cd(v(i), 0.0); return r; }
// ---- list-directed input (OPEN / READ(unit,*) / CLOSE of the interp adapter's setup): values are separated by blanks, commas
// or record ends; a READ takes as many records as its list needs and leaves the file at the start of the next record
inline std::map<int, FILE*>& F_units() { static std::map<int, FILE*> u; return u; }
inline void F_open(int unit, const char* name) {
  FILE* f = fopen(name, "r");
  if (!f) { fprintf(stderr, "f90rt: cannot open %s\n", name); abort(); }
  F_units()[unit] = f;
}
inline void F_close(int unit) { fclose(F_units()[unit]); F_units().erase(unit); }
struct F_Read {
  FILE* f;
  explicit F_Read(int unit) : f(F_units()[unit]) {}
  ~F_Read() { int c; while ((c = fgetc(f)) != EOF && c != '\n') {} }
  std::string token() {
    int c;
    while ((c = fgetc(f)) != EOF && (c == ' ' || c == '\t' || c == ',' || c == '\n' || c == '\r')) {}
    if (c == EOF) { fprintf(stderr, "f90rt: end of file in a READ\n"); abort(); }
    std::string t;
    while (c != EOF && c != ' ' && c != '\t' && c != ',' && c != '\n' && c != '\r') { t.push_back(c == 'D' || c == 'd' ? 'E' : (char)c); c = fgetc(f); }
    if (c != EOF) ungetc(c, f);
    return t;
  }
  void item(int& v) { v = (int)strtol(token().c_str(), nullptr, 10); }
  void item(double& v) { v = strtod(token().c_str(), nullptr); }  // decimal -> nearest double, as libgfortran's read_real does
  void item(const Arr4<double>& a) {  // a whole array: its elements in array-element order
    size_t n = (size_t)a.n1 * a.n2 * a.n3 * a.n4;
    for (size_t i = 0; i < n; ++i) item(a.p[i]);
  }
};
template <class T> T& F_tmp(T&& v) { return v; }
template <class T> T& F_tmp(T& v) { return v; }
template <class T> void F_move_alloc(T& from, T& to) { to = from; to.allocated = true; from = T(); }
