// TEST INFRASTRUCTURE ONLY -- part of the parity oracle, never linked into the product library.
//
// Clean-room restatement of the public semantics of the forward-mode AD library that the reference
// depends on but does not ship:  autodiff/autodiff pinned at cb2d6e8b9fa98b7b8cf3ffaefa231943c56361ae
// (/root/reference/.SUBMODULES.json:44-50, status "missing").  Only the API subset that the reference's
// hot path touches is provided; the call sites that define that subset are
//   include/teqp/types.hpp:51-64,89-121        (detail::isReal/isDual/isExpr, detail::val, detail::eval)
//   include/teqp/derivs.hpp:159-161,548-553    (Real<N,T>, derivatives(f, along(1), at(x)))
//   include/teqp/derivs.hpp:211-218,249-258    (HigherOrderDual<N,double>, Wrt<>, wrt(...), at(...))
//   include/teqp/derivs.hpp:1035-1064,1116-1122 (dual2nd, ArrayXdual(2nd), gradient(), hessian())
//   include/teqp/derivs.hpp:1276,1292          (structured binding of derivatives(f, wrt(a,b), at(a,b)))
//
// Semantics restated (from the library's published documentation):
//  * Real<N,T>: value and the N first directional derivatives along one direction.  derivatives() returns
//    d^k f / ds^k for k = 0..N (derivatives, NOT Taylor coefficients).  Internally this shim stores
//    normalised Taylor coefficients c_k = f^(k)/k! and converts on output.
//  * Dual<T,G>: first-order dual number {val, grad}; HigherOrderDual<N,T> nests it N times.
//    wrt(v1,...,vm) seeds the k-th listed variable at nesting level k (k = 1 is the outermost grad);
//    derivatives() returns {f, d f/dv1, d2 f/dv1 dv2, ..., d^m f/dv1..dvm}.
//  * No expression templates: eval() is the identity, isExpr is always false.
//
// Nothing here is copied from the autodiff sources (they are not available in this container).
#pragma once

#include <array>
#include <cmath>
#include <cstddef>
#include <tuple>
#include <type_traits>
#include <utility>
#include <ostream>

namespace autodiff {
namespace detail {

template<typename U> constexpr bool isArithmetic = std::is_arithmetic_v<std::decay_t<U>>;

// ------------------------------------------------------------------------------------------------
// Real<N,T>: truncated Taylor series in one direction, c[k] = f^(k)(s)/k!
// ------------------------------------------------------------------------------------------------
template<size_t N, typename T>
struct Real {
    std::array<T, N + 1> c{};
    constexpr Real() = default;
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0>
    constexpr Real(const U& v) { c[0] = static_cast<T>(v); }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0>
    Real& operator=(const U& v) { c.fill(T(0)); c[0] = static_cast<T>(v); return *this; }
    explicit operator T() const { return c[0]; }
    T& operator[](size_t i) { return c[i]; }
    const T& operator[](size_t i) const { return c[i]; }
    Real& operator+=(const Real& o) { for (size_t k = 0; k <= N; ++k) c[k] += o.c[k]; return *this; }
    Real& operator-=(const Real& o) { for (size_t k = 0; k <= N; ++k) c[k] -= o.c[k]; return *this; }
    Real& operator*=(const Real& o) { *this = (*this) * o; return *this; }
    Real& operator/=(const Real& o) { *this = (*this) / o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real& operator+=(const U& o) { c[0] += o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real& operator-=(const U& o) { c[0] -= o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real& operator*=(const U& o) { for (auto& v : c) v *= o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real& operator/=(const U& o) { for (auto& v : c) v /= o; return *this; }
};

template<size_t N, typename T> Real<N,T> operator+(const Real<N,T>& a) { return a; }
template<size_t N, typename T> Real<N,T> operator-(const Real<N,T>& a) { Real<N,T> r; for (size_t k = 0; k <= N; ++k) r.c[k] = -a.c[k]; return r; }
template<size_t N, typename T> Real<N,T> operator+(const Real<N,T>& a, const Real<N,T>& b) { Real<N,T> r; for (size_t k = 0; k <= N; ++k) r.c[k] = a.c[k] + b.c[k]; return r; }
template<size_t N, typename T> Real<N,T> operator-(const Real<N,T>& a, const Real<N,T>& b) { Real<N,T> r; for (size_t k = 0; k <= N; ++k) r.c[k] = a.c[k] - b.c[k]; return r; }
template<size_t N, typename T> Real<N,T> operator*(const Real<N,T>& a, const Real<N,T>& b) {
    Real<N,T> r;
    for (size_t k = 0; k <= N; ++k) { T s = 0; for (size_t j = 0; j <= k; ++j) s += a.c[j] * b.c[k - j]; r.c[k] = s; }
    return r;
}
template<size_t N, typename T> Real<N,T> operator/(const Real<N,T>& a, const Real<N,T>& b) {
    Real<N,T> r;
    for (size_t k = 0; k <= N; ++k) { T s = a.c[k]; for (size_t j = 1; j <= k; ++j) s -= b.c[j] * r.c[k - j]; r.c[k] = s / b.c[0]; }
    return r;
}
#define SHIM_REAL_SCALAR_OPS(OP) \
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> operator OP(const Real<N,T>& a, const U& b) { return a OP Real<N,T>(b); } \
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> operator OP(const U& a, const Real<N,T>& b) { return Real<N,T>(a) OP b; }
SHIM_REAL_SCALAR_OPS(+)
SHIM_REAL_SCALAR_OPS(-)
#undef SHIM_REAL_SCALAR_OPS
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> operator*(const Real<N,T>& a, const U& b) { Real<N,T> r; for (size_t k = 0; k <= N; ++k) r.c[k] = a.c[k] * b; return r; }
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> operator*(const U& a, const Real<N,T>& b) { Real<N,T> r; for (size_t k = 0; k <= N; ++k) r.c[k] = a * b.c[k]; return r; }
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> operator/(const Real<N,T>& a, const U& b) { Real<N,T> r; for (size_t k = 0; k <= N; ++k) r.c[k] = a.c[k] / b; return r; }
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> operator/(const U& a, const Real<N,T>& b) { return Real<N,T>(a) / b; }

#define SHIM_CMP(TYPEHEAD, TYPE, VAL) \
TYPEHEAD bool operator<(const TYPE& a, const TYPE& b) { return VAL(a) < VAL(b); } \
TYPEHEAD bool operator>(const TYPE& a, const TYPE& b) { return VAL(a) > VAL(b); } \
TYPEHEAD bool operator<=(const TYPE& a, const TYPE& b) { return VAL(a) <= VAL(b); } \
TYPEHEAD bool operator>=(const TYPE& a, const TYPE& b) { return VAL(a) >= VAL(b); } \
TYPEHEAD bool operator==(const TYPE& a, const TYPE& b) { return VAL(a) == VAL(b); } \
TYPEHEAD bool operator!=(const TYPE& a, const TYPE& b) { return VAL(a) != VAL(b); }
#define SHIM_CMP_SCALAR(TYPEHEAD, TYPE, VAL) \
TYPEHEAD bool operator<(const TYPE& a, const U& b) { return VAL(a) < b; }  TYPEHEAD bool operator<(const U& a, const TYPE& b) { return a < VAL(b); } \
TYPEHEAD bool operator>(const TYPE& a, const U& b) { return VAL(a) > b; }  TYPEHEAD bool operator>(const U& a, const TYPE& b) { return a > VAL(b); } \
TYPEHEAD bool operator<=(const TYPE& a, const U& b) { return VAL(a) <= b; }  TYPEHEAD bool operator<=(const U& a, const TYPE& b) { return a <= VAL(b); } \
TYPEHEAD bool operator>=(const TYPE& a, const U& b) { return VAL(a) >= b; }  TYPEHEAD bool operator>=(const U& a, const TYPE& b) { return a >= VAL(b); } \
TYPEHEAD bool operator==(const TYPE& a, const U& b) { return VAL(a) == b; }  TYPEHEAD bool operator==(const U& a, const TYPE& b) { return a == VAL(b); } \
TYPEHEAD bool operator!=(const TYPE& a, const U& b) { return VAL(a) != b; }  TYPEHEAD bool operator!=(const U& a, const TYPE& b) { return a != VAL(b); }

// innermost base value
template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> constexpr U val(const U& x) { return x; }
template<size_t N, typename T> auto val(const Real<N,T>& x) { return val(x.c[0]); }

#define SHIM_HEAD template<size_t N, typename T>
#define SHIM_TYPE Real<N,T>
SHIM_CMP(SHIM_HEAD, SHIM_TYPE, val)
#undef SHIM_HEAD
#define SHIM_HEAD template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0>
SHIM_CMP_SCALAR(SHIM_HEAD, SHIM_TYPE, val)
#undef SHIM_HEAD
#undef SHIM_TYPE

// elementary functions on Taylor series (standard recurrences)
template<size_t N, typename T> Real<N,T> exp(const Real<N,T>& a) {
    using std::exp; Real<N,T> r; r.c[0] = exp(a.c[0]);
    for (size_t k = 1; k <= N; ++k) { T s = 0; for (size_t j = 1; j <= k; ++j) s += T(j) * a.c[j] * r.c[k - j]; r.c[k] = s / T(k); }
    return r;
}
template<size_t N, typename T> Real<N,T> log(const Real<N,T>& a) {
    using std::log; Real<N,T> r; r.c[0] = log(a.c[0]);
    for (size_t k = 1; k <= N; ++k) { T s = 0; for (size_t j = 1; j < k; ++j) s += T(j) * r.c[j] * a.c[k - j]; r.c[k] = (a.c[k] - s / T(k)) / a.c[0]; }
    return r;
}
template<size_t N, typename T> Real<N,T> sqrt(const Real<N,T>& a) {
    using std::sqrt; Real<N,T> r; r.c[0] = sqrt(a.c[0]);
    for (size_t k = 1; k <= N; ++k) { T s = a.c[k]; for (size_t j = 1; j < k; ++j) s -= r.c[j] * r.c[k - j]; r.c[k] = s / (T(2) * r.c[0]); }
    return r;
}
template<size_t N, typename T> Real<N,T> powint(const Real<N,T>& a, long n) {
    if (n < 0) return Real<N,T>(1) / powint(a, -n);
    Real<N,T> r(1), b(a);
    while (n > 0) { if (n & 1) r = r * b; n >>= 1; if (n) b = b * b; }
    return r;
}
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> pow(const Real<N,T>& a, const U& e_) {
    using std::pow;
    const T e = static_cast<T>(e_);
    if (a.c[0] == T(0)) {  // only integer exponents are differentiable at a zero base
        const long n = static_cast<long>(e_);
        if (static_cast<T>(n) == e && n >= 0) return powint(a, n);
    }
    Real<N,T> r; r.c[0] = pow(a.c[0], e_);  // the arithmetic exponent goes down unchanged, so nested Real<N,Real<M,T>> keeps this rule at every level
    for (size_t k = 1; k <= N; ++k) { T s = 0; for (size_t j = 1; j <= k; ++j) s += (e * T(j) - T(k - j)) * a.c[j] * r.c[k - j]; r.c[k] = s / (T(k) * a.c[0]); }
    return r;
}
template<size_t N, typename T> Real<N,T> pow(const Real<N,T>& a, const Real<N,T>& b) { return exp(b * log(a)); }
template<size_t N, typename T, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Real<N,T> pow(const U& a, const Real<N,T>& b) { using std::log; return exp(b * log(static_cast<T>(a))); }
template<size_t N, typename T> Real<N,T> abs(const Real<N,T>& a) { return (val(a) < 0) ? -a : a; }
template<size_t N, typename T> void sincos_series(const Real<N,T>& a, Real<N,T>& s, Real<N,T>& c) {
    using std::sin; using std::cos; s.c[0] = sin(a.c[0]); c.c[0] = cos(a.c[0]);
    for (size_t k = 1; k <= N; ++k) { T ss = 0, cc = 0; for (size_t j = 1; j <= k; ++j) { ss += T(j) * a.c[j] * c.c[k - j]; cc -= T(j) * a.c[j] * s.c[k - j]; } s.c[k] = ss / T(k); c.c[k] = cc / T(k); }
}
template<size_t N, typename T> Real<N,T> sin(const Real<N,T>& a) { Real<N,T> s, c; sincos_series(a, s, c); return s; }
template<size_t N, typename T> Real<N,T> cos(const Real<N,T>& a) { Real<N,T> s, c; sincos_series(a, s, c); return c; }
template<size_t N, typename T> Real<N,T> sinh(const Real<N,T>& a) { return (exp(a) - exp(-a)) / T(2); }
template<size_t N, typename T> Real<N,T> cosh(const Real<N,T>& a) { return (exp(a) + exp(-a)) / T(2); }
template<size_t N, typename T> Real<N,T> tanh(const Real<N,T>& a) { auto p = exp(a), m = exp(-a); return (p - m) / (p + m); }
template<size_t N, typename T> std::ostream& operator<<(std::ostream& os, const Real<N,T>& a) { return os << a.c[0]; }

// ------------------------------------------------------------------------------------------------
// Dual<T,G>: first-order dual number; nest for higher order
// ------------------------------------------------------------------------------------------------
template<typename T, typename G> struct Dual;
template<typename T, typename G> auto val(const Dual<T,G>& x);

template<typename T, typename G>
struct Dual {
    T val{};
    G grad{};
    constexpr Dual() = default;
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0>
    constexpr Dual(const U& v) : val(v), grad(0) {}
    constexpr Dual(const T& v, const G& g) : val(v), grad(g) {}
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0>
    Dual& operator=(const U& v) { val = v; grad = 0; return *this; }
    Dual& operator+=(const Dual& o) { val += o.val; grad += o.grad; return *this; }
    Dual& operator-=(const Dual& o) { val -= o.val; grad -= o.grad; return *this; }
    Dual& operator*=(const Dual& o) { *this = (*this) * o; return *this; }
    Dual& operator/=(const Dual& o) { *this = (*this) / o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual& operator+=(const U& o) { val += o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual& operator-=(const U& o) { val -= o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual& operator*=(const U& o) { val *= o; grad *= o; return *this; }
    template<typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual& operator/=(const U& o) { val /= o; grad /= o; return *this; }
    explicit operator double() const { return static_cast<double>(detail::val(val)); }
};

template<typename T, typename G> auto val(const Dual<T,G>& x) { return detail::val(x.val); }

template<typename T, typename G> Dual<T,G> operator+(const Dual<T,G>& a) { return a; }
template<typename T, typename G> Dual<T,G> operator-(const Dual<T,G>& a) { return Dual<T,G>(-a.val, -a.grad); }
template<typename T, typename G> Dual<T,G> operator+(const Dual<T,G>& a, const Dual<T,G>& b) { return Dual<T,G>(a.val + b.val, a.grad + b.grad); }
template<typename T, typename G> Dual<T,G> operator-(const Dual<T,G>& a, const Dual<T,G>& b) { return Dual<T,G>(a.val - b.val, a.grad - b.grad); }
template<typename T, typename G> Dual<T,G> operator*(const Dual<T,G>& a, const Dual<T,G>& b) { return Dual<T,G>(a.val * b.val, a.grad * b.val + a.val * b.grad); }
template<typename T, typename G> Dual<T,G> operator/(const Dual<T,G>& a, const Dual<T,G>& b) {
    const T q = a.val / b.val;
    return Dual<T,G>(q, (a.grad - q * b.grad) / b.val);
}
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator+(const Dual<T,G>& a, const U& b) { return Dual<T,G>(a.val + b, a.grad); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator+(const U& a, const Dual<T,G>& b) { return Dual<T,G>(a + b.val, b.grad); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator-(const Dual<T,G>& a, const U& b) { return Dual<T,G>(a.val - b, a.grad); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator-(const U& a, const Dual<T,G>& b) { return Dual<T,G>(a - b.val, -b.grad); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator*(const Dual<T,G>& a, const U& b) { return Dual<T,G>(a.val * b, a.grad * b); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator*(const U& a, const Dual<T,G>& b) { return Dual<T,G>(a * b.val, a * b.grad); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator/(const Dual<T,G>& a, const U& b) { return Dual<T,G>(a.val / b, a.grad / b); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> operator/(const U& a, const Dual<T,G>& b) {
    const T q = a / b.val;
    return Dual<T,G>(q, -(q * b.grad) / b.val);
}

#define SHIM_HEAD template<typename T, typename G>
#define SHIM_TYPE Dual<T,G>
SHIM_CMP(SHIM_HEAD, SHIM_TYPE, val)
#undef SHIM_HEAD
#define SHIM_HEAD template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0>
SHIM_CMP_SCALAR(SHIM_HEAD, SHIM_TYPE, val)
#undef SHIM_HEAD
#undef SHIM_TYPE
#undef SHIM_CMP
#undef SHIM_CMP_SCALAR

template<typename T, typename G> Dual<T,G> exp(const Dual<T,G>& a) { using std::exp; const T e = exp(a.val); return Dual<T,G>(e, e * a.grad); }
template<typename T, typename G> Dual<T,G> log(const Dual<T,G>& a) { using std::log; return Dual<T,G>(log(a.val), a.grad / a.val); }
template<typename T, typename G> Dual<T,G> sqrt(const Dual<T,G>& a) { using std::sqrt; const T s = sqrt(a.val); return Dual<T,G>(s, a.grad / (2.0 * s)); }
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> pow(const Dual<T,G>& a, const U& e) {
    using std::pow;
    const double ed = static_cast<double>(e);
    if (ed == 0.0) return Dual<T,G>(1.0);
    return Dual<T,G>(pow(a.val, ed), (ed * pow(a.val, ed - 1.0)) * a.grad);
}
template<typename T, typename G> Dual<T,G> pow(const Dual<T,G>& a, const Dual<T,G>& b) {
    using std::pow; using std::log;
    const T p = pow(a.val, b.val);
    return Dual<T,G>(p, p * (b.grad * log(a.val) + b.val * a.grad / a.val));
}
template<typename T, typename G, typename U, std::enable_if_t<isArithmetic<U>, int> = 0> Dual<T,G> pow(const U& a, const Dual<T,G>& b) {
    using std::pow; using std::log;
    const T p = pow(static_cast<double>(a), b.val);
    return Dual<T,G>(p, p * b.grad * log(static_cast<double>(a)));
}
template<typename T, typename G> Dual<T,G> abs(const Dual<T,G>& a) { return (val(a) < 0) ? -a : a; }
template<typename T, typename G> Dual<T,G> sin(const Dual<T,G>& a) { using std::sin; using std::cos; return Dual<T,G>(sin(a.val), cos(a.val) * a.grad); }
template<typename T, typename G> Dual<T,G> cos(const Dual<T,G>& a) { using std::sin; using std::cos; return Dual<T,G>(cos(a.val), -(sin(a.val) * a.grad)); }
template<typename T, typename G> Dual<T,G> sinh(const Dual<T,G>& a) { using std::sinh; using std::cosh; return Dual<T,G>(sinh(a.val), cosh(a.val) * a.grad); }
template<typename T, typename G> Dual<T,G> cosh(const Dual<T,G>& a) { using std::sinh; using std::cosh; return Dual<T,G>(cosh(a.val), sinh(a.val) * a.grad); }
template<typename T, typename G> Dual<T,G> tanh(const Dual<T,G>& a) { using std::tanh; const T t = tanh(a.val); return Dual<T,G>(t, (1.0 - t * t) * a.grad); }
template<typename T, typename G> std::ostream& operator<<(std::ostream& os, const Dual<T,G>& a) { return os << val(a); }

template<size_t N, typename T> struct AuxHigherOrderDual { using inner = typename AuxHigherOrderDual<N - 1, T>::type; using type = Dual<inner, inner>; };
template<typename T> struct AuxHigherOrderDual<0, T> { using type = T; };
template<size_t N, typename T> using HigherOrderDual = typename AuxHigherOrderDual<N, T>::type;

// type predicates / traits
template<typename U> struct IsRealImpl : std::false_type {};
template<size_t N, typename T> struct IsRealImpl<Real<N,T>> : std::true_type {};
template<typename U> struct IsDualImpl : std::false_type {};
template<typename T, typename G> struct IsDualImpl<Dual<T,G>> : std::true_type {};
template<typename U> constexpr bool isReal = IsRealImpl<std::decay_t<U>>::value;
template<typename U> constexpr bool isDual = IsDualImpl<std::decay_t<U>>::value;
template<typename U> constexpr bool isExpr = false;

template<typename U> struct NumberTraits { using NumericType = U; static constexpr size_t Order = 0; };
template<size_t N, typename T> struct NumberTraits<Real<N,T>> { using NumericType = T; static constexpr size_t Order = N; };
template<typename T, typename G> struct NumberTraits<Dual<T,G>> { using NumericType = typename NumberTraits<T>::NumericType; static constexpr size_t Order = 1 + NumberTraits<T>::Order; };

template<typename U> decltype(auto) eval(U&& x) { return std::forward<U>(x); }

// seeding: level 1 is the outermost grad, level k>1 recurses into val
template<size_t level, typename D, typename S> void seed(D& x, const S& s) {
    static_assert(isDual<D>, "seed needs a dual number");
    if constexpr (level == 1) x.grad = s; else seed<level - 1>(x.val, s);
}
// derivative<order>: order 0 = base value, order k = k-fold descent through grad
template<size_t order, typename D> auto derivative(const D& x) {
    if constexpr (order == 0) return val(x);
    else if constexpr (order == 1) return val(x.grad);
    else return derivative<order - 1>(x.grad);
}

template<typename... Args> struct Wrt { std::tuple<Args...> args; };
template<typename... Args> struct At { std::tuple<Args...> args; };
template<typename... Args> struct Along { std::tuple<Args...> args; };
template<typename... Args> auto wrt(Args&&... args) { return Wrt<Args&&...>{ std::forward_as_tuple(std::forward<Args>(args)...) }; }
template<typename... Args> auto at(Args&&... args) { return At<Args&&...>{ std::forward_as_tuple(std::forward<Args>(args)...) }; }
template<typename... Args> auto along(Args&&... args) { return Along<Args&&...>{ std::forward_as_tuple(std::forward<Args>(args)...) }; }

template<typename Tuple, typename S, size_t... I> void seed_all(Tuple& t, const S& s, std::index_sequence<I...>) {
    (seed<I + 1>(std::get<I>(t), s), ...);
}
template<typename R, typename U, size_t... I> auto collect(const U& u, std::index_sequence<I...>) {
    return std::array<R, sizeof...(I)>{ static_cast<R>(derivative<I>(u))... };
}

// derivatives(f, wrt(v1..vm), at(a1..an)) for scalar dual variables
template<typename Fun, typename... Vars, typename... Args>
auto derivatives(const Fun& f, const Wrt<Vars...>& w, const At<Args...>& a) {
    constexpr size_t M = sizeof...(Vars);
    auto vars = w.args;  // tuple of references
    seed_all(vars, 1.0, std::make_index_sequence<M>{});
    auto u = std::apply(f, a.args);
    seed_all(vars, 0.0, std::make_index_sequence<M>{});
    using R = decltype(val(u));
    return collect<R>(u, std::make_index_sequence<M + 1>{});
}

// derivatives(f, along(v), at(x)) for one Real variable
template<typename Fun, typename V, typename Arg>
auto derivatives(const Fun& f, const Along<V>& al, const At<Arg>& a) {
    auto& x = std::get<0>(a.args);
    using RealT = std::decay_t<decltype(x)>;
    using T = typename NumberTraits<RealT>::NumericType;
    constexpr size_t N = NumberTraits<RealT>::Order;
    x[1] = static_cast<T>(std::get<0>(al.args));
    auto u = f(x);
    x[1] = T(0);
    std::array<T, N + 1> out{};
    T fact = 1;
    for (size_t k = 0; k <= N; ++k) { if (k > 0) fact *= T(k); out[k] = u[k] * fact; }
    return out;
}

} // namespace detail

using detail::Real;
using detail::Dual;
using detail::HigherOrderDual;
using detail::Wrt;
using detail::At;
using detail::Along;
using detail::wrt;
using detail::at;
using detail::along;
using detail::derivatives;
using detail::val;

using real = Real<1, double>;
using real1st = Real<1, double>;
using real2nd = Real<2, double>;
using real3rd = Real<3, double>;
using real4th = Real<4, double>;
using dual0th = double;
using dual = HigherOrderDual<1, double>;
using dual1st = HigherOrderDual<1, double>;
using dual2nd = HigherOrderDual<2, double>;
using dual3rd = HigherOrderDual<3, double>;
using dual4th = HigherOrderDual<4, double>;

} // namespace autodiff
