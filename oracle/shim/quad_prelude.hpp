// TEST INFRASTRUCTURE ONLY -- force-included ahead of oracle/ref_truth.cpp (and nothing else).
//
// Makes GCC's __float128 (libquadmath, 113-bit significand) usable as the `Scalar` of the reference's templates
// (/root/reference/include/teqp/derivs.hpp:145 `template<typename Model, typename Scalar, typename VectorType> struct TDXDerivatives`)
// and as the coefficient type of the AD shim's Real<N,T> / Dual<T,G>.  The reference's models call exp/log/pow/sqrt
// unqualified after `using std::exp` etc.; for a built-in type there is no ADL, so the quad overloads have to be visible as
// std::exp(__float128), ... .  Adding overloads to namespace std is formally not allowed; it is confined to this one test TU and
// exact-match overloads cannot change the meaning of any other call.  Compile with -std=gnu++20 -fext-numeric-literals.
#pragma once
#include <quadmath.h>

#include <cmath>
#include <complex>
#include <iostream>
#include <limits>

namespace std {
// libstdc++ 13 has no numeric_limits<__float128>.  The reference uses epsilon() of the BASE type as an algorithmic constant: the cutoff of
// the Barker-Henderson integral is where exp(-u/T) = EPS (saftvrmie.hpp:299-301).  The truth leg must evaluate the SAME formula the double
// reference evaluates (same cutoff, same 10-point Gauss-Legendre interval), only in wider arithmetic, so epsilon() is the DOUBLE epsilon here;
// with the binary128 epsilon the quadrature interval changes and the "truth" would be that of another formula (it moves d_ii by ~1e-7).
template<> struct numeric_limits<__float128> {
    static constexpr bool is_specialized = true, is_signed = true, is_integer = false, is_exact = false, has_infinity = true, has_quiet_NaN = true;
    static constexpr int digits = 113, digits10 = 33, max_digits10 = 36, radix = 2;
    static constexpr __float128 epsilon() noexcept { return 2.220446049250313080847263336181640625e-16Q; }
    static constexpr __float128 min() noexcept { return 3.36210314311209350626267781732175260e-4932Q; }
    static constexpr __float128 max() noexcept { return 1.18973149535723176508575932662800702e4932Q; }
    static constexpr __float128 lowest() noexcept { return -1.18973149535723176508575932662800702e4932Q; }
    static __float128 infinity() noexcept { return __builtin_huge_valq(); }
    static __float128 quiet_NaN() noexcept { return __builtin_nanq(""); }
};
inline __float128 exp(__float128 x) { return expq(x); }
inline __float128 log(__float128 x) { return logq(x); }
inline __float128 log10(__float128 x) { return log10q(x); }
inline __float128 sqrt(__float128 x) { return sqrtq(x); }
inline __float128 cbrt(__float128 x) { return cbrtq(x); }
inline __float128 pow(__float128 x, __float128 y) { return powq(x, y); }
inline __float128 pow(__float128 x, double y) { return powq(x, (__float128)y); }
inline __float128 pow(__float128 x, int y) { return powq(x, (__float128)y); }
inline __float128 pow(double x, __float128 y) { return powq((__float128)x, y); }
inline __float128 fabs(__float128 x) { return fabsq(x); }
inline __float128 sinh(__float128 x) { return sinhq(x); }
inline __float128 cosh(__float128 x) { return coshq(x); }
inline __float128 tanh(__float128 x) { return tanhq(x); }
inline __float128 sin(__float128 x) { return sinq(x); }
inline __float128 cos(__float128 x) { return cosq(x); }
inline __float128 atan(__float128 x) { return atanq(x); }
inline __float128 floor(__float128 x) { return floorq(x); }
inline __float128 ceil(__float128 x) { return ceilq(x); }
inline bool isfinite(__float128 x) { return finiteq(x) != 0; }
inline bool isnan(__float128 x) { return isnanq(x) != 0; }
inline bool isinf(__float128 x) { return isinfq(x) != 0; }
inline ostream& operator<<(ostream& os, __float128 x) { return os << (double)x; }
}  // namespace std

// the same prelude the other oracle TUs get (ref_prelude.hpp)
#include "boost/multiprecision/cpp_bin_float.hpp"
#include "teqp/finite_derivs.hpp"
