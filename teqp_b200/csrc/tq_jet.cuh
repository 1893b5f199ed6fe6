// Register-resident truncated-Taylor arithmetic ("jets") for the sm_100a kernels.
//
// This replaces the reference's forward-mode AD number types (autodiff::Real<N>, nested autodiff::Dual, multicomplex; call
// sites /root/reference/include/teqp/derivs.hpp:156-162, 209-219, 247-259, 543-569) with one template:
//
//   Tay<N,T>  = sum_{k<=N} c[k] s^k,  c[k] of type T, truncated at s^N.
//
// T is either double or another Tay<M,U>: nesting gives a multivariate jet with RECTANGULAR truncation (degree <= N in the
// outer variable, <= M in the next, ...), so ONE sweep through a model yields every mixed derivative d^(i+j+..) up to
// (N, M, ..) -- where the reference re-runs the whole model once per requested derivative with 2^(i+j+..)-wide nested duals.
//
// Every jet in a kernel has the SAME nesting depth (level 0 = 1/T, level 1 = rho, level 2 = a composition direction, see
// J3<> below); a quantity that does not depend on a level's variable carries degree 0 there (one coefficient, no cost).
// Binary operators accept operands whose degrees at each level are equal or 0 and return the level-wise maximum, so
// temperature-only sub-expressions (alpha functions, Barker-Henderson diameters, ...) stay narrow until they meet a density.
//
// Variables are seeded SCALED: rho(s) = rho0 (1+s), (1/T)(s) = (1/T0)(1+s), so c[k] = rho0^k f^(k)/k! and the reference's
// Lambda_nm = (1/T)^n rho^m d^(n+m) alphar / d(1/T)^n drho^m  (derivs.hpp:468-478) is n! m! c[n][m] with no extra powers.
//
// All loops have compile-time bounds and are fully unrolled: the coefficient arrays live in registers.
#pragma once
#include <cmath>
#include <type_traits>

#if defined(__CUDACC__)
#define TQ_HD __host__ __device__ __forceinline__
#else
#define TQ_HD inline
#endif

namespace tq {

template<int N, class T = double> struct Tay;

template<class T> struct is_tay { static constexpr bool value = false; };
template<int N, class T> struct is_tay<Tay<N, T>> { static constexpr bool value = true; };

// level-wise maximum of two jet types
template<class A, class B> struct Prom;
template<> struct Prom<double, double> { using type = double; };
template<int N, class T> struct Prom<Tay<N, T>, double> { using type = Tay<N, T>; };
template<int N, class T> struct Prom<double, Tay<N, T>> { using type = Tay<N, T>; };
template<int N1, class T1, int N2, class T2> struct Prom<Tay<N1, T1>, Tay<N2, T2>> {
    static_assert(N1 == N2 || N1 == 0 || N2 == 0, "jet degrees at one level must be equal or zero");
    using type = Tay<(N1 > N2 ? N1 : N2), typename Prom<T1, T2>::type>;
};
template<class A, class B> using prom_t = typename Prom<A, B>::type;

// ---- scalar math on the innermost type --------------------------------------------------------------------------------
TQ_HD double base(double x) { return x; }
TQ_HD double texp(double x) { return ::exp(x); }
TQ_HD double tlog(double x) { return ::log(x); }
TQ_HD double tsqrt(double x) { return ::sqrt(x); }
TQ_HD double tpow(double x, double e) { return ::pow(x, e); }
TQ_HD double trecip(double x) { return 1.0 / x; }

template<int N, class T>
struct Tay {
    static constexpr int order = N;
    using coeff_type = T;
    T c[N + 1];
    TQ_HD Tay() {}
    TQ_HD Tay(double v) {
        c[0] = T(v);
#pragma unroll
        for (int k = 1; k <= N; ++k) c[k] = T(0.0);
    }
    // variable seeded so that s is the RELATIVE perturbation: x(s) = x0 (1 + s)
    TQ_HD static Tay scaled_variable(double x0) {
        Tay r(x0);
        if (N >= 1) r.c[N >= 1 ? 1 : 0] = T(x0);
        return r;
    }
    // T(s) when (1/T)(s) = (1/T0)(1+s):  T0/(1+s) = T0 sum_k (-s)^k   (exact alternating series)
    TQ_HD static Tay reciprocal_scaled_variable(double x0) {
        Tay r;
        double v = x0;
#pragma unroll
        for (int k = 0; k <= N; ++k) { r.c[k] = T(v); v = -v; }
        return r;
    }
    // unit-seeded variable: x(s) = x0 + s  (composition directions: derivatives are NOT scaled by x_i, derivs.hpp:247-259)
    TQ_HD static Tay unit_variable(double x0) {
        Tay r(x0);
        if (N >= 1) r.c[N >= 1 ? 1 : 0] = T(1.0);
        return r;
    }
    TQ_HD static Tay lift(const T& v) {
        Tay r;
        r.c[0] = v;
#pragma unroll
        for (int k = 1; k <= N; ++k) r.c[k] = T(0.0);
        return r;
    }
};

template<int N, class T> TQ_HD double base(const Tay<N, T>& x) { return base(x.c[0]); }

// embed a narrower jet into a wider type R (structural zeros for the missing coefficients)
template<class R> TQ_HD R embed(double v) { return R(v); }
template<class R, int N, class T> TQ_HD R embed(const Tay<N, T>& a) {
    static_assert(is_tay<R>::value, "cannot embed a jet into double");
    static_assert(R::order == N || N == 0, "embed: incompatible degree");
    R r;
#pragma unroll
    for (int k = 0; k <= R::order; ++k) {
        if (k <= N) r.c[k] = embed<typename R::coeff_type>(a.c[k <= N ? k : 0]);
        else r.c[k] = typename R::coeff_type(0.0);
    }
    return r;
}

// ---- unary minus ----------------------------------------------------------------------------------------------------------
template<int N, class T> TQ_HD Tay<N, T> operator-(const Tay<N, T>& a) {
    Tay<N, T> r;
#pragma unroll
    for (int k = 0; k <= N; ++k) r.c[k] = -a.c[k];
    return r;
}

// ---- + and - (mixed degrees allowed) ------------------------------------------------------------------------------------
template<int N1, class T1, int N2, class T2>
TQ_HD prom_t<Tay<N1, T1>, Tay<N2, T2>> operator+(const Tay<N1, T1>& a, const Tay<N2, T2>& b) {
    using R = prom_t<Tay<N1, T1>, Tay<N2, T2>>;
    using RC = typename R::coeff_type;
    R r;
    r.c[0] = a.c[0] + b.c[0];
#pragma unroll
    for (int k = 1; k <= R::order; ++k) {
        if (N1 == 0) r.c[k] = embed<RC>(b.c[k <= N2 ? k : 0]);
        else if (N2 == 0) r.c[k] = embed<RC>(a.c[k <= N1 ? k : 0]);
        else r.c[k] = a.c[k <= N1 ? k : 0] + b.c[k <= N2 ? k : 0];
    }
    return r;
}
template<int N1, class T1, int N2, class T2>
TQ_HD prom_t<Tay<N1, T1>, Tay<N2, T2>> operator-(const Tay<N1, T1>& a, const Tay<N2, T2>& b) {
    using R = prom_t<Tay<N1, T1>, Tay<N2, T2>>;
    using RC = typename R::coeff_type;
    R r;
    r.c[0] = a.c[0] - b.c[0];
#pragma unroll
    for (int k = 1; k <= R::order; ++k) {
        if (N1 == 0) r.c[k] = embed<RC>(-b.c[k <= N2 ? k : 0]);
        else if (N2 == 0) r.c[k] = embed<RC>(a.c[k <= N1 ? k : 0]);
        else r.c[k] = a.c[k <= N1 ? k : 0] - b.c[k <= N2 ? k : 0];
    }
    return r;
}
template<int N, class T> TQ_HD Tay<N, T> operator+(const Tay<N, T>& a, double b) { Tay<N, T> r = a; r.c[0] = r.c[0] + b; return r; }
template<int N, class T> TQ_HD Tay<N, T> operator+(double b, const Tay<N, T>& a) { Tay<N, T> r = a; r.c[0] = r.c[0] + b; return r; }
template<int N, class T> TQ_HD Tay<N, T> operator-(const Tay<N, T>& a, double b) { Tay<N, T> r = a; r.c[0] = r.c[0] - b; return r; }
template<int N, class T> TQ_HD Tay<N, T> operator-(double b, const Tay<N, T>& a) { Tay<N, T> r = -a; r.c[0] = r.c[0] + b; return r; }

// ---- * --------------------------------------------------------------------------------------------------------------------
template<int N, class T> TQ_HD Tay<N, T> operator*(const Tay<N, T>& a, double b) {
    Tay<N, T> r;
#pragma unroll
    for (int k = 0; k <= N; ++k) r.c[k] = a.c[k] * b;
    return r;
}
template<int N, class T> TQ_HD Tay<N, T> operator*(double b, const Tay<N, T>& a) { return a * b; }
template<int N, class T> TQ_HD Tay<N, T> operator/(const Tay<N, T>& a, double b) {
    Tay<N, T> r;
#pragma unroll
    for (int k = 0; k <= N; ++k) r.c[k] = a.c[k] / b;
    return r;
}
template<int N1, class T1, int N2, class T2>
TQ_HD prom_t<Tay<N1, T1>, Tay<N2, T2>> operator*(const Tay<N1, T1>& a, const Tay<N2, T2>& b) {
    using R = prom_t<Tay<N1, T1>, Tay<N2, T2>>;
    R r;
    if (N1 == 0) {
#pragma unroll
        for (int k = 0; k <= N2; ++k) r.c[k] = a.c[0] * b.c[k];
    } else if (N2 == 0) {
#pragma unroll
        for (int k = 0; k <= N1; ++k) r.c[k] = a.c[k <= N1 ? k : 0] * b.c[0];
    } else {
#pragma unroll
        for (int k = 0; k <= R::order; ++k) {
            auto s = a.c[0] * b.c[k <= N2 ? k : 0];
#pragma unroll
            for (int j = 1; j <= k; ++j) s = s + a.c[j <= N1 ? j : 0] * b.c[(k - j) <= N2 ? (k - j) : 0];
            r.c[k] = s;
        }
    }
    return r;
}

// ---- reciprocal and division ---------------------------------------------------------------------------------------------
template<int N, class T> TQ_HD Tay<N, T> trecip(const Tay<N, T>& b) {
    Tay<N, T> r;
    const T inv = trecip(b.c[0]);
    r.c[0] = inv;
#pragma unroll
    for (int k = 1; k <= N; ++k) {
        T s = b.c[1 <= N ? 1 : 0] * r.c[k - 1];
#pragma unroll
        for (int j = 2; j <= k; ++j) s = s + b.c[j] * r.c[k - j];
        r.c[k] = -(s * inv);
    }
    return r;
}
template<int N, class T> TQ_HD Tay<N, T> operator/(const Tay<N, T>& a, const Tay<N, T>& b) {
    Tay<N, T> r;
    r.c[0] = a.c[0] / b.c[0];  // true division for the base value (bit-faithful to the reference's a/b); reciprocal for the rest
    if (N == 0) return r;
    const T inv = trecip(b.c[0]);
#pragma unroll
    for (int k = 1; k <= N; ++k) {
        T s = a.c[k];
#pragma unroll
        for (int j = 1; j <= k; ++j) s = s - b.c[j] * r.c[k - j];
        r.c[k] = s * inv;
    }
    return r;
}
template<int N1, class T1, int N2, class T2>
TQ_HD typename std::enable_if<!std::is_same<Tay<N1, T1>, Tay<N2, T2>>::value, prom_t<Tay<N1, T1>, Tay<N2, T2>>>::type
operator/(const Tay<N1, T1>& a, const Tay<N2, T2>& b) {
    return a * trecip(b);
}
template<int N, class T> TQ_HD Tay<N, T> operator/(double a, const Tay<N, T>& b) { return trecip(b) * a; }

// ---- elementary functions (standard Taylor recurrences) -------------------------------------------------------------------
template<int N, class T> TQ_HD Tay<N, T> texp(const Tay<N, T>& a) {
    Tay<N, T> r;
    r.c[0] = texp(a.c[0]);
#pragma unroll
    for (int k = 1; k <= N; ++k) {
        T s = (a.c[1 <= N ? 1 : 0] * r.c[k - 1]);
#pragma unroll
        for (int j = 2; j <= k; ++j) s = s + (a.c[j] * r.c[k - j]) * double(j);
        r.c[k] = s * (1.0 / double(k));
    }
    return r;
}
template<int N, class T> TQ_HD Tay<N, T> tlog(const Tay<N, T>& a) {
    Tay<N, T> r;
    r.c[0] = tlog(a.c[0]);
    if (N >= 1) {
        const T inv = trecip(a.c[0]);
#pragma unroll
        for (int k = 1; k <= N; ++k) {
            T s = a.c[k];
#pragma unroll
            for (int j = 1; j < k; ++j) s = s - (r.c[j] * a.c[k - j]) * (double(j) / double(k));
            r.c[k] = s * inv;
        }
    }
    return r;
}
template<int N, class T> TQ_HD Tay<N, T> tsqrt(const Tay<N, T>& a) {
    Tay<N, T> r;
    r.c[0] = tsqrt(a.c[0]);
    if (N >= 1) {
        const T inv = trecip(r.c[0] + r.c[0]);
#pragma unroll
        for (int k = 1; k <= N; ++k) {
            T s = a.c[k];
#pragma unroll
            for (int j = 1; j < k; ++j) s = s - r.c[j] * r.c[k - j];
            r.c[k] = s * inv;
        }
    }
    return r;
}
// a^e for real e (a's base value > 0)
template<int N, class T> TQ_HD Tay<N, T> tpow(const Tay<N, T>& a, double e) {
    Tay<N, T> r;
    r.c[0] = tpow(a.c[0], e);
    if (N >= 1) {
        const T inv = trecip(a.c[0]);
#pragma unroll
        for (int k = 1; k <= N; ++k) {
            T s = (a.c[1 <= N ? 1 : 0] * r.c[k - 1]) * (e - double(k - 1));
#pragma unroll
            for (int j = 2; j <= k; ++j) s = s + (a.c[j] * r.c[k - j]) * (e * double(j) - double(k - j));
            r.c[k] = (s * inv) * (1.0 / double(k));
        }
    }
    return r;
}
template<class T> TQ_HD T tsqr(const T& a) { return a * a; }
template<class T> TQ_HD T tcube(const T& a) { return a * a * a; }
// integer power by squaring, mirroring the special cases of teqp::powi (/root/reference/include/teqp/types.hpp:138-170)
template<class T> TQ_HD T tpowi(const T& x, int n) {
    switch (n) {
        case 0: return T(1.0);
        case 1: return x;
        case 2: return x * x;
        case 3: return x * x * x;
        case 4: { T y = x * x; return y * y; }
    }
    if (n < 0) return trecip(tpowi(x, -n));
    T y = x, r(1.0);
    int m = n;
    while (m > 0) {
        if (m & 1) r = r * y;
        m >>= 1;
        if (m) y = y * y;
    }
    return r;
}

// The three-level jet every generic model kernel uses: degree A in 1/T, B in rho, C in one composition direction.
template<int A, int B, int C> using J3 = Tay<A, Tay<B, Tay<C, double>>>;
template<int A, int B, int C> TQ_HD double coef(const J3<A, B, C>& x, int a, int b, int c) { return x.c[a].c[b].c[c]; }

// Five levels: 1/T, rho and THREE composition directions -- the cross-composition derivatives get_ATrhoXiXj / get_ATrhoXiXjXk
// (/root/reference/include/teqp/derivs.hpp:294-339, 423-465).  Every jet of one evaluation has this depth.
template<int A, int B, int C, int D, int E> using J5 = Tay<A, Tay<B, Tay<C, Tay<D, Tay<E, double>>>>>;

TQ_HD double tfact(int k) {
    double f = 1.0;
    for (int i = 2; i <= k; ++i) f *= double(i);
    return f;
}

}  // namespace tq
