// Ideal-gas Helmholtz energy alpha^ig(T, rho, x) = sum_i x_i (alpha^ig_i(T, rho) + ln x_i), written once over jet types.
//
// Reference arithmetic: the ten term classes of /root/reference/include/teqp/ideal_eosterms.hpp:14-239, PureIdealHelmholtz::alphaig :305-313,
// IdealHelmholtz::alphaig :337-353 (x_i == 0 skipped), and the GERG-2004 / GERG-2008 ideal part GERG200XAlphaig::alphaig_pure / alphaig
// /root/reference/include/teqp/models/GERG/GERG.hpp:396-422 (x_i <= 0 skipped).  The reference exposes these models through the same
// AbstractModel surface as the residual ones ("alphar" is an alias of alphaig, ideal_eosterms.hpp:355-360, GERG.hpp:1214-1223), so every entry
// point of the hot path -- Lambda tables, composition derivatives, the isochoric family -- serves them too.
//
// Every temperature term is a function of 1/T; the generic kernels seed T as the exact reciprocal series of the scaled variable 1/T, so
// trecip(T) is exact.  ln|sinh y| and ln cosh y are evaluated as |y| - ln 2 + ln(1 -/+ exp(-2|y|)): y is linear in 1/T, so every derivative beyond the
// first comes from the small correction term -- measured against a 60-digit evaluation this is closer to the exact Lambda_30 / Lambda_40 than
// taking the logarithm of a sinh / cosh jet (whose coefficients all grow like exp(y) and cancel in the log recurrence).
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kIdealMaxC = 21;       // GERG-2008 has 21 components
constexpr int kIdealMaxTerms = 224;  // flat list over all components

enum IdealTermType : int {
    IG_CONSTANT = 0,   // a
    IG_INVT,           // a / T
    IG_LOGT,           // a ln T
    IG_LOGRATIO,       // a ln(b / T)                        (GERG: n3 ln(Tc/T))
    IG_POWERT,         // a T^b
    IG_PLANCK,         // a ln(1 - exp(-b / T))
    IG_PLANCKGEN,      // a ln(c + d exp(b / T))
    IG_LNCOSH,         // a ln|cosh(b / T)|
    IG_LNSINH,         // a ln|sinh(b / T)|
    IG_CP0CONST,       // a ((T - b)/T - ln(T / b))
    IG_CP0POWT         // a (T^b (1/(b+1) - 1/b) - c / T + d),   c = T0^(b+1)/(b+1), d = T0^b / b
};

struct IdealTerm {
    int type;
    int pad_;
    double a, b, c, d;
};

struct IdealParams {
    int nc;
    int gerg_rule;                     // 1: x_i > 0 contributes (GERG.hpp:415); 0: x_i != 0 (ideal_eosterms.hpp:345)
    int first[kIdealMaxC + 1];         // terms of component i: [first[i], first[i+1])
    double lnrho_coef[kIdealMaxC];     // number of "Lead" terms (each contributes ln rho)
    double rho_div[kIdealMaxC];        // ln(rho / rho_div): 1 for Lead terms, rhoc_i for GERG (GERG.hpp:408)
    double R;
    IdealTerm terms[kIdealMaxTerms];
};

template<class TT>
TQ_HD TT ideal_term(const IdealTerm& t, const TT& T, const TT& invT, const TT& lnT) {
    switch (t.type) {
        case IG_CONSTANT: return TT(t.a);
        case IG_INVT: return t.a * invT;
        case IG_LOGT: return t.a * lnT;
        case IG_LOGRATIO: return t.a * tlog(t.b * invT);
        case IG_POWERT: return t.a * tpow(T, t.b);
        case IG_PLANCK: return t.a * tlog(1.0 - texp((-t.b) * invT));
        case IG_PLANCKGEN: return t.a * tlog(t.c + t.d * texp(t.b * invT));
        case IG_LNCOSH: {   // ln cosh y = |y| - ln 2 + ln(1 + exp(-2|y|))
            const TT y = fabs(t.b) * invT;
            return t.a * ((y - 0.6931471805599453094) + tlog(1.0 + texp(-2.0 * y)));
        }
        case IG_LNSINH: {   // ln|sinh y| = |y| - ln 2 + ln(1 - exp(-2|y|))
            const TT y = fabs(t.b) * invT;
            return t.a * ((y - 0.6931471805599453094) + tlog(1.0 - texp(-2.0 * y)));
        }
        case IG_CP0CONST: return t.a * ((T - t.b) * invT - tlog(T / t.b));
        default: /* IG_CP0POWT */ return t.a * (tpow(T, t.b) * (1.0 / (t.b + 1.0) - 1.0 / t.b) - t.c * invT + t.d);
    }
}

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> ideal_alphaig(const IdealParams& p, const TT& T, const RT& rho, const XT* x) {
    using TR = prom_t<TT, RT>;
    using F = prom_t<TR, XT>;
    const TT invT = trecip(T);
    const TT lnT = tlog(T);
    F alpha(0.0);
    for (int i = 0; i < MAXN; ++i) {
        if (i >= p.nc) break;
        const double xb = base(x[i]);
        if (p.gerg_rule ? !(xb > 0) : (xb == 0)) continue;   // lim x ln x = 0
        TT f(0.0);
        for (int k = p.first[i]; k < p.first[i + 1]; ++k) f = f + ideal_term(p.terms[k], T, invT, lnT);
        TR pure = embed<TR>(f);
        if (p.lnrho_coef[i] != 0) pure = pure + p.lnrho_coef[i] * tlog(rho / p.rho_div[i]);
        alpha = alpha + x[i] * (pure + tlog(x[i]));
    }
    return alpha;
}

struct IdealDev {
    using Params = IdealParams;
    static constexpr int kMaxC = kIdealMaxC;
    static constexpr bool kNoSmallVariants = true;   // one kernel variant per jet shape, component count at run time
    static constexpr int kCrossOrder = 2;  // full get_ATrhoXiXj table from one J5 sweep
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return ideal_alphaig<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
