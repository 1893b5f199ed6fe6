// Cubic equations of state (Peng-Robinson, Soave-Redlich-Kwong, generalized cubic with Twu / Mathias-Copeman alpha functions,
// van der Waals) -- device parameter block and the model arithmetic.
//
// Reference arithmetic: GenericCubic::alphar  /root/reference/include/teqp/models/cubics/simple_cubics.hpp:242-255,
//   get_a :217-231, get_b :233-240, alpha functions :28-93; vdWEOS1 / vdWEOS  /root/reference/include/teqp/models/vdW.hpp:32-35, 115-127.
//
//   alphar = Psi_minus(rho) - a(T,x)/(R T) * Psi_plus(rho)
//   Psi_minus = -ln(1 - b rho)
//   Psi_plus  = ln((Delta1 b rho + 1)/(Delta2 b rho + 1)) / (b (Delta1 - Delta2))      (cubic)
//             = rho                                                                      (van der Waals)
#pragma once
#include "tq_jet.cuh"
#include "tq_math.cuh"

namespace tq {

constexpr int kCubicMaxC = 24;
enum CubicAlphaKind : int { ALPHA_BASIC = 0, ALPHA_TWU = 1, ALPHA_MC = 2, ALPHA_NONE = 3 };

struct CubicParams {
    int nc;
    int is_vdw;  // 1: Psi_plus = rho, alpha_i = 1;  2: same, single fluid whose mole-fraction argument is ignored (vdWEOS1)
    double Delta1, Delta2, R;
    double ai[kCubicMaxC], bi[kCubicMaxC], Tci[kCubicMaxC];
    int alpha_kind[kCubicMaxC];
    double c0[kCubicMaxC], c1[kCubicMaxC], c2[kCubicMaxC];  // Basic: c0 = m_i
    double one_minus_k[kCubicMaxC * kCubicMaxC];           // (1 - k_ij), row-major
};

// alpha_i(T) in whatever jet type T has
template<class TT> TQ_HD TT cubic_alpha(const CubicParams& p, int i, const TT& T) {
    const int kind = p.alpha_kind[i];
    if (kind == ALPHA_BASIC) {
        TT s = 1.0 + p.c0[i] * (1.0 - tsqrt(T / p.Tci[i]));
        return s * s;
    } else if (kind == ALPHA_TWU) {
        TT Tr = T / p.Tci[i];
        return tpow(Tr, p.c2[i] * (p.c1[i] - 1.0)) * texp(p.c0[i] * (1.0 - tpow(Tr, p.c1[i] * p.c2[i])));
    } else if (kind == ALPHA_MC) {
        TT x = 1.0 - tsqrt(T / p.Tci[i]);
        TT paren = 1.0 + p.c0[i] * x + p.c1[i] * x * x + p.c2[i] * x * x * x;
        return paren * paren;
    }
    return TT(1.0);
}

// a(T, x): double sum with geometric-mean combining rule (simple_cubics.hpp:217-231); MAXN bounds the local array
template<int MAXN, class TT, class XT>
TQ_HD prom_t<TT, XT> cubic_a(const CubicParams& p, const TT& T, const XT* x) {
    using R = prom_t<TT, XT>;
    TT aal[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i)
        if (i < p.nc) aal[i] = p.ai[i] * cubic_alpha(p, i, T);
    R a(0.0);
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i) {
        if (i < p.nc) {
#pragma unroll(MAXN <= 4 ? MAXN : 1)
            for (int j = 0; j < MAXN; ++j) {
                if (j < p.nc) {
                    TT aij = p.one_minus_k[i * p.nc + j] * tsqrt(aal[i] * aal[j]);
                    a = a + (x[i] * x[j]) * aij;
                }
            }
        }
    }
    return a;
}

template<int MAXN, class XT> TQ_HD XT cubic_b(const CubicParams& p, const XT* x) {
    XT b(0.0);
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i)
        if (i < p.nc) b = b + x[i] * p.bi[i];
    return b;
}

// generic-jet alphar (used for composition derivatives and the isochoric family)
template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> cubic_alphar(const CubicParams& p, const TT& T, const RT& rho, const XT* x) {
    if (p.is_vdw == 2) {  // vdWEOS1 ignores the composition argument (vdW.hpp:32-35)
        using R = prom_t<prom_t<TT, RT>, XT>;
        return embed<R>(-tlog(1.0 - p.bi[0] * rho) - (p.ai[0] / (p.R * T)) * rho);
    }
    auto b = cubic_b<MAXN>(p, x);
    auto brho = b * rho;
    auto Psiminus = -tlog(1.0 - brho);
    auto a = cubic_a<MAXN>(p, T, x);
    auto q = a / (p.R * T);
    if (p.is_vdw) return Psiminus - q * rho;
    auto Psiplus = tlog(((p.Delta1 * b) * rho + 1.0) / ((p.Delta2 * b) * rho + 1.0)) / (b * (p.Delta1 - p.Delta2));
    return Psiminus - q * Psiplus;
}

// One state point of the Lambda^r_nm table (closed-form density derivatives, 1/T jet for a(T)/(RT)); writes
// out[(n*(ND+1)+m)*ostride] for n <= NT, m <= ND.
template<int NT, int MAXN>
TQ_HD void cubic_table_point(const CubicParams& p, const double T, const double rho, double* x, const int ND, double* out, const long long ostride) {
    if (p.is_vdw == 2) x[0] = 1.0;  // vdWEOS1 ignores its composition argument
    using TJ = Tay<NT, double>;
    const TJ Tj = TJ::reciprocal_scaled_variable(T);
    const double b = cubic_b<MAXN>(p, x);
    const TJ q = cubic_a<MAXN>(p, Tj, x) / (p.R * Tj);  // c[n] = (1/T)^n d^n q / d(1/T)^n / n!

    // the m = 0 values follow the reference's operation order (simple_cubics.hpp:250-253): at low density -ln(1-b rho) and the
    // Psi_plus logarithm amplify a 1-ulp difference in their arguments by 1/(b rho)
    // Only the m = 0 entries keep IEEE divisions; the ratios that feed the m >= 1 closed forms use a Newton-refined reciprocal (no
    // slow-path guards: 1 - y, n1, n2 and b (Delta1 - Delta2) are O(1) and away from zero for every physical state)
    const double y = b * rho;
    const double u = y * fast_rcp(1.0 - y);
    double v1 = 0.0, v2 = 0.0, invbd = 0.0;
    double psi_m = -log(1.0 - y);
    double psi_p;
    if (p.is_vdw) {
        psi_p = rho;
    } else {
        const double n1 = p.Delta1 * b * rho + 1.0, n2 = p.Delta2 * b * rho + 1.0;
        const double bd = b * (p.Delta1 - p.Delta2);
        invbd = fast_rcp(bd);
        psi_p = log(n1 / n2) / bd;
        v1 = p.Delta1 * y * fast_rcp(n1);
        v2 = p.Delta2 * y * fast_rcp(n2);
    }
    double um = 1.0, v1m = 1.0, v2m = 1.0, fact = 1.0;  // fact = (m-1)!
    for (int m = 0; m <= ND; ++m) {
        if (m >= 1) {
            um *= u; v1m *= v1; v2m *= v2;
            if (m >= 2) fact *= double(m - 1);
            psi_m = fact * um;
            if (p.is_vdw) psi_p = (m == 1) ? rho : 0.0;
            else psi_p = ((m & 1) ? fact : -fact) * (v1m - v2m) * invbd;
        }
        double nfact = 1.0;
#pragma unroll
        for (int n = 0; n <= NT; ++n) {
            if (n >= 2) nfact *= double(n);
            const double val = (n == 0 ? psi_m : 0.0) - (nfact * q.c[n]) * psi_p;
            out[(long long)(n * (ND + 1) + m) * ostride] = val;
        }
    }
}

struct CubicDev {
    using Params = CubicParams;
    static constexpr int kMaxC = kCubicMaxC;
    static constexpr int kCrossOrder = 2;  // order of x_i and x_j in the get_ATrhoXiXj jet (2: full table; 1: mixed entries only)
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return cubic_alphar<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
