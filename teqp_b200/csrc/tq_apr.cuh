// Peng-Robinson with the EoS/aE mixing rule of Jaubert and co-workers and a Wilson residual excess Helmholtz energy, written once over jet types.
//
// Reference arithmetic: AdvancedPRaEres  /root/reference/include/teqp/models/cubics/advancedmixing_cubics.hpp:53-218 -- a/b = aE_res/CEoS + sum_i x_i a_i(T)/b_i
// (:165-173), b by the quadratic rule with exponent s or linearly (:175-200), alphar :202-217; WilsonResidualHelmholtzOverRT
// /root/reference/include/teqp/models/activity/activity_models.hpp:40-101 (total - combinatorial).  Model kind "advancedPRaEres"
// (/root/reference/interface/CPP/teqp_impl_factory.cpp:59); its factory make_AdvancedPRaEres (:220-250) passes a 2 x 2 zero l_ij matrix, so the
// reference only builds binaries.
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kAprMaxC = 2;

struct AprParams {
    int nc, linear_b;
    int alpha_kind[kAprMaxC], pad_[2];                   // CubicAlphaKind numbering: 0 Basic (c0 = m), 1 Twu, 2 Mathias-Copeman
    double R, CEoS;
    double Tc[kAprMaxC], ai[kAprMaxC], bi[kAprMaxC], c0[kAprMaxC], c1[kAprMaxC], c2[kAprMaxC];
    double bij[kAprMaxC * kAprMaxC];                     // (1 - l_ij) ((b_i^(1/s) + b_j^(1/s)) / 2)^s
    double v[kAprMaxC];                                  // Wilson volumes: OmegaB R_CODATA Tc / pc (advancedmixing_cubics.hpp:113-117, 231-233)
    double vratio[kAprMaxC * kAprMaxC];                  // v_j / v_i at [i * nc + j]
    double m[kAprMaxC * kAprMaxC], n[kAprMaxC * kAprMaxC];
};

template<class TT> TQ_HD TT apr_alpha(const AprParams& p, int i, const TT& T) {
    // alpha functions of simple_cubics.hpp:28-93
    if (p.alpha_kind[i] == 0) {
        const TT s = 1.0 + p.c0[i] * (1.0 - tsqrt(T / p.Tc[i]));
        return s * s;
    }
    if (p.alpha_kind[i] == 1) {
        const TT Tr = T / p.Tc[i];
        return tpow(Tr, p.c2[i] * (p.c1[i] - 1.0)) * texp(p.c0[i] * (1.0 - tpow(Tr, p.c1[i] * p.c2[i])));
    }
    const TT q = 1.0 - tsqrt(T / p.Tc[i]);
    const TT paren = 1.0 + p.c0[i] * q + p.c1[i] * q * q + p.c2[i] * q * q * q;
    return paren * paren;
}

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> apr_alphar(const AprParams& p, const TT& T, const RT& rho, const XT* x) {
    using TX = prom_t<TT, XT>;
    const int N = p.nc;
    // covolume
    XT b(0.0);
    if (p.linear_b) {
        for (int i = 0; i < N; ++i) b = b + x[i] * p.bi[i];
    } else {
        for (int i = 0; i < N; ++i)
            for (int j = 0; j < N; ++j) b = b + (x[i] * x[j]) * p.bij[i * N + j];
    }
    // Wilson aE_res / RT = total - combinatorial
    XT Vtot(0.0);
    for (int i = 0; i < N; ++i) Vtot = Vtot + x[i] * p.v[i];
    XT comb(0.0);
    TX total(0.0);
    for (int i = 0; i < N; ++i) {
        comb = comb + x[i] * tlog(p.v[i] / Vtot);
        TX sj(0.0);
        for (int j = 0; j < N; ++j) {
            const TT Aij = p.m[i * N + j] * T + p.n[i * N + j];
            const TT Omega = p.vratio[i * N + j] * texp(-(Aij / T));
            sj = sj + x[j] * Omega;
        }
        total = total + x[i] * tlog(sj);
    }
    const TX aEresRT = (-total) - comb;
    TX a_over_b = aEresRT * ((p.R * T) / p.CEoS);
    for (int i = 0; i < N; ++i) a_over_b = a_over_b + x[i] * ((p.ai[i] * apr_alpha(p, i, T)) / p.bi[i]);
    const TX a = a_over_b * b;
    const double Delta1 = 1.0 + 1.4142135623730951, Delta2 = 1.0 - 1.4142135623730951;
    const auto brho = b * rho;
    const auto Psiminus = -tlog(1.0 - brho);
    const auto Psiplus = tlog((Delta1 * brho + 1.0) / (Delta2 * brho + 1.0)) / (b * (Delta1 - Delta2));
    return Psiminus - (a / (p.R * T)) * Psiplus;
}

struct AprDev {
    using Params = AprParams;
    static constexpr int kMaxC = kAprMaxC;
    static constexpr bool kNoSmallVariants = true;
    static constexpr int kCrossOrder = 2;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return apr_alphar<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
