// RK-PR cubic of Cismondi and Mollerup (Fluid Phase Equilib. 232 (2005) 74), written once over jet types.
//
// Reference arithmetic: RKPRCismondi2005  /root/reference/include/teqp/models/cubics/simple_cubics.hpp:593-690 -- a_i(T) = a_c,i (3 / (2 + T/Tc_i))^k_i (:648),
// constant covolumes (:643), quadratic mixing with k_ij and l_ij (:652-671), and a third parameter Delta_1 = sum_i x_i delta_1,i (mole-fraction average,
// :680-682) with Delta_2 = (1 - Delta_1) / (1 + Delta_1).  Model kind "RKPRCismondi2005" (/root/reference/interface/CPP/teqp_impl_factory.cpp:61).
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kRkprMaxC = 8;

struct RkprParams {
    int nc, pad_;
    double R;
    double Tc[kRkprMaxC], ac[kRkprMaxC], bc[kRkprMaxC], k[kRkprMaxC], delta1[kRkprMaxC];
    double one_minus_k[kRkprMaxC * kRkprMaxC], one_minus_l[kRkprMaxC * kRkprMaxC];   // row-major [i * nc + j]
};

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> rkpr_alphar(const RkprParams& p, const TT& T, const RT& rho, const XT* x) {
    using TX = prom_t<TT, XT>;
    const int N = p.nc;
    XT Delta1(0.0);
    TT ai[MAXN];
    for (int i = 0; i < MAXN; ++i) {
        if (i >= N) break;
        Delta1 = Delta1 + x[i] * p.delta1[i];
        ai[i] = p.ac[i] * tpow(3.0 / (2.0 + T / p.Tc[i]), p.k[i]);
    }
    TX a(0.0);
    XT b(0.0);
    for (int i = 0; i < MAXN; ++i) {
        if (i >= N) break;
        for (int j = 0; j < MAXN; ++j) {
            if (j >= N) break;
            const XT w = x[i] * x[j];
            a = a + w * (tsqrt(ai[i] * ai[j]) * p.one_minus_k[i * N + j]);
            b = b + w * ((p.bc[i] + p.bc[j]) / 2.0 * p.one_minus_l[i * N + j]);
        }
    }
    const XT Delta2 = (1.0 - Delta1) / (1.0 + Delta1);
    const auto brho = b * rho;
    const auto Psiminus = -tlog(1.0 - brho);
    const auto Psiplus = tlog((Delta1 * brho + 1.0) / (Delta2 * brho + 1.0)) / (b * (Delta1 - Delta2));
    return Psiminus - (a / (p.R * T)) * Psiplus;
}

struct RkprDev {
    using Params = RkprParams;
    static constexpr int kMaxC = kRkprMaxC;
    static constexpr bool kNoSmallVariants = true;
    static constexpr int kCrossOrder = 2;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return rkpr_alphar<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
