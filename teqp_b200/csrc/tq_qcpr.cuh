// Quantum-corrected Peng-Robinson (Aasen, Hammer, Lasala, Jaubert, Wilhelmsen, Fluid Phase Equilib. 524 (2020) 112790), written once over jet types.
//
// Reference arithmetic: QuantumCorrectedPR  /root/reference/include/teqp/models/cubics/simple_cubics.hpp:447-557 -- temperature-dependent covolumes
// b_i(T) = 0.07780 R Tc_i / pc_i * ((1 + A_i/(T + B_i)) / (1 + A_i/(Tc_i + B_i)))^3 (:486-491), Twu alpha functions (:49-75, coefficients L, M, N),
// quadratic mixing with k_ij and l_ij (:503-520), Peneloux volume translation rho -> 1/(1/rho + c(x)) (:521-541).  Model kind "QCPRAasen"
// (/root/reference/interface/CPP/teqp_impl_factory.cpp:60).
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kQcprMaxC = 8;

struct QcprParams {
    int nc, pad_;
    double R;
    double Tc[kQcprMaxC], a0[kQcprMaxC], b0[kQcprMaxC], bden[kQcprMaxC];   // a0 = OmegaA (Tc R)^2/pc, b0 = 0.07780 Tc R/pc, bden = (1 + A/(Tc + B))^3
    double L[kQcprMaxC], M[kQcprMaxC], N[kQcprMaxC], A[kQcprMaxC], B[kQcprMaxC], c[kQcprMaxC];
    double one_minus_k[kQcprMaxC * kQcprMaxC], one_minus_l[kQcprMaxC * kQcprMaxC];   // row-major [i * nc + j]
};

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> qcpr_alphar(const QcprParams& p, const TT& T, const RT& rhoinit, const XT* x) {
    using TX = prom_t<TT, XT>;
    using RX = prom_t<RT, XT>;
    const int N = p.nc;
    XT c(0.0);
    TT ai[MAXN], bi[MAXN];
    for (int i = 0; i < MAXN; ++i) {
        if (i >= N) break;
        c = c + x[i] * p.c[i];
        const TT Tr = T / p.Tc[i];
        const TT alpha = tpow(Tr, p.N[i] * (p.M[i] - 1.0)) * texp(p.L[i] * (1.0 - tpow(Tr, p.M[i] * p.N[i])));   // Twu, coefficients (L, M, N)
        ai[i] = p.a0[i] * alpha;
        const TT q = 1.0 + p.A[i] / (T + p.B[i]);
        bi[i] = (p.b0[i] * tcube(q)) / p.bden[i];
    }
    TX a(0.0), b(0.0);
    for (int i = 0; i < MAXN; ++i) {
        if (i >= N) break;
        for (int j = 0; j < MAXN; ++j) {
            if (j >= N) break;
            const XT w = x[i] * x[j];
            b = b + w * ((bi[i] + bi[j]) / 2.0 * p.one_minus_l[i * N + j]);
            a = a + w * (tsqrt(ai[i] * ai[j]) * p.one_minus_k[i * N + j]);
        }
    }
    const RX rho = trecip(trecip(rhoinit) + c);   // volume translation: 1/(1/rho + c)
    const double Delta1 = 1.0 + 1.4142135623730951, Delta2 = 1.0 - 1.4142135623730951;
    const auto brho = b * rho;
    const auto Psiminus = -tlog(1.0 - brho);
    const auto Psiplus = tlog((Delta1 * brho + 1.0) / (Delta2 * brho + 1.0)) / (b * (Delta1 - Delta2));
    return Psiminus - (a / (p.R * T)) * Psiplus;
}

struct QcprDev {
    using Params = QcprParams;
    static constexpr int kMaxC = kQcprMaxC;
    static constexpr bool kNoSmallVariants = true;
    static constexpr int kCrossOrder = 2;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return qcpr_alphar<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
