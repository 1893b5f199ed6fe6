// GERG-2004 / GERG-2008 residual Helmholtz energy over jet types: the SAME factorised term program the headline kernel
// interprets (tq_gerg.cuh), evaluated in truncated-Taylor arithmetic in 1/T, rho and one composition direction, so that the
// generic jet-sweep kernels (tq_generic.cuh) serve the composition-derivative entry points of the GERG families:
// get_ATrhoXi (derivs.hpp:247-292), and anything else that needs derivatives the analytic kernels do not provide.
//
// Reference arithmetic: GERG200XReducing::Y (GERG.hpp:179-205, pairs whose denominator's base value is zero are skipped, :189),
// pure terms :103-122, departure terms :257-274, GERG2008ResidualModel::alphar :1178-1191.
//     alphar = sum_g  delta^d_g exp(h_g(delta)) * [ sum_{t in g} tau^t * ( sum_{terms (g,t)} n * W_w(x) ) ]
// Only products of delta and one exponential per density shape are taken in the (rho, x) jet, tau^t = exp(t ln tau) in the (T, x)
// jet, and the innermost sums in the composition jet alone.
#pragma once
#include "tq_gerg.cuh"

namespace tq {

// the part after the reducing functions: alphar(tau, delta, x) (MultiFluid::alphar_taudelta, multifluid.hpp:179-191)
template<int MAXN, class TX, class RX, class XT>
TQ_HD prom_t<TX, RX> gerg_alphar_taudelta_jet(const GergProgram& P, const TX& tau, const RX& delta, const XT* x) {
    using F = prom_t<TX, RX>;
    const int nc = P.nc;
    const TX lntau = tlog(tau);
    // weights
    XT W[kGergMaxW];
    for (int i = 0; i < nc; ++i) W[i] = x[i];
    if (P.pure_shortcut) W[0] = XT(1.0);
    for (int q = 0; q < P.ndep; ++q) W[nc + q] = (x[P.dep_i[q]] * x[P.dep_j[q]]) * P.depF[q];

    F A(0.0);
    RX Eh(1.0), dd(1.0);
    int rec = 0;
    for (int g = 0; g < P.ngroup; ++g) {
        const GergGroup& gr = P.group[g];
        if (gr.new_h) {
            if (gr.kind == GG_POLY) {
                Eh = RX(1.0);
            } else if (gr.kind == GG_EXP) {
                Eh = texp(tpowi(delta, gr.l) * (-gr.c));
            } else {
                const RX de = delta - gr.eps;
                Eh = texp((de * de) * (-gr.eta) - (delta - gr.gamma) * gr.beta);
            }
            dd = RX(1.0);
        }
        for (int s = 0; s < gr.dstep; ++s) dd = dd * delta;
        const RX G = Eh * dd;
        TX C(0.0);
        for (int q = 0; q < gr.ntq; ++q) {
            const GergRec hd = P.rec[rec];
            const double n0 = P.rec[rec + 1].val;
            const int tfi = P.rec[rec + 1].a;
            rec += 2;
            XT c = W[hd.b / (kGergWS * 8)] * n0;
            for (int k = 0; k < 2 * hd.a; ++k, ++rec) {
                const GergRec r = P.rec[rec];
                c = c + W[r.a / (kGergWS * 8)] * r.val;  // padding records carry n = 0
            }
            TX arg = lntau * hd.val;
            if (tfi > 0) {
                const GergTauFn f = P.tf[tfi - 1];
                if (f.kind == GT_GAUSS) { const TX dt = tau - f.p2; arg = arg - (dt * dt) * f.p1; }
                else if (f.kind == GT_GAOB) { const TX dt = tau - f.p2; arg = arg + 1.0 / ((dt * dt) * f.p1 + f.p3); }
                else arg = arg - texp(lntau * f.p2) * f.p1;
            }
            C = C + texp(arg) * c;
        }
        A = A + C * G;
    }
    // non-analytic terms (multifluid_eosterms.hpp:335-362); a non-finite value (the critical point itself) counts as zero, as in the reference
    for (int q = 0; q < P.nna; ++q) {
        const GergNA t = P.na[q];
        const RX dm1 = delta - 1.0;
        const RX d2 = dm1 * dm1;
        const TX tm1 = tau - 1.0;
        const F Psi = texp(embed<F>(d2 * (-t.C)) - embed<F>((tm1 * tm1) * t.D));
        const F theta = embed<F>(1.0 - tau) + embed<F>(tpow(d2, 1.0 / (2.0 * t.beta)) * t.A);
        const F Delta = theta * theta + embed<F>(tpow(d2, t.a) * t.B);
        F term = (tpow(Delta, t.b) * Psi) * delta;
        const double bv = base(term);
        if (!(bv - bv == 0.0)) term = F(0.0);
        A = A + term * (W[t.w] * t.n);
    }
    return A;
}

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> gerg_alphar_jet(const GergProgram& P, const TT& T, const RT& rho, const XT* x) {
    using TX = prom_t<TT, XT>;
    using RX = prom_t<RT, XT>;
    const int nc = P.nc;
    // reducing functions
    XT YT(0.0), YV(0.0);
    {
        int pidx = 0;
        for (int i = 0; i < nc; ++i) {
            const XT xi2 = x[i] * x[i];
            YT = YT + xi2 * P.Tc[i];
            YV = YV + xi2 * P.vc[i];
            for (int j = i + 1; j < nc; ++j, ++pidx) {
                const GergRedPair pr = P.pair[pidx];
                const XT dT = x[i] * pr.beta2T + x[j];
                const XT dV = x[i] * pr.beta2V + x[j];
                const XT num = (x[i] * x[j]) * (x[i] + x[j]);
                // zero denominator (x_i = x_j = 0): GERG skips the pair (GERG.hpp:189); multifluid adds the limit form
                // Y_ij (x_i x_j + x_i^2 (1 - beta^2)) (multifluid_reducing.hpp:234-237), whose first composition derivatives differ
                if (base(dT) != 0.0) YT = YT + (num / dT) * pr.YT2;
                else if (P.limit_form) YT = YT + (x[i] * x[j] + xi2 * (1.0 - pr.beta2T)) * pr.YT2;
                if (base(dV) != 0.0) YV = YV + (num / dV) * pr.YV2;
                else if (P.limit_form) YV = YV + (x[i] * x[j] + xi2 * (1.0 - pr.beta2V)) * pr.YV2;
            }
        }
    }
    const TX tau = YT * trecip(T);
    const RX delta = rho * YV;
    return gerg_alphar_taudelta_jet<MAXN>(P, tau, delta, x);
}

struct GergJetDev {
    using Params = GergProgram;
    static constexpr int kMaxC = kGergMaxC;
    static constexpr int kCrossOrder = 2;  // order of x_i and x_j in the get_ATrhoXiXj jet (2: full table; 1: mixed entries only)
    static constexpr bool kNoSmallVariants = true;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return gerg_alphar_jet<MAXN>(p, T, rho, x);
    }
    // gas constant: a plain number for the GERG kinds, mole-fraction weighted for the JSON-driven multifluid models (multifluid.hpp:973)
    template<class XT> static TQ_HD XT R(const Params& p, const XT* x) {
        if (!p.r_weighted) return XT(p.Rgas);
        XT r(0.0);
        for (int i = 0; i < p.nc; ++i) r = r + x[i] * p.Ri[i];
        return r;
    }
};

// alphar as a function of (tau, delta, x): the generic kernels are driven with "T" := 1/tau and "rho" := delta, so their scaled
// (1/T)^n d^n/d(1/T)^n and rho^m d^m/drho^m are tau^n d^n/dtau^n and delta^m d^m/ddelta^m -- get_AtaudeltaXi / XiXj / XiXjXk
// (derivs.hpp:341-414)
struct GergTauDeltaDev {
    using Params = GergProgram;
    static constexpr int kMaxC = kGergMaxC;
    static constexpr int kCrossOrder = 2;
    static constexpr bool kNoSmallVariants = true;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        using TX = prom_t<TT, XT>;
        using RX = prom_t<RT, XT>;
        return gerg_alphar_taudelta_jet<MAXN>(p, embed<TX>(trecip(T)), embed<RX>(rho), x);
    }
    template<class XT> static TQ_HD XT R(const Params& p, const XT* x) { return GergJetDev::R(p, x); }
};

}  // namespace tq
