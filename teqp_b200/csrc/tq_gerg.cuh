// GERG-2004 / GERG-2008 (and, later, JSON-driven multifluid) residual Helmholtz energy -- the "term program" a model is compiled
// into on the host and that the sm_100a kernel interprets, one state point per thread.
//
// Mathematics (reference: /root/reference/include/teqp/models/GERG/GERG.hpp:103-122 pure terms, :179-205 reducing functions,
// :257-274 departure function, :322-340 departure sum, :1178-1191 GERG2008ResidualModel::alphar):
//     tau = T_r(x)/T, delta = rho/rho_r(x)
//     alphar = sum_i x_i sum_k n_ik tau^t_ik delta^d_ik exp(-c_ik delta^l_ik)
//            + sum_{i<j, F_ij != 0} x_i x_j F_ij sum_k n_ijk tau^t delta^d exp(-eta (delta-eps)^2 - beta (delta-gamma))
//
// The reference evaluates every one of the ~300 terms as exp(t ln tau + d ln delta - ...) on a Taylor number.  Here the sum is
// FACTORISED before it reaches the device.  Every term is  weight(x) * n * tau^t * g(delta)  where g belongs to a small set of
// distinct "density groups" (25 distinct (d,l) shapes + ~15 distinct Gaussian shapes in GERG-2008), so
//     alphar = sum_g  G_g(delta) * [ sum_{t in g} tau^t * ( sum_{terms with (g,t)} n * W[w] ) ]
// with W = (x_1..x_N, x_i x_j F_ij ...).  Only G_g needs Taylor arithmetic in delta (one exp per group), tau^t needs one exp
// per distinct (g,t), and each individual term costs ONE fused multiply-add.  Temperature derivatives are free:
// tau^n d^n(tau^t)/dtau^n = t(t-1)..(t-n+1) tau^t.
//
// Groups are sorted by exponent shape and then by d, so that the Taylor coefficients of exp(h(delta)) are computed once per
// distinct shape (new_h flag) and delta^d is a running product.  Exponentials use tq::fast_exp (tq_math.cuh).
#pragma once
#include "tq_jet.cuh"
#include "tq_math.cuh"

namespace tq {

constexpr int kGergMaxC = 21;
constexpr int kGergMaxPair = kGergMaxC * (kGergMaxC - 1) / 2;
constexpr int kGergMaxDep = 24;
constexpr int kGergMaxW = kGergMaxC + kGergMaxDep;
constexpr int kGergMaxGroup = 56;
constexpr int kGergMaxRec = 800;  // (g,t) headers + terms
constexpr int kGergND = 6;        // binomial tables cover derivative orders up to this
constexpr int kGergWS = 128;      // stride between consecutive weights in a thread's column of the staging area (= CTA size)

struct GergRedPair {       // reducing-function data of the pair (i<j), in lexicographic pair order
    double beta2T, YT2;  // beta_T,ij^2 ; 2 * beta_T gamma_T sqrt(Tc_i Tc_j)
    double beta2V, YV2;  // beta_v,ij^2 ; 2 * (1/8) beta_v gamma_v (vc_i^(1/3) + vc_j^(1/3))^3
};

enum GergGroupKind : int { GG_POLY = 0, GG_EXP = 1, GG_GAUSS = 2 };

// Temperature shape of a term: F(tau) = tau^t exp(h(tau)).  GERG-2004/2008 only has GT_PLAIN; the JSON-driven multifluid models
// (/root/reference/include/teqp/models/multifluid_eosterms.hpp) add
//   GT_GAUSS   h = -p1 (tau - p2)^2               GaussianEOSTerm :139-165 (beta, gamma)
//   GT_GAOB    h = 1 / (p1 (tau - p2)^2 + p3)     GaoBEOSTerm :199-243 (beta, gamma, b)
//   GT_POWEXP  h = -p1 tau^p2                      Lemmon2005EOSTerm :171-193 (1, m), DoubleExponentialEOSTerm :108-133 (gt, lt)
enum GergTauKind : int { GT_PLAIN = 0, GT_GAUSS = 1, GT_GAOB = 2, GT_POWEXP = 3 };
struct GergTauFn { int kind, pad_; double p1, p2, p3; };
constexpr int kGergMaxTF = 40;   // distinct non-plain temperature shapes per model
// Non-analytic critical-region terms (NonAnalyticEOSTerm, multifluid_eosterms.hpp:335-362; CO2, water):
//   n delta Delta^b exp(-C (delta-1)^2 - D (tau-1)^2),  Delta = theta^2 + B ((delta-1)^2)^a,  theta = (1 - tau) + A ((delta-1)^2)^(1/(2 beta))
// tau and delta do not separate, so these are evaluated by the jet sweeps only (a model that has them runs no generated / analytic kernel)
struct GergNA { double n, A, B, C, D, a, b, beta; int w, pad_; };
constexpr int kGergMaxNA = 8;

struct GergGroup {
    int kind;        // GG_POLY: delta^d ; GG_EXP: delta^d exp(-c delta^l) ; GG_GAUSS: delta^d exp(-eta(delta-eps)^2 - beta(delta-gamma))
    int d, l;        // integer exponents
    int ntq;         // number of distinct tau exponents in this group
    int new_h;       // 1: the exponent h(delta) differs from the previous group's -> recompute the Taylor coefficients of exp(h)
    int dstep;       // multiplications by delta that take the running delta^d from the previous group's d (0 after new_h) to this d
    int pad_[2];
    double c, eta, eps, beta, gamma;
    double bd[kGergND];  // binom(d, m), m = 1..6
    double bl[kGergND];  // binom(l, m), m = 1..6
};

// The (g,t) sums are one stream of 16-byte units, read front to back with warp-uniform addresses:
//   (g,t) header, 2 units: {val = t, a = number of term PAIRS that follow, b = byte offset of the first term's weight}
//                          {val = n of the first term, a = 1 + index into tf[] of the temperature shape, 0 = plain tau^t}
//   then per further term  {val = n, a = byte offset of its weight}, padded with {0, 0} to an even count
// (62 of the 109 (g,t) sums of the 10-component natural gas have a single term: no inner loop for those)
struct alignas(16) GergRec {
    double val;
    int a;
    int b;
};

struct GergProgram {
    int nc, nw, ndep, npair, ngroup, ntq, nterm, nrec;
    double Rgas;
    double Tc[kGergMaxC], vc[kGergMaxC];
    double depF[kGergMaxDep];
    unsigned char dep_i[kGergMaxDep], dep_j[kGergMaxDep];
    GergRedPair pair[kGergMaxPair];
    GergGroup group[kGergMaxGroup];
    GergRec rec[kGergMaxRec];
    // ---- JSON-driven multifluid extensions (all zero for the GERG kinds) ----
    int ntf;            // non-plain temperature shapes in tf[]
    int r_weighted;     // 1: R(x) = sum_i x_i Ri[i] (multifluid.hpp:973); 0: Rgas
    int limit_form;     // 1: a reducing pair whose denominator is zero contributes its limit form (multifluid_reducing.hpp:234-237); 0: skipped (GERG.hpp:189)
    int pure_shortcut;  // 1: single component: the mole fraction does not multiply the pure-fluid sum (multifluid.hpp:172-175)
    double Ri[kGergMaxC];
    GergTauFn tf[kGergMaxTF];
    int nna, pad2_;     // non-analytic terms in na[]
    GergNA na[kGergMaxNA];
};
static_assert(sizeof(GergProgram) <= 32500, "GergProgram must fit in the 32 KB kernel-parameter space (32764 B)");

TQ_HD double wload(const double* W, int byte_offset) {
    return *reinterpret_cast<const double*>(reinterpret_cast<const char*>(W) + byte_offset);
}
constexpr double kGergMaxAbsT = 30.0;  // largest |t| the device term program accepts (GERG-2008: 30)

TQ_HD double ipow(double x, int n) {  // x^n, n >= 0 small: repeated multiplication (n is warp-uniform)
    double r = 1.0;
    for (int k = 0; k < n; ++k) r *= x;
    return r;
}


// One state point.  W is this thread's column of the weight staging area (element w at W[w*kGergWS]); on entry W[0..nc) holds the
// mole fractions (WS = kGergWS).  On exit A[n][m] = tau^n delta^m d^(n+m) alphar / dtau^n ddelta^m / (n! m!).
template<int NT, int ND>
TQ_HD void gerg_point(const GergProgram& P, const Exp2K& K, const double T, const double rho, double* W, double (&A)[NT + 1][ND + 1]) {
    // ---- reducing functions (GERG.hpp:179-205), pairs in the reference's order; the two quotients of a pair share one
    // reciprocal: num/dT = num*dV/(dT*dV).  A pair is skipped when its denominator is zero (x_i = x_j = 0), GERG.hpp:189.
    double s1T = 0.0, s1V = 0.0, s2T = 0.0, s2V = 0.0;
    {
        int pidx = 0;
        for (int i = 0; i < P.nc; ++i) {
            const double xi = W[i * kGergWS];
            const double xi2 = xi * xi;
            s1T = fma(xi2, P.Tc[i], s1T);
            s1V = fma(xi2, P.vc[i], s1V);
            for (int j = i + 1; j < P.nc; ++j, ++pidx) {
                const double xj = W[j * kGergWS];
                const GergRedPair pr = P.pair[pidx];
                const double dT = fma(pr.beta2T, xi, xj);
                const double dV = fma(pr.beta2V, xi, xj);
                const double den = dT * dV;
                const double q = (xi * xj) * (xi + xj) * fast_rcp(den != 0.0 ? den : 1.0);  // numerator is 0 whenever den is
                s2T = fma(q * dV, pr.YT2, s2T);
                s2V = fma(q * dT, pr.YV2, s2V);
            }
        }
    }
    const double Tr = s1T + s2T;
    const double Yv = s1V + s2V;  // 1/rho_r
    const double tau = Tr / T;
    const double delta = rho * Yv;
    // tau^t = 2^(t log2 tau); |t| <= kGergMaxAbsT is checked when the program is built, log2 tau is clamped so that the
    // product stays inside fast_exp2_bounded's range.  nanfix turns a non-finite log2 tau (T <= 0, NaN) into NaN results.
    double l2tau = log2(tau);
    const double nanfix = l2tau - l2tau;
    l2tau = (l2tau > 30.0) ? 30.0 : l2tau;
    l2tau = (l2tau < -30.0) ? -30.0 : l2tau;
    for (int q = 0; q < P.ndep; ++q) W[(P.nc + q) * kGergWS] = W[P.dep_i[q] * kGergWS] * W[P.dep_j[q] * kGergWS] * P.depF[q];

    // ---- factorised term sum ------------------------------------------------------------------------------------
#pragma unroll
    for (int n = 0; n <= NT; ++n)
#pragma unroll
        for (int m = 0; m <= ND; ++m) A[n][m] = nanfix;

    int rec = 0;
    double E[ND + 1];   // delta^m/m! d^m/ddelta^m exp(h(delta)), shared by consecutive groups with the same h
    double dd = 1.0;    // running delta^d
    for (int g = 0; g < P.ngroup; ++g) {
        const GergGroup& gr = P.group[g];
        if (gr.new_h) {
            double h[ND + 1];
            if (gr.kind == GG_POLY) {
#pragma unroll
                for (int m = 0; m <= ND; ++m) h[m] = 0.0;
            } else if (gr.kind == GG_EXP) {
                const double cdl = -gr.c * ipow(delta, gr.l);
                h[0] = cdl;
#pragma unroll
                for (int m = 1; m <= ND; ++m) h[m] = gr.bl[m - 1] * cdl;
            } else {
                const double de = delta - gr.eps;
                h[0] = -gr.eta * (de * de) - gr.beta * (delta - gr.gamma);
                if (ND >= 1) h[ND >= 1 ? 1 : 0] = delta * (-2.0 * gr.eta * de - gr.beta);
                if (ND >= 2) h[ND >= 2 ? 2 : 0] = -gr.eta * (delta * delta);
#pragma unroll
                for (int m = 3; m <= ND; ++m) h[m] = 0.0;
            }
            E[0] = (gr.kind == GG_POLY) ? 1.0 : fast_exp(h[0]);
#pragma unroll
            for (int m = 1; m <= ND; ++m) {
                double s = h[1 <= ND ? 1 : 0] * E[m - 1];
#pragma unroll
                for (int j = 2; j <= m; ++j) s = fma(double(j) * h[j], E[m - j], s);
                E[m] = s * (1.0 / double(m));
            }
            dd = 1.0;
        }
#pragma unroll 1
        for (int s = 0; s < gr.dstep; ++s) dd *= delta;
        // G[m] = delta^m/m! d^m/ddelta^m [ delta^d exp(h(delta)) ] = delta^d sum_j binom(d,j) E[m-j]
        double G[ND + 1];
#pragma unroll
        for (int m = 0; m <= ND; ++m) {
            double s = E[m];
#pragma unroll
            for (int j = 1; j <= m; ++j) s = fma(gr.bd[j - 1], E[m - j], s);
            G[m] = s * dd;
        }
        // C[n] = tau^n/n! d^n/dtau^n sum_t c_t tau^t
        double C[NT + 1];
#pragma unroll
        for (int n = 0; n <= NT; ++n) C[n] = 0.0;
        // (g,t) sums two at a time: the two 2^(t log2 tau) Horner chains are independent and interleave, which hides the DFMA
        // latency that a single 11-deep chain exposes (r01g: "wait" + "short scoreboard" stalls of 5 cycles per issue)
        auto term_sum = [&](int r, const GergRec& hd, const double n0) {
            double c = wload(W, hd.b) * n0;
#pragma unroll 1
            for (int k = 0; k < hd.a; ++k) {
                const GergRec r0 = P.rec[r], r1 = P.rec[r + 1];
                r += 2;
                c = fma(wload(W, r0.a), r0.val, c);
                c = fma(wload(W, r1.a), r1.val, c);
            }
            return c;
        };
        auto accumulate = [&](const double t, double v) {
            C[0] += v;
#pragma unroll
            for (int n = 1; n <= NT; ++n) {
                v *= (t - double(n - 1)) * (1.0 / double(n));
                C[n] += v;
            }
        };
        int q = 0;
#pragma unroll 1
        for (; q + 1 < gr.ntq; q += 2) {
            const GergRec hA = P.rec[rec];
            const double nA = P.rec[rec + 1].val;
            const int recB = rec + 2 + 2 * hA.a;
            const GergRec hB = P.rec[recB];
            const double nB = P.rec[recB + 1].val;
            const double eA = fast_exp2_prod(hA.val, l2tau, K);
            const double eB = fast_exp2_prod(hB.val, l2tau, K);
            const double cA = term_sum(rec + 2, hA, nA);
            const double cB = term_sum(recB + 2, hB, nB);
            rec = recB + 2 + 2 * hB.a;
            accumulate(hA.val, cA * eA);
            accumulate(hB.val, cB * eB);
        }
        if (q < gr.ntq) {
            const GergRec hd = P.rec[rec];
            const double n0 = P.rec[rec + 1].val;
            const double e = fast_exp2_prod(hd.val, l2tau, K);
            const double c = term_sum(rec + 2, hd, n0);
            rec += 2 + 2 * hd.a;
            accumulate(hd.val, c * e);
        }
#pragma unroll
        for (int n = 0; n <= NT; ++n)
#pragma unroll
            for (int m = 0; m <= ND; ++m) A[n][m] = fma(C[n], G[m], A[n][m]);
    }
}

}  // namespace tq
