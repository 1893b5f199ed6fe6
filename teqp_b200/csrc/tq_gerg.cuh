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
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kGergMaxC = 21;
constexpr int kGergMaxPair = kGergMaxC * (kGergMaxC - 1) / 2;
constexpr int kGergMaxDep = 24;
constexpr int kGergMaxW = kGergMaxC + kGergMaxDep;
constexpr int kGergMaxGroup = 64;
constexpr int kGergMaxTQ = 320;
constexpr int kGergMaxTerm = 520;
constexpr int kGergND = 6;  // binomial tables cover derivative orders up to this

struct GergRedPair {       // reducing-function data of the pair (i<j), in lexicographic pair order
    double beta2T, YT2;  // beta_T,ij^2 ; 2 * beta_T gamma_T sqrt(Tc_i Tc_j)
    double beta2V, YV2;  // beta_v,ij^2 ; 2 * (1/8) beta_v gamma_v (vc_i^(1/3) + vc_j^(1/3))^3
};

enum GergGroupKind : int { GG_POLY = 0, GG_EXP = 1, GG_GAUSS = 2 };

struct GergGroup {
    int kind;        // GG_POLY: delta^d ; GG_EXP: delta^d exp(-c delta^l) ; GG_GAUSS: delta^d exp(-eta(delta-eps)^2 - beta(delta-gamma))
    int d, l;        // integer exponents
    int tq_end;      // one past this group's last entry in tq[]
    double c, eta, eps, beta, gamma;
    double bd[kGergND];  // binom(d, m), m = 1..6
    double bl[kGergND];  // binom(l, m), m = 1..6
};

struct GergTQ {       // a distinct tau exponent inside a group
    double t;
    int term_end;    // one past the last term
};

struct GergProgram {
    int nc, nw, ndep, npair, ngroup, ntq, nterm;
    int pad_;
    double Tc[kGergMaxC], vc[kGergMaxC];
    GergRedPair pair[kGergMaxPair];
    double depF[kGergMaxDep];
    unsigned char dep_i[kGergMaxDep], dep_j[kGergMaxDep];
    GergGroup group[kGergMaxGroup];
    GergTQ tq[kGergMaxTQ];
    double term_n[kGergMaxTerm];
    unsigned char term_w[kGergMaxTerm];
};
static_assert(sizeof(GergProgram) <= 32000, "GergProgram must fit in the 32 KB kernel-parameter space");

TQ_HD double ipow(double x, int n) {  // x^n, n >= 0 small (same special cases as teqp::powi, types.hpp:138-170)
    switch (n) {
        case 0: return 1.0;
        case 1: return x;
        case 2: return x * x;
        case 3: return x * x * x;
        case 4: { const double y = x * x; return y * y; }
    }
    double y = x, r = 1.0;
    while (n > 0) {
        if (n & 1) r *= y;
        n >>= 1;
        if (n) y *= y;
    }
    return r;
}


// One state point.  W is this thread's column of the weight staging area (element w at W[w*ws]); on entry W[0..nc) holds the
// mole fractions.  On exit A[n][m] = tau^n delta^m d^(n+m) alphar / dtau^n ddelta^m / (n! m!).
template<int NT, int ND>
TQ_HD void gerg_point(const GergProgram& P, const double T, const double rho, double* W, const int ws, double (&A)[NT + 1][ND + 1]) {
    // ---- reducing functions (GERG.hpp:179-205), same summation order as the reference --------------------------
    double s1T = 0.0, s1V = 0.0, s2T = 0.0, s2V = 0.0;
    {
        int pidx = 0;
        for (int i = 0; i < P.nc; ++i) {
            const double xi = W[i * ws];
            const double xi2 = xi * xi;
            s1T += xi2 * P.Tc[i];
            s1V += xi2 * P.vc[i];
            for (int j = i + 1; j < P.nc; ++j, ++pidx) {
                const double xj = W[j * ws];
                const double num = xi * xj * (xi + xj);
                const double dT = P.pair[pidx].beta2T * xi + xj;
                const double dV = P.pair[pidx].beta2V * xi + xj;
                if (dT != 0.0) s2T += num / dT * P.pair[pidx].YT2;
                if (dV != 0.0) s2V += num / dV * P.pair[pidx].YV2;
            }
        }
    }
    const double Tr = s1T + s2T;
    const double Yv = s1V + s2V;  // 1/rho_r
    const double tau = Tr / T;
    const double delta = rho * Yv;
    const double lntau = log(tau);
    for (int q = 0; q < P.ndep; ++q) W[(P.nc + q) * ws] = W[P.dep_i[q] * ws] * W[P.dep_j[q] * ws] * P.depF[q];

    // ---- factorised term sum ------------------------------------------------------------------------------------
#pragma unroll
    for (int n = 0; n <= NT; ++n)
#pragma unroll
        for (int m = 0; m <= ND; ++m) A[n][m] = 0.0;

    int tq = 0, term = 0;
    for (int g = 0; g < P.ngroup; ++g) {
        const GergGroup& gr = P.group[g];
        // G[m] = delta^m/m! d^m/ddelta^m [ delta^d exp(h(delta)) ]
        double G[ND + 1];
        const double dd = ipow(delta, gr.d);
        if (gr.kind == GG_POLY) {
            G[0] = dd;
#pragma unroll
            for (int m = 1; m <= ND; ++m) G[m] = gr.bd[m - 1] * dd;
        } else {
            double h[ND + 1];
            if (gr.kind == GG_EXP) {
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
            double E[ND + 1];
            E[0] = exp(h[0]);
#pragma unroll
            for (int m = 1; m <= ND; ++m) {
                double s = h[1 <= ND ? 1 : 0] * E[m - 1];
#pragma unroll
                for (int j = 2; j <= m; ++j) s += (double(j) * h[j]) * E[m - j];
                E[m] = s * (1.0 / double(m));
            }
#pragma unroll
            for (int m = 0; m <= ND; ++m) {
                double s = E[m];  // binom(d,0) = 1
#pragma unroll
                for (int j = 1; j <= m; ++j) s += gr.bd[j - 1] * E[m - j];
                G[m] = s * dd;
            }
        }
        // C[n] = tau^n/n! d^n/dtau^n sum_t c_t tau^t
        double C[NT + 1];
#pragma unroll
        for (int n = 0; n <= NT; ++n) C[n] = 0.0;
        const int tq_end = gr.tq_end;
        for (; tq < tq_end; ++tq) {
            const double t = P.tq[tq].t;
            const int term_end = P.tq[tq].term_end;
            double c0 = 0.0, c1 = 0.0;
            for (; term + 1 < term_end; term += 2) {
                c0 = fma(W[P.term_w[term] * ws], P.term_n[term], c0);
                c1 = fma(W[P.term_w[term + 1] * ws], P.term_n[term + 1], c1);
            }
            if (term < term_end) { c0 = fma(W[P.term_w[term] * ws], P.term_n[term], c0); ++term; }
            double v = (c0 + c1) * exp(t * lntau);
            C[0] += v;
#pragma unroll
            for (int n = 1; n <= NT; ++n) {
                v *= (t - double(n - 1)) * (1.0 / double(n));
                C[n] += v;
            }
        }
#pragma unroll
        for (int n = 0; n <= NT; ++n)
#pragma unroll
            for (int m = 0; m <= ND; ++m) A[n][m] = fma(C[n], G[m], A[n][m]);
    }
}

}  // namespace tq
