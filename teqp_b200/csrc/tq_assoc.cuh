// Association contribution (Wertheim TPT1 with site-type index compression) and the genericSAFT sum nonpolar + association,
// written once over jet types.
//
// Reference arithmetic: teqp::association::Association  /root/reference/include/teqp/models/association/association.hpp
//   site indexing :86-111, counting matrix D_IJ :117-151, Delta_IJ (CR1 and Dufal rules) :274-371, successive substitution
//   :373-430 (damped fixed point on the FULL derivative type, convergence tested on base values, the iterate BEFORE the one that
//   passes the test is returned), alphar :436-461;  GenericSAFT::alphar  /root/reference/include/teqp/models/saft/genericsaft.hpp:65-73.
// Mixing rules for the Dufal data: association_types.hpp:57-74 (evaluated on the host, tq_gsaft.cu).
//
// The fixed point is reproduced step for step (same damping, same test, same returned iterate) because its derivative
// coefficients lag the base value by the same number of contractions as the reference's.  What is precomputed on the host:
// the site tables, and the 66 coefficients a_ij(lambda_r) = sum_k b_ijk lambda_r^k of the Dufal kernel for each component pair.
#pragma once
#include "tq_jet.cuh"
#include "tq_mie.cuh"

namespace tq {

constexpr int kAssocMaxS = 6;                                // unique (component, site type) groups in the mixture
constexpr int kAssocMaxC = kMieMaxC;
constexpr int kAssocMaxPair = kAssocMaxC * (kAssocMaxC + 1) / 2;
constexpr int kDufalNA = 66;                                 // (i, j) with i + j <= 10

struct AssocParams {
    int nsite, rule;              // rule: 0 = CR1, 1 = Dufal
    int max_iters, explicit_ok;   // explicit_ok: the closed form of association.hpp:403-418 applies (pure fluid, two site types)
    double alpha, rtol, atol;
    int comp[kAssocMaxS];         // component of each site group
    int nsites_of[kAssocMaxC];    // number of site groups on each component (groups of one component are consecutive)
    int first_site[kAssocMaxC];
    double count[kAssocMaxS];     // multiplicity of each site group
    double D[kAssocMaxS * kAssocMaxS];
    // per component pair (full matrices, [i * kAssocMaxC + j])
    double pref[kAssocMaxC * kAssocMaxC];    // CR1: b_ij beta_ij ; Dufal: K_HB,ij / m^3
    double eps1[kAssocMaxC * kAssocMaxC];    // CR1: epsilon_ij / R / K ; Dufal: epsilon_HB,ij / k / K
    double epsk[kAssocMaxC * kAssocMaxC];    // Dufal: epsilon_ij / k / K (T* = T / epsk)
    double sigma3[kAssocMaxC * kAssocMaxC];  // Dufal: ((sigma_i + sigma_j)/2)^3 / m^3
    double b[kAssocMaxC];                    // CR1: covolumes / m^3/mol
    double a[kAssocMaxPair][kDufalNA];       // Dufal: a_ij(lambda_r,pair), packed i-major (i = power of rho*, j = power of T*)
};

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> assoc_alphar(const AssocParams& p, const int N, const TT& T, const RT& rhomolar, const XT* x) {
    using RX = prom_t<RT, XT>;
    using F = prom_t<prom_t<TT, XT>, RT>;
    constexpr double N_A = 6.02214076e23;
    const int S = p.nsite;
    if (S == 0) return F(0.0);
    const TT invT = trecip(T);

    // ---- Delta for each pair of components that carry sites (association.hpp:274-371)
    RX g(0.0), rhostar(0.0);
    if (p.rule == 0) {
        XT bmix(0.0);
        for (int i = 0; i < N; ++i) bmix = bmix + x[i] * p.b[i];
        const RX eta = (bmix * rhomolar) * 0.25;
        const RX ome = 1.0 - eta;
        g = (2.0 - eta) / (2.0 * (ome * ome * ome));  // Carnahan-Starling
    } else {
        XT s3(0.0);
        for (int i = 0; i < N; ++i)
            for (int j = 0; j < N; ++j) s3 = s3 + (x[i] * x[j]) * p.sigma3[i * kAssocMaxC + j];
        rhostar = (rhomolar * N_A) * s3;
    }
    // Dufal rule, reference quirk (measured here): DufalData keeps its pair tables in 1-D Eigen arrays that
    // are indexed (i, j); Eigen ignores the column index of a vector, so a read of pair (i, j) returns what was stored last for
    // row i -- the value of pair (i, N-1).  Delta_IJ therefore depends on the ROW component only.  Reproduced on purpose.
    F Delta[kAssocMaxC * kAssocMaxC];
    for (int i = 0; i < N; ++i) {
        if (p.nsites_of[i] == 0) continue;
        if (p.rule == 0) {
            for (int j = i; j < N; ++j) {
                if (p.nsites_of[j] == 0) continue;
                const int ij = i * kAssocMaxC + j;
                const F Dl = (g * p.pref[ij]) * (texp(invT * p.eps1[ij]) - 1.0) * (1.0 / N_A);
                Delta[ij] = Dl;
                Delta[j * kAssocMaxC + i] = Dl;
            }
        } else {
            const int ij = i * kAssocMaxC + (N - 1);
            const TT Tstar = T * (1.0 / p.epsk[ij]);
            const double* a = p.a[mie_pair_index(i, N - 1, N)];
            F summer(0.0);
            RX ri(1.0);
            int q = 0;
            for (int ii = 0; ii <= 10; ++ii) {
                TT tj(1.0);
                for (int jj = 0; ii + jj <= 10; ++jj, ++q) {
                    summer = summer + (ri * a[q]) * tj;
                    tj = tj * Tstar;
                }
                ri = ri * rhostar;
            }
            const F Dl = ((texp(invT * p.eps1[ij]) - 1.0) * p.pref[ij]) * summer;
            for (int j = 0; j < N; ++j) Delta[i * kAssocMaxC + j] = Dl;
        }
    }
    // ---- rDDX(I, J) = rho N_A Delta_IJ D_IJ x_comp(J)
    const RT rhoN = rhomolar * N_A;
    F M[kAssocMaxS * kAssocMaxS];
    for (int I = 0; I < S; ++I)
        for (int J = 0; J < S; ++J)
            M[I * kAssocMaxS + J] = (rhoN * (Delta[p.comp[I] * kAssocMaxC + p.comp[J]] * p.D[I * kAssocMaxS + J])) * x[p.comp[J]];

    F X[kAssocMaxS], Xn[kAssocMaxS];
    for (int I = 0; I < S; ++I) X[I] = F(1.0);
    bool solved = false;
    if (p.explicit_ok && base(M[0]) == 0.0 && base(M[kAssocMaxS + 1]) == 0.0) {
        // pure fluid with two site types that only bond to each other: closed form (association.hpp:403-418)
        const F Dl = Delta[0];
        const F kA = (rhoN * p.count[0]) * Dl, kB = (rhoN * p.count[1]) * Dl;
        const F disc = tsqrt(kA * kA - 2.0 * (kA * kB) + 2.0 * kA + kB * kB + 2.0 * kB + 1.0);
        const F i2kA = trecip(2.0 * kA);
        const F XA1 = (kA - kB - disc - 1.0) * i2kA, XA2 = (kA - kB + disc - 1.0) * i2kA;
        if (base(XA1) < 0 && base(XA2) > 0) X[0] = XA2;
        else if (base(XA1) > 0 && base(XA2) < 0) X[0] = XA1;
        X[1] = trecip(1.0 + kA * X[0]);
        solved = true;
    }
    if (!solved) {
        for (int counter = 0; counter < p.max_iters; ++counter) {
            bool all_ok = true;
            for (int I = 0; I < S; ++I) {
                F s = M[I * kAssocMaxS] * X[0];
                for (int J = 1; J < S; ++J) s = s + M[I * kAssocMaxS + J] * X[J];
                Xn[I] = p.alpha * X[I] + (1.0 - p.alpha) * trecip(1.0 + s);
                const double diff = fabs(base(Xn[I]) - base(X[I]));
                const double tol = p.rtol * base(Xn[I]) + p.atol;
                if (!(diff < tol)) all_ok = false;
            }
            if (all_ok) break;
            for (int I = 0; I < S; ++I) X[I] = Xn[I];
        }
    }
    // ---- alphar = sum_i x_i sum_sites n_s (ln X - X/2 + 1/2)   (association.hpp:448-459)
    F alpha(0.0);
    for (int i = 0; i < N; ++i) {
        if (p.nsites_of[i] == 0) continue;
        F summer(0.0);
        for (int s = 0; s < p.nsites_of[i]; ++s) {
            const int I = p.first_site[i] + s;
            summer = summer + (tlog(X[I]) - X[I] * 0.5 + 0.5) * p.count[I];
        }
        alpha = alpha + x[i] * summer;
    }
    return alpha;
}

// genericSAFT = SAFT-VR-Mie (nonpolar) [+ association]; also serves the plain "SAFT-VR-Mie" kind (has_assoc = 0)
struct GsaftParams {
    MieParams mie;
    AssocParams assoc;
    int has_assoc, pad_;
};

struct GsaftDev {
    using Params = GsaftParams;
    static constexpr int kMaxC = kMieMaxC;
    static constexpr int kCrossOrder = 1;  // order of x_i and x_j in the get_ATrhoXiXj jet (2: full table; 1: mixed entries only)
    static constexpr bool kNoSmallVariants = true;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        auto a = mie_alphar<MAXN>(p.mie, T, rho, x);
        if (p.has_assoc) a = a + assoc_alphar<MAXN>(p.assoc, p.mie.nc, T, rho, x);
        return a;
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.mie.R; }
};

}  // namespace tq
