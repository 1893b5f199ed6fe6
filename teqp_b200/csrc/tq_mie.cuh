// SAFT-VR-Mie (Lafitte et al., J. Chem. Phys. 139 (2013) 154504), monomer + chain, written once over jet types.
//
// Reference arithmetic: SAFTVRMieChainContributionTerms  /root/reference/include/teqp/models/saftvrmie.hpp
//   pair tables :106-237, Barker-Henderson diameter d_ii(T) :296-376 (cut-off by <= 4 Halley steps, then 10-point Gauss-Legendre),
//   get_core_calcs :395-639, K_HS / a_HS / Bhat / a1Shat helpers :641-770, SAFTVRMieNonpolarMixture::alphar :906-915.
//
// What is different here (B200-first, same mathematics):
//   * x0_ij = sigma_ij / d_ij depends on temperature ONLY, so all its powers are kept in the temperature jet type: in a density
//     sweep (get_Ar0n) they are plain doubles, where the reference promotes them to the full derivative type;
//   * the ~40 pow(x0, lambda) per pair of the reference collapse to TWO exponentials: with L = ln x0,
//     x0^la = e^(la L), x0^lr = e^(lr L), and x0^(2la), x0^(2lr), x0^(la+lr), x0^(3-l), x0^(4-l) follow by products and quotients;
//   * the coefficients of I and J in Bhat and rho dBhat/drho depend on zeta_x only and are formed once per state point;
//   * ln g_ii^Mie = (k0 + k1 x0 + k2 x0^2 + k3 x0^3) + (beta eps g1 + (beta eps)^2 g2)/g_d^HS  -- no exp-then-log round trip.
// Pair tables are built on the host (tq_mie.cu) exactly as the reference builds them.
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kMieMaxC = 6;
constexpr int kMieMaxPair = kMieMaxC * (kMieMaxC + 1) / 2;

struct MiePairData {           // pair (i <= j)
    double la, lr;             // lambda_a,ij, lambda_r,ij  (Eq. A48)
    double C, eps, sigma3;     // C_ij (A3), epsilon_ij/k (A55 or Lorentz-Berthelot), sigma_ij^3 / A^3
    double sigma;              // sigma_ij / A
    double c[5][4];            // Eq. A18 coefficients for lambda = la, lr, 2la, 2lr, la+lr
    double f[6];               // f_1..f_6(alpha_ij)  (A26)
    double g7;                 // phi_7,0 (1 - tanh(phi_7,1 (phi_7,2 - alpha_ii)))  (A37), used for i == j
    double pad_;
};

struct MieParams {
    int nc, pad_;
    double m[kMieMaxC], sigma[kMieMaxC], eps[kMieMaxC], la[kMieMaxC], lr[kMieMaxC];  // sigma / A, eps/k / K
    double Cii[kMieMaxC], epsii[kMieMaxC];                                            // C_ii, epsilon_ii (with 1 - k_ii)
    MiePairData pair[kMieMaxPair];                                                    // packed upper triangle, row-major
    double phi73, phi74;                                                              // phi_7,3 and phi_7,4 of A37
    double R;
};

TQ_HD int mie_pair_index(int i, int j, int n) { return i * n - i * (i - 1) / 2 + (j - i); }  // i <= j

// 10-point Gauss-Legendre nodes and weights in the reference's order (/root/reference/include/teqp/math/quadrature.hpp:44-48)
TQ_HD double mie_gl_x(int q) {
    const double x[10] = {-0.1488743389816312, 0.1488743389816312, -0.4333953941292472, 0.4333953941292472, -0.6794095682990244,
                          0.6794095682990244, -0.8650633666889845, 0.8650633666889845, -0.9739065285171717, 0.9739065285171717};
    return x[q];
}
TQ_HD double mie_gl_w(int q) {
    const double w[10] = {0.2955242247147529, 0.2955242247147529, 0.2692667193099963, 0.2692667193099963, 0.2190863625159820,
                          0.2190863625159820, 0.1494513491505806, 0.1494513491505806, 0.0666713443086881, 0.0666713443086881};
    return w[q];
}

// Barker-Henderson diameter d_ii(T) / A  (saftvrmie.hpp:296-376).  Failure cases that make the reference throw IterationFailure
// (cut-off j < 1, d > sigma) give NaN here.
template<class TT>
TQ_HD TT mie_dii(const MieParams& p, const int i, const TT& T, const TT& invT) {
    const double EPS = 2.220446049250313e-16;
    const double lnEPS = -36.043653389117154;  // log(EPS)
    const double lr = p.lr[i], la = p.la[i];
    const double kappa = p.Cii[i] * p.epsii[i];
    const TT K = tlog(T * (-lnEPS / kappa));
    TT j = texp(K / lr);
    for (int counter = 0; counter <= 3; ++counter) {
        const TT jlr = tpow(j, lr), jla = tpow(j, la);
        const TT ij = trecip(j);
        const TT u = (jlr - jla) * kappa;
        const TT up = (jlr * lr - jla * la) * ij * kappa;
        const TT upp = (jlr * (lr * (lr - 1.0)) - jla * (la * (la - 1.0))) * (ij * ij) * kappa;
        const TT R = -(u * invT) - lnEPS;
        const TT Rp = -(up * invT);
        const TT Rpp = -(upp * invT);
        const TT den = 2.0 * (Rp * Rp) - R * Rpp;
        if (base(den) < EPS) break;
        j = j - (2.0 * (R * Rp)) / den;
    }
    const double sig = p.sigma[i];
    const TT rcut = sig * trecip(j);
    const TT half = (sig - rcut) * 0.5, mid = (rcut + sig) * 0.5;
    const double Ce = p.Cii[i] * p.eps[i];
    TT sum(0.0);
    for (int q = 0; q < 10; ++q) {
        const TT r = half * mie_gl_x(q) + mid;
        const TT rs = sig * trecip(r);
        const TT u = (tpow(rs, lr) - tpow(rs, la)) * Ce;
        sum = sum + (1.0 - texp(-(u * invT))) * mie_gl_w(q);
    }
    TT d = rcut + half * sum;
    if (base(j) < 1.0 || base(d) > sig) d = d * NAN;
    return d;
}

// alphar = alphar_mono + alphar_chain  (saftvrmie.hpp:395-639, 906-915)
template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> mie_alphar(const MieParams& p, const TT& T, const RT& rhomolar, const XT* x) {
    using TX = prom_t<TT, XT>;
    using RX = prom_t<RT, XT>;
    using F = prom_t<TX, RT>;
    constexpr double N_A = 6.02214076e23;
    constexpr double PI = 3.141592653589793238462643383279502884;
    constexpr double pi6 = PI / 6.0;
    const int N = p.nc;
    const TT invT = trecip(T);

    TT d[MAXN];
    XT xs[MAXN];
    XT mbar(0.0);
    for (int i = 0; i < N; ++i) {
        d[i] = mie_dii(p, i, T, invT);
        mbar = mbar + x[i] * p.m[i];
    }
    const XT imbar = trecip(mbar);
    for (int i = 0; i < N; ++i) xs[i] = (x[i] * p.m[i]) * imbar;
    const RX rhos = (rhomolar * (N_A / 1e30)) * mbar;  // segments / A^3

    TX D[4];
    F zeta[4];
    {
        TX s0(0.0), s1(0.0), s2(0.0), s3(0.0);
        for (int i = 0; i < N; ++i) {
            const TT d2 = d[i] * d[i];
            s0 = s0 + embed<TX>(xs[i]);
            s1 = s1 + xs[i] * d[i];
            s2 = s2 + xs[i] * d2;
            s3 = s3 + xs[i] * (d2 * d[i]);
        }
        D[0] = s0 * pi6; D[1] = s1 * pi6; D[2] = s2 * pi6; D[3] = s3 * pi6;
#pragma unroll
        for (int l = 0; l < 4; ++l) zeta[l] = D[l] * rhos;
    }
    TX szx(0.0);
    XT szb(0.0);
    for (int i = 0; i < N; ++i) {
        for (int j = i; j < N; ++j) {
            const MiePairData& pr = p.pair[mie_pair_index(i, j, N)];
            const double fac = (i == j) ? 1.0 : 2.0;
            const TT dij = (i == j) ? d[i] : (d[i] + d[j]) * 0.5;
            const XT xx = (xs[i] * xs[j]) * fac;
            szx = szx + xx * (dij * dij * dij);
            szb = szb + xx * pr.sigma3;
        }
    }
    const F zx = (szx * pi6) * rhos;     // zeta_x, Eq. A13
    const RX zb = (szb * pi6) * rhos;    // zeta_x_bar, Eq. A23
    const RX zb2 = zb * zb;
    const RX zb5 = zb2 * zb2 * zb;
    const RX zb8 = zb5 * zb2 * zb;
    const F zx2 = zx * zx, zx3 = zx2 * zx, zx4 = zx2 * zx2;
    const F omz = 1.0 - zx;
    const F iomz = trecip(omz);
    const F iX2 = iomz * iomz, iX3 = iX2 * iomz, iX4 = iX2 * iX2;
    // g_d^HS exponent coefficients, Eq. A30-A33
    const F k0 = -tlog(omz) + (42.0 * zx - 39.0 * zx2 + 9.0 * zx3 - 2.0 * zx4) * iX3 * (1.0 / 6.0);
    const F k1 = (zx4 + 6.0 * zx2 - 12.0 * zx) * iX3 * 0.5;
    const F k2 = zx2 * iX2 * (-3.0 / 8.0);
    const F k3 = (3.0 * zx2 + 3.0 * zx - zx4) * iX3 * (1.0 / 6.0);
    // K_HS (A21) and rho_s dK_HS/drho_s
    const F kden = 1.0 + 4.0 * zx + 4.0 * zx2 - 4.0 * zx3 + zx4;
    const F ikden = trecip(kden);
    const F omz2 = omz * omz;
    const F KHS = (omz2 * omz2) * ikden;
    const F rdK = ((omz2 * omz) * (zx2 - 5.0 * zx - 2.0)) * (ikden * ikden) * zx * 4.0;  // -4 (zx-1)^3 = 4 (1-zx)^3
    // Bhat = cI I - cJ J ;  zeta_x dBhat/dzeta_x = eI I + eJ J
    const F cI = (1.0 - 0.5 * zx) * iX3;
    const F cJ = (zx * (1.0 + zx)) * iX3 * 4.5;
    const F eI = (zx * (5.0 - 2.0 * zx)) * iX4 * 0.5;
    const F eJ = (zx * (zx2 + 4.0 * zx + 1.0)) * iX4 * (-4.5);

    F a1(0.0), a2(0.0);
    RX a3(0.0);
    F chain(0.0);
    for (int i = 0; i < N; ++i) {
        for (int j = i; j < N; ++j) {
            const MiePairData& pr = p.pair[mie_pair_index(i, j, N)];
            const bool diag = (i == j);
            const TT dij = diag ? d[i] : (d[i] + d[j]) * 0.5;
            const TT x0 = pr.sigma * trecip(dij);
            const TT L = tlog(x0);
            const TT x03 = x0 * x0 * x0;
            const TT dij3 = dij * dij * dij;
            // powers x0^lambda for lambda = la, lr, 2la, 2lr, la+lr
            TT pw[5];
            pw[0] = texp(L * pr.la);
            pw[1] = texp(L * pr.lr);
            pw[2] = pw[0] * pw[0];
            pw[3] = pw[1] * pw[1];
            pw[4] = pw[0] * pw[1];
            const double lam[5] = {pr.la, pr.lr, 2.0 * pr.la, 2.0 * pr.lr, pr.la + pr.lr};
            F ot[5];    // x0^lambda (Bhat + a1Shat)
            F dot_[5];  // x0^lambda (rho d a1Shat/drho + rho dBhat/drho)   (diagonal pairs only)
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const double l = lam[q];
                const TT q3 = x03 * trecip(pw[q]);  // x0^(3-lambda)
                const TT I = (q3 - 1.0) * (-1.0 / (l - 3.0));                                                   // Eq. A14
                const TT J = ((q3 * x0) * (l - 3.0) - q3 * (l - 4.0) - 1.0) * (-1.0 / ((l - 3.0) * (l - 4.0)));   // Eq. A15
                const F zeff = pr.c[q][0] * zx + pr.c[q][1] * zx2 + pr.c[q][2] * zx3 + pr.c[q][3] * zx4;
                const F u = trecip(1.0 - zeff);
                const F u3 = u * u * u;
                const F a1S = ((1.0 - 0.5 * zeff) * u3) * (-1.0 / (l - 3.0));                                     // Eq. A16 / (2 pi eps d^3 rho_s)
                const F Bh = cI * I - cJ * J;
                ot[q] = pw[q] * (Bh + a1S);
                if (diag) {
                    const F dzeff = pr.c[q][0] + (2.0 * pr.c[q][1]) * zx + (3.0 * pr.c[q][2]) * zx2 + (4.0 * pr.c[q][3]) * zx3;
                    const F da1S = a1S + (((2.0 * zeff - 5.0) * dzeff) * (u3 * u)) * zx * (0.5 / (l - 3.0));
                    const F dB = Bh + (eI * I + eJ * J);
                    dot_[q] = pw[q] * (da1S + dB);
                }
            }
            const double fac = diag ? 1.0 : 2.0;
            const XT xx = (xs[i] * xs[j]) * fac;
            const RX kap = rhos * (2.0 * PI * pr.eps);                       // 2 pi rho_s eps_ij (times d_ij^3 below)
            const F a1ij = (kap * dij3) * ((ot[0] - ot[1]) * pr.C);
            a1 = a1 + xx * a1ij;
            const RX chi = pr.f[0] * zb + pr.f[1] * zb5 + pr.f[2] * zb8;
            const F s2 = ot[2] - 2.0 * ot[4] + ot[3];
            const F a2ij = (KHS * (1.0 + chi)) * ((kap * dij3) * s2) * (0.5 * pr.eps * pr.C * pr.C);
            a2 = a2 + xx * a2ij;
            const RX a3ij = (zb * texp(pr.f[4] * zb + pr.f[5] * zb2)) * (-(pr.eps * pr.eps * pr.eps) * pr.f[3]);
            a3 = a3 + xx * a3ij;
            if (diag) {
                const F lng = k0 + k1 * x0 + k2 * (x0 * x0) + k3 * x03;   // ln g_d^HS(sigma), Eq. A29
                const F gd = texp(lng);
                const F g1 = (3.0 * pr.C) * (dot_[0] - dot_[1]) - pr.C * (pr.la * ot[0] - pr.lr * ot[1]);
                const F g2nd = (KHS * (pr.la * ot[2] - (pr.la + pr.lr) * ot[4] + pr.lr * ot[3])) * (-(pr.C * pr.C));
                const F da2 = (rdK * s2 + KHS * (dot_[2] - 2.0 * dot_[4] + dot_[3])) * (0.5 * pr.C * pr.C);
                const F g2mca = 3.0 * da2 + g2nd;
                const TT be = invT * pr.eps;
                const TT theta = texp(be) - 1.0;
                const RX gz = (zb * texp(p.phi73 * zb + p.phi74 * zb2)) * pr.g7;
                const F g2 = (1.0 + gz * theta) * g2mca;
                const F lngMie = lng + (be * g1 + (be * be) * g2) * trecip(gd);
                chain = chain - (x[i] * (p.m[i] - 1.0)) * lngMie;
            }
        }
    }
    // hard-sphere term (A6) with the rho_s == 0 series branch of the reference (saftvrmie.hpp:663-698)
    F ahs;
    const F Ups = 1.0 - zeta[3];
    if (base(rhos) == 0.0) {
        ahs = 3.0 * (D[1] / D[0]) * zeta[2] / Ups + (D[2] * D[2]) * zeta[2] / (D[3] * D[0] * (Ups * Ups)) - tlog(Ups) +
              ((D[2] * D[2] * D[2]) / (D[3] * D[3] * D[0])) * tlog(Ups);
    } else {
        const F z23 = zeta[2] * zeta[2] * zeta[2];
        ahs = (3.0 * zeta[1] * zeta[2] / Ups + z23 / (zeta[3] * (Ups * Ups)) + (z23 / (zeta[3] * zeta[3]) - zeta[0]) * tlog(Ups)) *
              trecip(rhos * (PI / 6.0));
    }
    const TT invT2 = invT * invT;
    const F mono = mbar * (ahs + a1 * invT + a2 * invT2 + a3 * (invT2 * invT));
    return mono + chain;
}

struct MieDev {
    using Params = MieParams;
    static constexpr int kMaxC = kMieMaxC;
    static constexpr bool kNoSmallVariants = true;  // one kernel variant (runtime component count): keeps the build time down
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return mie_alphar<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
