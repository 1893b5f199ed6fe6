// PC-SAFT (Gross & Sadowski, Ind. Eng. Chem. Res. 40 (2001) 1244), hard chain + dispersion, written once over jet types.
//
// Reference arithmetic: PCSAFTHardChainContribution::eval  /root/reference/include/teqp/models/pcsaft.hpp:204-294,
//   C1 :87-93, get_alphar_hs :104-148 (zeta_3 == 0 series branch :134-141), gij_HS :150-161, PCSAFTMixture::alphar :433-451.
// Universal constants a, b: published tables (G&S 2001 Table 1; Liang et al. IECR 2012 / 2014), as selected by "ab" in the JSON
// (/root/reference/include/teqp/models/pcsaft.hpp:460-476, values in /root/reference/src/data/PCSAFT.cpp:8-44).
#pragma once
#include "tq_jet.cuh"

namespace tq {

constexpr int kPcsaftMaxC = 8;

struct PcsaftParams {
    int nc;
    int pad_;
    double m[kPcsaftMaxC], sigma[kPcsaftMaxC], eps[kPcsaftMaxC];   // segments, sigma / Angstrom, epsilon/k / K
    double kmat[kPcsaftMaxC * kPcsaftMaxC];
    double a[3][7], b[3][7];
    double R;
};

template<int MAXN, class TT, class RT, class XT>
TQ_HD prom_t<prom_t<TT, RT>, XT> pcsaft_alphar(const PcsaftParams& p, const TT& T, const RT& rhomolar, const XT* x) {
    using TX = prom_t<TT, XT>;
    using F = prom_t<TX, RT>;
    constexpr double N_A = 6.02214076e23;
    constexpr double PI = 3.141592653589793238462643383279502884;
    const int N = p.nc;
    TT d[MAXN];
    TX m2es3(0.0), m2e2s3(0.0);
    const TT invT = trecip(T);
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i) {
        if (i < N) {
            d[i] = p.sigma[i] * (1.0 - 0.12 * texp((-3.0 * p.eps[i]) * invT));
#pragma unroll(MAXN <= 4 ? MAXN : 1)
            for (int j = 0; j < MAXN; ++j) {
                if (j < N) {
                    const double sigma_ij = 0.5 * p.sigma[i] + 0.5 * p.sigma[j];
                    const double eij = sqrt(p.eps[i] * p.eps[j]) * (1.0 - p.kmat[i * N + j]);
                    const double s3 = sigma_ij * sigma_ij * sigma_ij;
                    const TT ekT = eij * invT;
                    const XT w = x[i] * x[j] * (p.m[i] * p.m[j]);
                    m2es3 = m2es3 + w * (ekT * s3);
                    m2e2s3 = m2e2s3 + w * ((ekT * ekT) * s3);
                }
            }
        }
    }
    XT mbar(0.0);
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i)
        if (i < N) mbar = mbar + x[i] * p.m[i];

    const RT rho_A3 = rhomolar * (N_A * 1e-30);
    const double pi6 = PI / 6.0;
    TX D[4];
    F zeta[4];
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        TX s(0.0);
#pragma unroll(MAXN <= 4 ? MAXN : 1)
        for (int i = 0; i < MAXN; ++i) {
            if (i < N) {
                const XT xm = x[i] * p.m[i];
                if (n == 0) s = s + embed<TX>(xm);
                else if (n == 1) s = s + xm * d[i];
                else if (n == 2) s = s + xm * (d[i] * d[i]);
                else s = s + xm * (d[i] * d[i] * d[i]);
            }
        }
        D[n] = s * pi6;
        zeta[n] = D[n] * rho_A3;
    }
    const F eta = zeta[3];
    const XT mm1 = (mbar - 1.0) / mbar;
    const XT mm2 = mm1 * ((mbar - 2.0) / mbar);
    // I1, I2 = sum_k (a_k(mbar), b_k(mbar)) eta^k
    F I1(0.0), I2(0.0), etak(1.0);
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const XT ak = p.a[0][k] + mm1 * p.a[1][k] + mm2 * p.a[2][k];
        const XT bk = p.b[0][k] + mm1 * p.b[1][k] + mm2 * p.b[2][k];
        I1 = I1 + ak * etak;
        I2 = I2 + bk * etak;
        if (k < 6) etak = etak * eta;
    }
    // hard sphere (Eq. A.6), with the zeta_3 == 0 series branch of the reference
    const F Ups = 1.0 - zeta[3];
    F ahs;
    if (base(zeta[3]) == 0.0) {
        ahs = 3.0 * (D[1] / D[0]) * zeta[2] / Ups + (D[2] * D[2]) * zeta[2] / (D[3] * D[0] * (Ups * Ups)) - tlog(Ups) +
              ((D[2] * D[2] * D[2]) / (D[3] * D[3] * D[0])) * tlog(Ups);
    } else {
        const F z23 = zeta[2] * zeta[2] * zeta[2];
        ahs = (3.0 * zeta[1] * zeta[2] / Ups + z23 / zeta[3] / Ups / Ups + (z23 / (zeta[3] * zeta[3]) - zeta[0]) * tlog(Ups)) / zeta[0];
    }
    // chain term: sum x_i (m_i - 1) ln g_ii^hs (Eq. A.4, A.7)
    F chain(0.0);
    const F iU = trecip(Ups);
    const F iU2 = iU * iU;
    const F iU3 = iU2 * iU;
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i) {
        if (i < N) {
            const TT dh = (d[i] * d[i]) / (d[i] + d[i]);
            const F g = iU + dh * 3.0 * zeta[2] * iU2 + (dh * dh) * 2.0 * (zeta[2] * zeta[2]) * iU3;
            chain = chain + (x[i] * (p.m[i] - 1.0)) * tlog(g);
        }
    }
    const F alphar_hc = mbar * ahs - chain;
    // dispersion (Eq. A.10), C1 from Eq. A.11
    const F oe = 1.0 - eta;
    const F oe2 = oe * oe;
    const F te = (1.0 - eta) * (2.0 - eta);
    const F eta2 = eta * eta;
    const F C1 = trecip(1.0 + mbar * (8.0 * eta - 2.0 * eta2) / (oe2 * oe2) +
                        (1.0 - mbar) * (20.0 * eta - 27.0 * eta2 + 12.0 * (eta2 * eta) - 2.0 * (eta2 * eta2)) / (te * te));
    const F alphar_disp = (-2.0 * PI) * rho_A3 * I1 * m2es3 - PI * rho_A3 * mbar * C1 * I2 * m2e2s3;
    return alphar_hc + alphar_disp;
}

struct PcsaftDev {
    using Params = PcsaftParams;
    static constexpr int kMaxC = kPcsaftMaxC;
    static constexpr int kCrossOrder = 1;  // order of x_i and x_j in the get_ATrhoXiXj jet (2: full table; 1: mixed entries only)
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        return pcsaft_alphar<MAXN>(p, T, rho, x);
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.R; }
};

}  // namespace tq
