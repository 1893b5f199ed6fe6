// GERG-2004 / GERG-2008: the isochoric family -- Psi^r(T, rhovec) = alphar * R * T * rho, its gradient and Hessian in the
// molar concentrations, and ln(phi) -- by an ANALYTIC chain rule over the factorised term program, one state point per thread.
//
// Reference: /root/reference/include/teqp/derivs.hpp:1031-1084 (build_Psir_Hessian_autodiff / build_Psir_fgradHessian_autodiff:
// Nc(Nc+1)/2 second-order dual sweeps of the whole model), :1115-1124 (gradient: Nc first-order sweeps), :1208-1230
// (ln phi_i = dPsir/drho_i /(RT) - ln(1 + Ar01)).  For the 21-component mixture that is 231 sweeps through 436 exponential
// terms per state point.  Here the structure of the model is used instead:
//
//   * the reducing functions are homogeneous of degree 2 in x, so with x = rhovec/rho
//         tau = Y_T(rhovec) / (rho^2 T),   delta = Y_V(rhovec) / rho          (GERG.hpp:179-205 evaluated at rhovec)
//   * Phi := rho * alphar = sum_w w * a_w(tau, delta),  weights w = rho_i (pure fluids) and F_ij rho_i rho_j / rho (departure)
//   * with s = ln tau, r = ln delta:
//         dPhi/drho_k        = Phi_s t_k + Phi_r d_k + sum_w a_w w_k
//         d2Phi/drho_k drho_l = Phi_ss t_k t_l + Phi_sr (t_k d_l + t_l d_k) + Phi_rr d_k d_l + Phi_s t_kl + Phi_r d_kl
//                               + sum_w (a_w,s t_l + a_w,r d_l) w_k + sum_w (a_w,s t_k + a_w,r d_k) w_l + sum_w a_w w_kl
//     t_k = dln(tau)/drho_k, d_k = dln(delta)/drho_k, t_kl, d_kl their derivatives.
//
// One pass over the term program gives the totals Phi_{s^n r^m} (n, m <= 2) and, per weight, a_w, a_w,s, a_w,r (three FMAs per
// term); one pass over the Nc(Nc-1)/2 reducing pairs gives gradients and diagonal second derivatives of Y_T, Y_V; the Hessian
// is then assembled pair by pair (the mixed second derivative of a pair's reducing term is recomputed where it is used) and
// written straight to global memory.  Cost ~ 4e4 flops per point for Nc = 21, against ~6e7 for the sweep formulation.
//
// Per-thread state lives in a column of the CTA's shared staging area (row stride WS = CTA size), rows:
//   [0,nw) weights | [nw,2nw) a_w | [2nw,3nw) a_w,s | [3nw,4nw) a_w,r | then Nc-vectors t, d, diagT, diagV, E, P, Q
#pragma once
#include "tq_common.hpp"
#include "tq_gerg.cuh"

#ifndef TQ_ISO_UNROLL
#define TQ_ISO_UNROLL 1   // unroll factor of the pair loops of passes A and C.  Measured on B200 (profiles/r02g, C5 10^6 points): 1 -> 120.5,
                          // 2 -> 123.5, 4 -> 121.7 Mpt/s: the loop bodies branch (zero denominators, beta = 1 pairs), so unrolling does not
                          // interleave iterations; one warp per scheduler stays latency-bound
#endif

namespace tq {
constexpr int kIsoUnroll = TQ_ISO_UNROLL;

// rows of the staging area.  Hessian path: weights, three per-weight arrays, the Nc-vectors dlnY_T, dlnY_V and the running diagonal
// sums; the vector E lives in rows that are free once the departure weights have been reduced (2 ndep of them), plus nc - 2 ndep
// extra rows for mixtures with few departure functions.
TQ_HD constexpr int gerg_iso_extra_rows(int nc, int nw) { return nc - 2 * (nw - nc) > 0 ? nc - 2 * (nw - nc) : 0; }
TQ_HD constexpr int gerg_iso_rows(int nc, int nw, bool want_h) { return want_h ? 4 * nw + 3 * nc + gerg_iso_extra_rows(nc, nw) : 2 * nw + 2 * nc; }

// value, first and (H) second derivatives of one reducing pair term f = ri rj (ri + rj) / (b2 ri + rj); inv = 1/(b2 ri + rj)
struct GergPairD { double f, fi, fj; };
TQ_HD GergPairD gerg_pair_first(double b2, double ri, double rj, double inv) {
    GergPairD r;
    r.f = (ri * rj) * (ri + rj) * inv;
    r.fi = fma(-r.f, b2, rj * fma(2.0, ri, rj)) * inv;
    r.fj = (ri * fma(2.0, rj, ri) - r.f) * inv;
    return r;
}

// The derived isochoric properties (IsoPropsOut; derivs.hpp:962-978, 1196-1201, 1264-1437, 1482-1494) come out of the same pass: with
//   g_i = dPhi/drho_i,  P_i = d g_i / d ln tau = Phi_ss t_i + Phi_sr d_i + sum_w a_w,s w_i,  Q_i = Phi_sr t_i + Phi_rr d_i + sum_w a_w,r w_i
// Euler's theorem for the scaling rhovec -> lambda rhovec (x, tau fixed; delta -> lambda delta) gives (H rhovec)_i = RT Q_i, so nothing has to
// be accumulated over the Hessian rows:
//   d2Psir/dT drho_i = R (g_i - P_i)                       dp/drho_i = RT (1 + Q_i)             v_i = (1 + Q_i) / (rho (1 + 2 Ar01 + Ar02))
//   d ln phi_i / dT  = -P_i / T + Ar11 / (T Z)             d ln phi_i / drho = (Q_i - (Ar01 + Ar02) / Z) / rho
//   d ln phi_i / dx_j = rho H_ij / RT - (dZ/dx_j) / Z,     dZ/dx_j = rho d2 alphar / drho dx_j = Q_j + sum_q a_q,r w_q / rho + (2 Phi_sr + Phi_rr) / rho
//   p_r = RT Phi_r,  s+ = (Phi - Phi_s) / rho,  dp/dT = R (rho + Phi_r - Phi_sr)
// (the reference forms the full Hessian by Nc (Nc + 1) / 2 second-order sweeps and multiplies it by a vector for each of these).
template<int WS, bool WANT_H>
TQ_HD void gerg_iso_point(const GergProgram& P, const Exp2K& K, const double Rgas, const double T, double* W, const IsochoricOut& o,
                          const long long k, const long long npts, const IsoPropsOut* ipp = nullptr) {
    constexpr int SH = (WS == 128) ? 0 : (WS == 64) ? 1 : (WS == 32) ? 2 : (WS == 16) ? 3 : (WS == 8) ? 4 : (WS == 4) ? 5 : (WS == 2) ? 6 : 7;
    static_assert((kGergWS >> SH) == WS, "WS must be a power of two <= kGergWS");
    constexpr int NO = WANT_H ? 2 : 1;  // order of the total-derivative table
    const int nc = P.nc, nw = P.nw;
    double* aw = W + (size_t)nw * WS;
    double* as = WANT_H ? aw + (size_t)nw * WS : aw;
    double* ar = WANT_H ? as + (size_t)nw * WS : aw;
    double* vt = WANT_H ? ar + (size_t)nw * WS : aw + (size_t)nw * WS;
    double* vd = vt + (size_t)nc * WS;
    double* hd = vd + (size_t)nc * WS;  // only touched when WANT_H: running sums of the diagonal second derivatives of Y_T, Y_V
    double* Ex = hd + (size_t)nc * WS;  // extra rows of E (mixtures with fewer than nc/2 departure functions)
    const int ndep = P.ndep;
    // E_k: rows of the departure weights (W) and of their a_w,s sums (as) are dead after the reductions; E moves in there
    auto Eref = [&](const int kk) -> double& {
        return (kk < ndep) ? W[(nc + kk) * WS] : (kk < 2 * ndep) ? as[(nc + kk - ndep) * WS] : Ex[(kk - 2 * ndep) * WS];
    };

    // ---- pass A: Y_T, Y_V at rhovec, gradients, diagonal second derivatives --------------------------------------------
    double rhotot = 0.0, YT = 0.0, YV = 0.0;
    for (int i = 0; i < nc; ++i) {
        const double ri = W[i * WS];
        rhotot += ri;
        YT = fma(ri * ri, P.Tc[i], YT);
        YV = fma(ri * ri, P.vc[i], YV);
        vt[i * WS] = 2.0 * ri * P.Tc[i];
        vd[i * WS] = 2.0 * ri * P.vc[i];
    }
    {
        int pidx = 0;
        for (int i = 0; i < nc; ++i) {
            const double ri = W[i * WS];
            double gTi = 0.0, gVi = 0.0;
#pragma unroll kIsoUnroll
            for (int j = i + 1; j < nc; ++j, ++pidx) {
                const double rj = W[j * WS];
                const GergRedPair pr = P.pair[pidx];
                if (pr.beta2T == 1.0 && pr.beta2V == 1.0) {
                    // beta_T = beta_v = 1 (154 of the 210 GERG-2008 pairs): the pair term ri rj (ri + rj) / (ri + rj) is ri rj -- gradient
                    // (rj, ri), mixed second derivative 1, no diagonal part.  The branch is uniform over the warp (same pair for every thread).
                    if (ri + rj == 0.0) continue;
                    const double pij = ri * rj;
                    YT = fma(pr.YT2, pij, YT);
                    YV = fma(pr.YV2, pij, YV);
                    gTi = fma(pr.YT2, rj, gTi);
                    gVi = fma(pr.YV2, rj, gVi);
                    vt[j * WS] = fma(pr.YT2, ri, vt[j * WS]);
                    vd[j * WS] = fma(pr.YV2, ri, vd[j * WS]);
                    continue;
                }
                const double dT = fma(pr.beta2T, ri, rj), dV = fma(pr.beta2V, ri, rj);
                const double den = dT * dV;
                if (den == 0.0) continue;  // both concentrations zero: the reference skips the pair (GERG.hpp:189)
                const double inv = fast_rcp(den);
                const double invT = inv * dV, invV = inv * dT;
                const GergPairD a = gerg_pair_first(pr.beta2T, ri, rj, invT);
                const GergPairD b = gerg_pair_first(pr.beta2V, ri, rj, invV);
                YT = fma(pr.YT2, a.f, YT);
                YV = fma(pr.YV2, b.f, YV);
                gTi = fma(pr.YT2, a.fi, gTi);
                gVi = fma(pr.YV2, b.fi, gVi);
                vt[j * WS] = fma(pr.YT2, a.fj, vt[j * WS]);
                vd[j * WS] = fma(pr.YV2, b.fj, vd[j * WS]);
            }
            vt[i * WS] += gTi;
            vd[i * WS] += gVi;
        }
    }
    const double invr = 1.0 / rhotot;
    const double iYT = 1.0 / YT, iYV = 1.0 / YV;
    const double tau = YT * (invr * invr) / T;
    const double delta = YV * invr;
    // vt, vd -> dlnY/drho_k (the -2/rho, -1/rho of ln tau, ln delta are added where they are used)
    for (int i = 0; i < nc; ++i) {
        vt[i * WS] *= iYT;
        vd[i * WS] *= iYV;
    }
    for (int q = 0; q < P.ndep; ++q) W[(nc + q) * WS] = W[P.dep_i[q] * WS] * W[P.dep_j[q] * WS] * P.depF[q] * invr;
    for (int w = 0; w < nw; ++w) {
        aw[w * WS] = 0.0;
        if (WANT_H) { as[w * WS] = 0.0; ar[w * WS] = 0.0; }
    }

    // ---- pass B: the term program: totals A[n][m] = Phi_{tau^n delta^m} tau^n delta^m / (n! m!) and per-weight sums ------
    double l2tau = log2(tau);
    const double nanfix = l2tau - l2tau;
    l2tau = (l2tau > 30.0) ? 30.0 : l2tau;
    l2tau = (l2tau < -30.0) ? -30.0 : l2tau;
    double A[NO + 1][NO + 1];
#pragma unroll
    for (int n = 0; n <= NO; ++n)
#pragma unroll
        for (int m = 0; m <= NO; ++m) A[n][m] = nanfix;
    {
        int rec = 0;
        double E[NO + 1];
        double dd = 1.0;
        for (int g = 0; g < P.ngroup; ++g) {
            const GergGroup& gr = P.group[g];
            if (gr.new_h) {
                double h[NO + 1];
                if (gr.kind == GG_POLY) {
#pragma unroll
                    for (int m = 0; m <= NO; ++m) h[m] = 0.0;
                } else if (gr.kind == GG_EXP) {
                    const double cdl = -gr.c * ipow(delta, gr.l);
                    h[0] = cdl;
#pragma unroll
                    for (int m = 1; m <= NO; ++m) h[m] = gr.bl[m - 1] * cdl;
                } else {
                    const double de = delta - gr.eps;
                    h[0] = -gr.eta * (de * de) - gr.beta * (delta - gr.gamma);
                    h[1] = delta * (-2.0 * gr.eta * de - gr.beta);
                    if (NO >= 2) h[NO >= 2 ? 2 : 0] = -gr.eta * (delta * delta);
                }
                E[0] = (gr.kind == GG_POLY) ? 1.0 : fast_exp(h[0]);
#pragma unroll
                for (int m = 1; m <= NO; ++m) {
                    double s = h[1] * E[m - 1];
#pragma unroll
                    for (int j = 2; j <= m; ++j) s = fma(double(j) * h[j], E[m - j], s);
                    E[m] = s * (1.0 / double(m));
                }
                dd = 1.0;
            }
#pragma unroll 1
            for (int s = 0; s < gr.dstep; ++s) dd *= delta;
            double G[NO + 1];
#pragma unroll
            for (int m = 0; m <= NO; ++m) {
                double s = E[m];
#pragma unroll
                for (int j = 1; j <= m; ++j) s = fma(gr.bd[j - 1], E[m - j], s);
                G[m] = s * dd;
            }
            double C[NO + 1];
#pragma unroll
            for (int n = 0; n <= NO; ++n) C[n] = 0.0;
            for (int q = 0; q < gr.ntq; ++q) {
                const GergRec hd = P.rec[rec];
                const double n0 = P.rec[rec + 1].val;
                rec += 2;
                const double t = hd.val;
                const double e = fast_exp2_prod(t, l2tau, K);
                const double eG0 = e * G[0], eG1 = e * G[1], teG0 = t * eG0;
                auto acc = [&](const int off, const double n) {
                    double* p = reinterpret_cast<double*>(reinterpret_cast<char*>(aw) + off);
                    *p = fma(n, eG0, *p);
                    if (WANT_H) {
                        double* ps = reinterpret_cast<double*>(reinterpret_cast<char*>(as) + off);
                        double* pr = reinterpret_cast<double*>(reinterpret_cast<char*>(ar) + off);
                        *ps = fma(n, teG0, *ps);
                        *pr = fma(n, eG1, *pr);
                    }
                };
                const int off0 = hd.b >> SH;
                double c = wload(W, off0) * n0;
                acc(off0, n0);
#pragma unroll 1
                for (int kk = 0; kk < hd.a; ++kk) {
                    const GergRec r0 = P.rec[rec], r1 = P.rec[rec + 1];
                    rec += 2;
                    c = fma(wload(W, r0.a >> SH), r0.val, c);
                    acc(r0.a >> SH, r0.val);
                    c = fma(wload(W, r1.a >> SH), r1.val, c);
                    acc(r1.a >> SH, r1.val);  // padding record: n = 0 -> adds 0 to a_0
                }
                double v = c * e;
                C[0] += v;
#pragma unroll
                for (int n = 1; n <= NO; ++n) {
                    v *= (t - double(n - 1)) * (1.0 / double(n));
                    C[n] += v;
                }
            }
#pragma unroll
            for (int n = 0; n <= NO; ++n)
#pragma unroll
                for (int m = 0; m <= NO; ++m) A[n][m] = fma(C[n], G[m], A[n][m]);
        }
    }

    // ---- reductions over the departure weights and the outputs -------------------------------------------------------------
    const double RT = Rgas * T;
    const double Fs = A[1][0], Fr = A[0][1];
    double S0 = 0.0, Ss = 0.0, Sr = 0.0;
    for (int q = 0; q < P.ndep; ++q) {
        const int i = P.dep_i[q], j = P.dep_j[q];
        const double wq = W[(nc + q) * WS], ri = W[i * WS], rj = W[j * WS];
        const double F = P.depF[q];
        const double m0 = F * aw[(nc + q) * WS];
        S0 = fma(aw[(nc + q) * WS], wq, S0);
        aw[(nc + q) * WS] = m0;  // M_ij
        aw[i * WS] = fma(m0 * invr, rj, aw[i * WS]);
        aw[j * WS] = fma(m0 * invr, ri, aw[j * WS]);
        if (WANT_H) {
            const double ms = F * as[(nc + q) * WS] * invr, mr = F * ar[(nc + q) * WS] * invr;
            Ss = fma(as[(nc + q) * WS], wq, Ss);
            Sr = fma(ar[(nc + q) * WS], wq, Sr);
            as[i * WS] = fma(ms, rj, as[i * WS]);
            as[j * WS] = fma(ms, ri, as[j * WS]);
            ar[i * WS] = fma(mr, rj, ar[i * WS]);
            ar[j * WS] = fma(mr, ri, ar[j * WS]);
        }
    }
    if (o.Psir) o.Psir[k] = RT * A[0][0];
    double lnZ = 0.0;
    if (o.lnphi || o.phi) lnZ = log(1.0 + Fr * invr);  // Z = 1 + Ar01, Ar01 = Phi_r / rho (derivs.hpp:1221-1225)
    const double Fss = WANT_H ? 2.0 * A[NO][0] + Fs : 0.0, Fsr = WANT_H ? A[1][1] : 0.0, Frr = WANT_H ? 2.0 * A[0][NO] + Fr : 0.0;
    const double S0r = S0 * invr, Ssr = Ss * invr, Srr = Sr * invr;
    const double invr2 = invr * invr;
    for (int i = 0; i < nc; ++i) {
        const double tk = vt[i * WS] - 2.0 * invr, dk = vd[i * WS] - invr;
        const double c0 = aw[i * WS] - S0r;
        const double g = fma(Fs, tk, fma(Fr, dk, c0));
        if (o.grad) o.grad[(long long)i * npts + k] = RT * g;
        if (o.lnphi) o.lnphi[(long long)i * npts + k] = g - lnZ;
        if (o.phi) o.phi[(long long)i * npts + k] = exp(g - lnZ);
        if (WANT_H) {
            const double cs = as[i * WS] - Ssr, cr = ar[i * WS] - Srr;
            as[i * WS] = cs;
            ar[i * WS] = cr;
        }
    }
    if (!WANT_H) return;
    const bool want_ip = ipp != nullptr;
    const double Z = 1.0 + Fr * invr;
    const double Ar02 = 2.0 * A[0][NO] * invr, Ar11 = Fsr * invr;
    const double dZx0 = Srr + (2.0 * Fsr + Frr) * invr;   // dZ/dx_j = Q_j + dZx0
    if (want_ip) {
        const IsoPropsOut& ip = *ipp;
        if (ip.pr) ip.pr[k] = RT * Fr;
        if (ip.splus) ip.splus[k] = (A[0][0] - Fs) * invr;
        if (ip.dpdT) ip.dpdT[k] = Rgas * (rhotot + Fr - Fsr);
    }

    // ---- E_k = -(1/rho^2) sum_{q contains k} M_q rho_other, into the rows the departure weights no longer need -------------
    for (int i = 0; i < nc; ++i) Eref(i) = 0.0;
    for (int q = 0; q < ndep; ++q) {
        const int i = P.dep_i[q], j = P.dep_j[q];
        const double m0 = -invr2 * aw[(nc + q) * WS];
        Eref(i) = fma(m0, W[j * WS], Eref(i));
        Eref(j) = fma(m0, W[i * WS], Eref(j));
    }
    // ---- pass C: Hessian.  Each pair's reducing term is evaluated again here for its three second derivatives; the diagonal
    // ones are summed as they come (hd[j] collects what rows i < j contribute to H_jj) ---------------------------------------
    const double FsY = Fs * iYT, FrY = Fr * iYV;  // Phi_s / Y_T, Phi_r / Y_V
    for (int i = 0; i < nc; ++i) hd[i * WS] = 2.0 * fma(FsY, P.Tc[i], FrY * P.vc[i]);
    const double const0 = (2.0 * S0r + 2.0 * Fs * invr + Fr * invr) * invr;
    const long long ncl = nc;
    int pidx = 0, qn = 0;
    for (int i = 0; i < nc; ++i) {
        const double ri = W[i * WS];
        const double gTi = vt[i * WS], gVi = vd[i * WS];
        const double ti = gTi - 2.0 * invr, di = gVi - invr;
        const double csi = as[i * WS], cri = ar[i * WS], Ei = Eref(i);
        const double Pi = fma(Fss, ti, fma(Fsr, di, csi)), Qi = fma(Fsr, ti, fma(Frr, di, cri));
        if (want_ip) {
            const IsoPropsOut& ip = *ipp;
            const double gi = fma(Fs, ti, fma(Fr, di, aw[i * WS] - S0r));
            const long long oi = (long long)i * npts + k;
            const double R1 = Rgas * (1.0 + log(ri));   // R (rho_ref + ln(rho_i / rho_ref)), rho_ref = 1 (derivs.hpp:1201, 1414)
            if (ip.d2PsirdTdrhoi) ip.d2PsirdTdrhoi[oi] = Rgas * (gi - Pi);
            if (ip.dpdrhovec) ip.dpdrhovec[oi] = fma(RT, Qi, RT);
            if (ip.pmv) ip.pmv[oi] = (1.0 + Qi) / (rhotot * (1.0 + 2.0 * Fr * invr + Ar02));
            if (ip.chempotVLE) ip.chempotVLE[oi] = fma(RT, gi, T * R1);
            if (ip.dchempotdT) ip.dchempotdT[oi] = Rgas * (gi - Pi) + R1;
            if (ip.dlnphi_dT) ip.dlnphi_dT[oi] = (Ar11 / Z - Pi) / T;
            const double dr = (Qi - (Fr * invr + Ar02) / Z) * invr;
            if (ip.dlnphi_drho) ip.dlnphi_drho[oi] = dr;
            if (ip.dlnphi_dv) ip.dlnphi_dv[oi] = -dr * (rhotot * rhotot);
        }
        const double dZxi = Qi + dZx0;
        double hdi = hd[i * WS];
#pragma unroll kIsoUnroll
        for (int j = i + 1; j < nc; ++j, ++pidx) {
            const double rj = W[j * WS];
            const GergRedPair pr = P.pair[pidx];
            const double dT = fma(pr.beta2T, ri, rj), dV = fma(pr.beta2V, ri, rj);
            const double den = dT * dV;
            double yT = 0.0, yV = 0.0;  // Y_ij / Y
            if (pr.beta2T == 1.0 && pr.beta2V == 1.0) {
                if (den != 0.0) { yT = pr.YT2 * iYT; yV = pr.YV2 * iYV; }
            } else if (den != 0.0) {
                const double inv = fast_rcp(den);
                const double invT = inv * dV, invV = inv * dT;
                const GergPairD a = gerg_pair_first(pr.beta2T, ri, rj, invT);
                const GergPairD b = gerg_pair_first(pr.beta2V, ri, rj, invV);
                const double s2 = 2.0 * (ri + rj);
                const double wT = pr.YT2 * FsY * invT, wV = pr.YV2 * FrY * invV;
                yT = pr.YT2 * iYT * (fma(-a.fj, pr.beta2T, s2 - a.fi) * invT);
                yV = pr.YV2 * iYV * (fma(-b.fj, pr.beta2V, s2 - b.fi) * invV);
                hdi = fma(wT, 2.0 * fma(-a.fi, pr.beta2T, rj), fma(wV, 2.0 * fma(-b.fi, pr.beta2V, rj), hdi));
                hd[j * WS] = fma(wT, 2.0 * (ri - a.fj), fma(wV, 2.0 * (ri - b.fj), hd[j * WS]));
            }
            const double gTj = vt[j * WS], gVj = vd[j * WS];
            const double tj = gTj - 2.0 * invr, dj = gVj - invr;
            double h = fma(tj, Pi, fma(dj, Qi, fma(ti, as[j * WS], di * ar[j * WS])));
            h = fma(Fs, fma(-gTi, gTj, yT), h);
            h = fma(Fr, fma(-gVi, gVj, yV), h);
            h += Ei + Eref(j) + const0;
            if (qn < ndep && P.dep_i[qn] == i && P.dep_j[qn] == j) { h = fma(aw[(nc + qn) * WS], invr, h); ++qn; }
            if (want_ip && ipp->dlnphi_dx) {
                const double Qj = fma(Fsr, tj, fma(Frr, dj, ar[j * WS]));
                ipp->dlnphi_dx[(ncl * i + j) * npts + k] = fma(rhotot, h, -(Qj + dZx0) / Z);   // d ln phi_i / d x_j
                ipp->dlnphi_dx[(ncl * j + i) * npts + k] = fma(rhotot, h, -dZxi / Z);          // d ln phi_j / d x_i
            }
            h *= RT;
            if (o.hess) {
                o.hess[(ncl * i + j) * npts + k] = h;
                o.hess[(ncl * j + i) * npts + k] = h;
            }
        }
        {
            double h = fma(ti, Pi, fma(di, Qi, fma(ti, csi, di * cri)));
            h = fma(-Fs * gTi, gTi, h);
            h = fma(-Fr * gVi, gVi, h);
            h += hdi + 2.0 * Ei + const0;
            if (want_ip && ipp->dlnphi_dx) ipp->dlnphi_dx[(ncl * i + i) * npts + k] = fma(rhotot, h, -dZxi / Z);
            if (o.hess) o.hess[(ncl * i + i) * npts + k] = RT * h;
        }
    }
}

// host launcher of the jet-sweep get_ATrhoXi table for the GERG families (tq_gerg_jet.cu)
void gerg_launch_ATrhoXi(const GergProgram& P, const DeviceBatch& in, double* out, cudaStream_t st);
// host launcher (tq_gerg_iso.cu)
void gerg_launch_isochoric(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st);
// the derived isochoric properties from the same analytic pass (tau^t temperature shapes, constant R)
void gerg_launch_isoprops(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st);

}  // namespace tq
