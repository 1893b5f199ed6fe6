// Batched pure-fluid saturation solver (first caller of the hot path): many independent isotherms of
// teqp::pure_VLE_T (/root/reference/include/teqp/algorithms/VLE_pure.hpp:18-143) advanced together.
//
// The reference solves, per temperature, the 2x2 system  p(T, rhoL) = p(T, rhoV),  mu(T, rhoL) = mu(T, rhoV)  by Newton's method with the
// residuals and Jacobian of IsothermPureVLEResiduals::call (:60-92), each iteration costing two get_Ar02n calls.  Here one launch of the model's
// Lambda table kernel evaluates Ar00, Ar01, Ar02 for the liquid AND vapour densities of every isotherm (a 2n-point batch), and one small kernel
// applies the Newton update with the reference's stopping rules (:121-139): a point whose step or residuals fall below epsilon keeps the
// densities it had BEFORE that step, exactly as the reference's `break` does; the other points go on.  No host round trip inside the loop.
#pragma once
#include "tq_jet.cuh"

namespace tq {

// one Newton step of do_pure_VLE_T for one isotherm.  A0, A1, A2: Ar00, Ar01, Ar02 at (liquid, vapour).  Returns true when the reference would
// have stopped at this iteration (densities unchanged).
TQ_HD bool pure_vle_newton_point(const double A0L, const double A1L, const double A2L, const double A0V, const double A1V, const double A2V,
                                 double& rhoL, double& rhoV) {
    const double R0_over_Rr = 1.0;   // VLE_pure.hpp:69 (R0 and Rr are both model.R(z))
    const double pRTL = rhoL * (R0_over_Rr + A1L);
    const double dpRTLdrhoL = R0_over_Rr + 2 * A1L + A2L;
    const double hatmurL = A1L + A0L + R0_over_Rr * log(rhoL);
    const double dhatmurLdrho = (2 * A1L + A2L) / rhoL + R0_over_Rr / rhoL;
    const double pRTV = rhoV * (R0_over_Rr + A1V);
    const double dpRTVdrhoV = R0_over_Rr + 2 * A1V + A2V;
    const double hatmurV = A1V + A0V + R0_over_Rr * log(rhoV);
    const double dhatmurVdrho = (2 * A1V + A2V) / rhoV + R0_over_Rr / rhoV;
    const double y0 = pRTL - pRTV, y1 = hatmurL - hatmurV;
    double a = dpRTLdrhoL, b = -dpRTVdrhoV, c = dhatmurLdrho, d = -dhatmurVdrho;
    // J v = -y by elimination with row pivoting (the reference uses a column-pivoted Householder QR: same solution to rounding)
    double r0 = -y0, r1 = -y1;
    if (fabs(c) > fabs(a)) { double t = a; a = c; c = t; t = b; b = d; d = t; t = r0; r0 = r1; r1 = t; }
    const double f = c / a;
    const double vV = (r1 - f * r0) / (d - f * b);
    const double vL = (r0 - b * vV) / a;
    const double newL = rhoL + vL, newV = rhoV + vV;
    const double eps = 2.220446049250313e-16;   // std::numeric_limits<double>::epsilon(), VLE_pure.hpp:130
    if (fabs(newL - rhoL) < eps && fabs(newV - rhoV) < eps) return true;
    if (fabs(y0) < eps && fabs(y1) < eps) return true;
    rhoL = newL; rhoV = newV;
    return false;
}

#if defined(__CUDACC__)
// T2, rho2: the 2n-point batch handed to the Lambda table kernel ([0, n) liquid, [n, 2n) vapour); done[n] (as doubles: scratch rows are doubles)
__global__ void __launch_bounds__(256) pure_vle_init_kernel(const long long n, const double* __restrict__ T, const double* __restrict__ rhoL0,
                                                            const double* __restrict__ rhoV0, double* __restrict__ T2, double* __restrict__ rho2,
                                                            double* __restrict__ done) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const double t = T[k];
        T2[k] = t; T2[n + k] = t;
        rho2[k] = rhoL0[k]; rho2[n + k] = rhoV0[k];
        done[k] = 0.0;
    }
}
// A: [3][2n] = Ar00, Ar01, Ar02 of the 2n-point batch
__global__ void __launch_bounds__(256) pure_vle_update_kernel(const long long n, const double* __restrict__ A, double* __restrict__ rho2,
                                                              double* __restrict__ done) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n2 = 2 * n;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        if (done[k] != 0.0) continue;
        double rL = rho2[k], rV = rho2[n + k];
        if (pure_vle_newton_point(A[k], A[n2 + k], A[2 * n2 + k], A[n + k], A[n2 + n + k], A[2 * n2 + n + k], rL, rV)) done[k] = 1.0;
        else { rho2[k] = rL; rho2[n + k] = rV; }
    }
}
// results; B: [2][2][2n] Lambda_nm (n, m <= 1) at the final densities when psat / dpsatdT are wanted (dpsatdT_pure, VLE_pure.hpp:158-171)
__global__ void __launch_bounds__(256) pure_vle_finish_kernel(const long long n, const double* __restrict__ T, const double* __restrict__ rho2,
                                                              const double* __restrict__ B, const double R, double* __restrict__ rhoL,
                                                              double* __restrict__ rhoV, double* __restrict__ psat, double* __restrict__ dpsatdT) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n2 = 2 * n;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const double rL = rho2[k], rV = rho2[n + k];
        rhoL[k] = rL; rhoV[k] = rV;
        if (B) {
            // B row (i*2 + j): Lambda_ij
            const double A01L = B[1 * n2 + k], A10L = B[2 * n2 + k], A01V = B[1 * n2 + n + k], A10V = B[2 * n2 + n + k];
            if (psat) psat[k] = rV * R * T[k] * (1.0 + A01V);
            if (dpsatdT) {
                const double hrVLERTV = A01V + A10V, hrVLERTL = A01L + A10L;
                dpsatdT[k] = (R * (hrVLERTV - hrVLERTL)) / (1 / rV - 1 / rL);
            }
        }
    }
}
#endif

// ---------------------------------------------------------------------------------------------------------------------------------------
// Binary mixture, T and liquid composition given: teqp::mix_VLE_Tx (/root/reference/include/teqp/algorithms/VLE.hpp:76-192).  Unknowns
// (rhoL_0, rhoL_1, rhoV_0, rhoV_1); residuals: equal chemical potentials (2), equal pressure, x_0 of the liquid = xspec_0; Jacobian from the Hessians of
// Psi^r (VLE.hpp:96-138).  One Newton step with the reference's step limiter (:148-156), frozen zero components (:158-164) and stopping rules (:168-189).
// f, g, H: Psi^r, its gradient [2] and Hessian [4, row-major] for (liquid, vapour).  code: VLE_return_code (VLE_types.hpp:44): 0 unset (still iterating),
// 1 xtol_satisfied, 2 functol_satisfied, 4 maxiter_met, 5 notfinite_step.  Returns true when the point is finished.
TQ_HD bool mix_vle_newton_point(const double RT, const double fL, const double* gL, const double* HL, const double fV, const double* gV, const double* HV,
                                const double* xspec, const double atol, const double reltol, const double axtol, const double relxtol, const bool last_iter,
                                double* x /* rhoL0, rhoL1, rhoV0, rhoV1 */, double& code) {
    const double rhoL = x[0] + x[1], rhoV = x[2] + x[3];
    const double pL = rhoL * RT - fL + (x[0] * gL[0] + x[1] * gL[1]);
    const double pV = rhoV * RT - fV + (x[2] * gV[0] + x[3] * gV[1]);
    const double dpL0 = RT + (HL[0] * x[0] + HL[1] * x[1]), dpL1 = RT + (HL[2] * x[0] + HL[3] * x[1]);
    const double dpV0 = RT + (HV[0] * x[2] + HV[1] * x[3]), dpV1 = RT + (HV[2] * x[2] + HV[3] * x[3]);
    const bool nz0 = x[0] > 0 && x[2] > 0, nz1 = x[1] > 0 && x[3] > 0;
    double r[4];
    r[0] = nz0 ? gL[0] + RT * log(x[0]) - (gV[0] + RT * log(x[2])) : gL[0] - gV[0];
    r[1] = nz1 ? gL[1] + RT * log(x[1]) - (gV[1] + RT * log(x[3])) : gL[1] - gV[1];
    r[2] = pL - pV;
    r[3] = x[0] / rhoL - xspec[0];
    double J[4][5] = {
        {HL[0] + (nz0 ? RT / x[0] : 0.0), HL[1], -(HV[0] + (nz0 ? RT / x[2] : 0.0)), -HV[1], -r[0]},
        {HL[2], HL[3] + (nz1 ? RT / x[1] : 0.0), -HV[2], -(HV[3] + (nz1 ? RT / x[3] : 0.0)), -r[1]},
        {dpL0, dpL1, -dpV0, -dpV1, -r[2]},
        {(rhoL - x[0]) / (rhoL * rhoL), -x[0] / (rhoL * rhoL), 0.0, 0.0, -r[3]}};
    // J dx = -r: elimination with row pivoting on the augmented matrix (the reference: column-pivoted Householder QR, VLE.hpp:141)
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int q = c + 1; q < 4; ++q) if (fabs(J[q][c]) > fabs(J[piv][c])) piv = q;
        if (piv != c) for (int k = 0; k < 5; ++k) { const double t = J[c][k]; J[c][k] = J[piv][k]; J[piv][k] = t; }
        for (int q = c + 1; q < 4; ++q) {
            const double m = J[q][c] / J[c][c];
            for (int k = c; k < 5; ++k) J[q][k] -= m * J[c][k];
        }
    }
    double dx[4];
    for (int c = 3; c >= 0; --c) {
        double s = J[c][4];
        for (int k = c + 1; k < 4; ++k) s -= J[c][k] * dx[k];
        dx[c] = s / J[c][c];
    }
    bool none_finite = true;
    for (int c = 0; c < 4; ++c) if (dx[c] - dx[c] == 0.0) none_finite = false;   // finite <=> x - x == 0
    if (none_finite) { code = 5.0; return true; }
    bool neg = false;
    for (int c = 0; c < 4; ++c) if (x[c] + dx[c] < 0) neg = true;
    if (neg) {   // at most half way to the most constraining zero concentration
        double fmin = dx[0] / (-x[0]);
        for (int c = 1; c < 4; ++c) { const double q = dx[c] / (-x[c]); if (q < fmin) fmin = q; }
        for (int c = 0; c < 4; ++c) dx[c] *= fmin / 2;
    }
    for (int i = 0; i < 2; ++i) if (xspec[i] == 0) { dx[i] = 0; dx[i + 2] = 0; }
    for (int c = 0; c < 4; ++c) x[c] += dx[c];
    bool all_x = true, all_r = true, all_tiny = true;
    for (int c = 0; c < 4; ++c) {
        if (!(fabs(dx[c]) < axtol + relxtol * fabs(x[c]))) all_x = false;
        if (!(fabs(r[c]) < atol + reltol * fabs(r[c]))) all_r = false;
        if (!(fabs(x[c] - dx[c]) < 2.2250738585072014e-308)) all_tiny = false;
    }
    if (all_x) { code = 1.0; return true; }
    if (all_r) { code = 2.0; return true; }
    if (all_tiny) { code = 1.0; return true; }
    if (last_iter) { code = 4.0; return true; }
    return false;
}

// teqp::mix_VLE_Tp (VLE.hpp:196-380): T and p given, unknowns the four molar concentrations; residuals of hybrj_functor__mix_VLE_Tp::operator() (:204-233: equal
// chemical potentials, (p_V - p)/p, (p_L - p)/p), Jacobian of ::df (:234-272), and the plain Newton loop of :351-367 -- step limiter only, no stopping rule:
// every point takes maxiter steps, as the reference does.
TQ_HD void mix_vle_tp_newton_point(const double RT, const double p, const double fL, const double* gL, const double* HL, const double fV, const double* gV,
                                   const double* HV, double* x /* rhoL0, rhoL1, rhoV0, rhoV1 */) {
    const double rhoL = x[0] + x[1], rhoV = x[2] + x[3];
    const double pL = rhoL * RT - fL + (x[0] * gL[0] + x[1] * gL[1]);
    const double pV = rhoV * RT - fV + (x[2] * gV[0] + x[3] * gV[1]);
    const double dpL0 = RT + (HL[0] * x[0] + HL[1] * x[1]), dpL1 = RT + (HL[2] * x[0] + HL[3] * x[1]);
    const double dpV0 = RT + (HV[0] * x[2] + HV[1] * x[3]), dpV1 = RT + (HV[2] * x[2] + HV[3] * x[3]);
    const bool nz0 = x[0] > 0 && x[2] > 0, nz1 = x[1] > 0 && x[3] > 0;
    const double r0 = nz0 ? gL[0] + RT * log(x[0]) - (gV[0] + RT * log(x[2])) : gL[0] - gV[0];
    const double r1 = nz1 ? gL[1] + RT * log(x[1]) - (gV[1] + RT * log(x[3])) : gL[1] - gV[1];
    const double r2 = (pV - p) / p, r3 = (pL - p) / p;
    double J[4][5] = {
        {HL[0] + (nz0 ? RT / x[0] : 0.0), HL[1], -(HV[0] + (nz0 ? RT / x[2] : 0.0)), -HV[1], -r0},
        {HL[2], HL[3] + (nz1 ? RT / x[1] : 0.0), -HV[2], -(HV[3] + (nz1 ? RT / x[3] : 0.0)), -r1},
        {0.0, 0.0, dpV0 / p, dpV1 / p, -r2},
        {dpL0 / p, dpL1 / p, 0.0, 0.0, -r3}};
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int q = c + 1; q < 4; ++q) if (fabs(J[q][c]) > fabs(J[piv][c])) piv = q;
        if (piv != c) for (int k = 0; k < 5; ++k) { const double t = J[c][k]; J[c][k] = J[piv][k]; J[piv][k] = t; }
        for (int q = c + 1; q < 4; ++q) {
            const double m = J[q][c] / J[c][c];
            for (int k = c; k < 5; ++k) J[q][k] -= m * J[c][k];
        }
    }
    double dx[4];
    for (int c = 3; c >= 0; --c) {
        double s = J[c][4];
        for (int k = c + 1; k < 4; ++k) s -= J[c][k] * dx[k];
        dx[c] = s / J[c][c];
    }
    bool neg = false;
    for (int c = 0; c < 4; ++c) if (x[c] + dx[c] < 0) neg = true;
    if (neg) {
        double fmin = dx[0] / (-x[0]);
        for (int c = 1; c < 4; ++c) { const double q = dx[c] / (-x[c]); if (q < fmin) fmin = q; }
        for (int c = 0; c < 4; ++c) dx[c] *= fmin / 2;
    }
    for (int c = 0; c < 4; ++c) x[c] += dx[c];
}

#if defined(__CUDACC__)
// rv2: [2][2n] molar concentrations of the 2n-point isochoric batch ([0, n) liquid, [n, 2n) vapour)
__global__ void __launch_bounds__(256) mix_vle_init_kernel(const long long n, const double* __restrict__ T, const double* __restrict__ rvL0,
                                                           const double* __restrict__ rvV0, double* __restrict__ T2, double* __restrict__ rv2,
                                                           double* __restrict__ code) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const double t = T[k];
        T2[k] = t; T2[n + k] = t;
        for (int i = 0; i < 2; ++i) { rv2[(long long)i * 2 * n + k] = rvL0[(long long)i * n + k]; rv2[(long long)i * 2 * n + n + k] = rvV0[(long long)i * n + k]; }
        code[k] = 0.0;
    }
}
__global__ void __launch_bounds__(256) mix_vle_update_kernel(const long long n, const double R, const double* __restrict__ T, const double* __restrict__ xspec,
                                                             const double* __restrict__ f, const double* __restrict__ g, const double* __restrict__ H,
                                                             const double atol, const double reltol, const double axtol, const double relxtol,
                                                             const int last_iter, double* __restrict__ rv2, double* __restrict__ code) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n2 = 2 * n;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        if (code[k] != 0.0) continue;
        double x[4] = {rv2[k], rv2[n2 + k], rv2[n + k], rv2[n2 + n + k]};
        const double gL[2] = {g[k], g[n2 + k]}, gV[2] = {g[n + k], g[n2 + n + k]};
        double HL[4], HV[4];
        for (int q = 0; q < 4; ++q) { HL[q] = H[(long long)q * n2 + k]; HV[q] = H[(long long)q * n2 + n + k]; }
        const double xs[2] = {xspec[k], xspec[n + k]};
        double c = 0.0;
        mix_vle_newton_point(R * T[k], f[k], gL, HL, f[n + k], gV, HV, xs, atol, reltol, axtol, relxtol, last_iter != 0, x, c);
        if (c != 5.0) { rv2[k] = x[0]; rv2[n2 + k] = x[1]; rv2[n + k] = x[2]; rv2[n2 + n + k] = x[3]; }   // notfinite_step leaves the densities unchanged
        code[k] = c;
    }
}
__global__ void __launch_bounds__(256) mix_vle_tp_update_kernel(const long long n, const double R, const double* __restrict__ T, const double* __restrict__ p,
                                                                const double* __restrict__ f, const double* __restrict__ g, const double* __restrict__ H,
                                                                double* __restrict__ rv2) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n2 = 2 * n;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        double x[4] = {rv2[k], rv2[n2 + k], rv2[n + k], rv2[n2 + n + k]};
        const double gL[2] = {g[k], g[n2 + k]}, gV[2] = {g[n + k], g[n2 + n + k]};
        double HL[4], HV[4];
        for (int q = 0; q < 4; ++q) { HL[q] = H[(long long)q * n2 + k]; HV[q] = H[(long long)q * n2 + n + k]; }
        mix_vle_tp_newton_point(R * T[k], p[k], f[k], gL, HL, f[n + k], gV, HV, x);
        rv2[k] = x[0]; rv2[n2 + k] = x[1]; rv2[n + k] = x[2]; rv2[n2 + n + k] = x[3];
    }
}
__global__ void __launch_bounds__(256) mix_vle_finish_kernel(const long long n, const double* __restrict__ rv2, double* __restrict__ rvL, double* __restrict__ rvV) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride)
        for (int i = 0; i < 2; ++i) { rvL[(long long)i * n + k] = rv2[(long long)i * 2 * n + k]; rvV[(long long)i * n + k] = rv2[(long long)i * 2 * n + n + k]; }
}
#endif

}  // namespace tq
