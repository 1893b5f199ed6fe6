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

}  // namespace tq
