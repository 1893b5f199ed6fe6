// Cross-composition derivative tables (get_ATrhoXiXj, get_ATrhoXiXjXk; derivs.hpp:294-339, 423-465) by DIRECTIONAL jets, for the
// families whose model code NVVM cannot compile over the five-level jets of tq_generic.cuh (PC-SAFT, SAFT-VR-Mie + association).
//
// A three-level jet J3<1,1,N> whose composition level is seeded along a direction d (x(s) = x + s d) holds c_n(d) = D_d^n f / n!.
// Mixed partials follow from pure directional ones by polarisation -- exact identities, no step size:
//     f_ij  = c_2(e_i + e_j) - c_2(e_i) - c_2(e_j)
//     f_ijk = c_3(e_i+e_j+e_k) - c_3(e_i+e_j) - c_3(e_i+e_k) - c_3(e_j+e_k) + c_3(e_i) + c_3(e_j) + c_3(e_k)
// so get_ATrhoXiXj takes 3 sweeps of J3<1,1,2> (12 coefficients) instead of one of J5<1,1,2,2,0> (36), and get_ATrhoXiXjXk 7 sweeps of
// J3<1,1,3> (16) instead of one of J5<1,1,1,1,1> (32).  The differences cancel to eps x (largest pure directional derivative).
// Entries of the tables that the reference does not expose and these sweeps do not give (x_i^2 x_j, x_i x_j^2, x_i^2 x_j^2) are NaN.
#pragma once
#include "tq_generic.cuh"

namespace tq {

template<class M, int MAXN, int NX>
TQ_HD J3<1, 1, NX> dir_jet(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin, const int a, const int b,
                           const int c) {
    using XT = J3<0, 0, NX>;
    using TT = J3<1, 0, 0>;
    using RT = J3<0, 1, 0>;
    const TT T = TT::reciprocal_scaled_variable(T0);
    RT rho;
    rho.c[0] = Tay<1, Tay<0, double>>::scaled_variable(rho0);
    XT x[MAXN];
    for (int k = 0; k < nc; ++k) {
        x[k] = XT(xin[k]);
        if (k == a || k == b || k == c) x[k].c[0].c[0].c[1] = 1.0;
    }
    return embed<J3<1, 1, NX>>(M::template alphar<MAXN>(p, T, rho, x));
}

// out[((NT*2+ND)*3+NXi)*3+NXj][npts]
template<class M, int MAXN>
TQ_HD void dir_ATrhoXiXj_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin, const int i, const int j,
                               double* out, const long long ostride) {
    const J3<1, 1, 2> ai = dir_jet<M, MAXN, 2>(p, nc, T0, rho0, xin, i, -1, -1);
    const J3<1, 1, 2> aj = dir_jet<M, MAXN, 2>(p, nc, T0, rho0, xin, j, -1, -1);
    const J3<1, 1, 2> aij = dir_jet<M, MAXN, 2>(p, nc, T0, rho0, xin, i, j, -1);
    const double nan = ai.c[0].c[0].c[0] * 0.0 + (0.0 / 0.0);   // quiet NaN without a library call on either side
#pragma unroll
    for (int nt = 0; nt <= 1; ++nt)
#pragma unroll
        for (int nd = 0; nd <= 1; ++nd) {
            double* o = out + (long long)((nt * 2 + nd) * 9) * ostride;
            const double ci1 = ai.c[nt].c[nd].c[1], ci2 = ai.c[nt].c[nd].c[2], cj1 = aj.c[nt].c[nd].c[1], cj2 = aj.c[nt].c[nd].c[2];
            o[0 * ostride] = ai.c[nt].c[nd].c[0];                         // (0,0)
            o[1 * ostride] = cj1;                                         // (0,1)
            o[2 * ostride] = 2.0 * cj2;                                   // (0,2)
            o[3 * ostride] = ci1;                                         // (1,0)
            o[4 * ostride] = aij.c[nt].c[nd].c[2] - ci2 - cj2;            // (1,1)
            o[5 * ostride] = nan;                                         // (1,2)
            o[6 * ostride] = 2.0 * ci2;                                   // (2,0)
            o[7 * ostride] = nan;                                         // (2,1)
            o[8 * ostride] = nan;                                         // (2,2)
        }
}

// out[(((NT*2+ND)*2+NXi)*2+NXj)*2+NXk][npts]
template<class M, int MAXN>
TQ_HD void dir_ATrhoXiXjXk_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin, const int i, const int j,
                                 const int kk, double* out, const long long ostride) {
    const J3<1, 1, 3> ai = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, i, -1, -1);
    const J3<1, 1, 3> aj = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, j, -1, -1);
    const J3<1, 1, 3> ak = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, kk, -1, -1);
    const J3<1, 1, 3> aij = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, i, j, -1);
    const J3<1, 1, 3> aik = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, i, kk, -1);
    const J3<1, 1, 3> ajk = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, j, kk, -1);
    const J3<1, 1, 3> aijk = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, i, j, kk);
#pragma unroll
    for (int nt = 0; nt <= 1; ++nt)
#pragma unroll
        for (int nd = 0; nd <= 1; ++nd) {
            double* o = out + (long long)((nt * 2 + nd) * 8) * ostride;
#define TQ_C(a, n) a.c[nt].c[nd].c[n]
            o[0 * ostride] = TQ_C(ai, 0);                                                                  // 000
            o[1 * ostride] = TQ_C(ak, 1);                                                                  // 001
            o[2 * ostride] = TQ_C(aj, 1);                                                                  // 010
            o[3 * ostride] = TQ_C(ajk, 2) - TQ_C(aj, 2) - TQ_C(ak, 2);                                     // 011
            o[4 * ostride] = TQ_C(ai, 1);                                                                  // 100
            o[5 * ostride] = TQ_C(aik, 2) - TQ_C(ai, 2) - TQ_C(ak, 2);                                     // 101
            o[6 * ostride] = TQ_C(aij, 2) - TQ_C(ai, 2) - TQ_C(aj, 2);                                     // 110
            o[7 * ostride] = TQ_C(aijk, 3) - TQ_C(aij, 3) - TQ_C(aik, 3) - TQ_C(ajk, 3) + TQ_C(ai, 3) + TQ_C(aj, 3) + TQ_C(ak, 3);   // 111
#undef TQ_C
        }
}

#if defined(__CUDACC__)
template<class M, int WHICH>
__global__ void __launch_bounds__(64) dir_cross_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in, const int i,
                                                       const int j, const int kk, double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double x[M::kMaxC];
        load_x<M, M::kMaxC>(x, nc, in, k);
        if constexpr (WHICH == 2) dir_ATrhoXiXj_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, i, j, out + k, npts);
        else dir_ATrhoXiXjXk_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, i, j, kk, out + k, npts);
    }
}
template<class M, int WHICH>
void dir_launch_cross(const typename M::Params& p, int nc, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    dir_cross_kernel<M, WHICH><<<grid_for(in.npts, block, sm_count() * 32), block, 0, st>>>(p, nc, in, i, j, k, out);
    TQ_CUDA(cudaGetLastError());
}
#endif

}  // namespace tq
