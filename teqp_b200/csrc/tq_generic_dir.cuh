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

// The sweeps over the directions are a ROLLED loop (#pragma unroll 1) with the seeds chosen at run time, so the model code is instantiated
// once per kernel -- inlining it 3 or 7 times is what NVVM did not finish compiling for these families.

// out[((NT*2+ND)*3+NXi)*3+NXj][npts]
template<class M, int MAXN>
TQ_HD void dir_ATrhoXiXj_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin, const int i, const int j,
                               double* out, const long long ostride) {
    double c[3][4][3];   // [direction: e_i, e_j, e_i + e_j][(nt, nd)][order]
#pragma unroll 1
    for (int s = 0; s < 3; ++s) {
        const J3<1, 1, 2> a = dir_jet<M, MAXN, 2>(p, nc, T0, rho0, xin, s == 1 ? j : i, s == 2 ? j : -1, -1);
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int n = 0; n < 3; ++n) c[s][q][n] = a.c[q >> 1].c[q & 1].c[n];
    }
    const double nan = c[0][0][0] * 0.0 + (0.0 / 0.0);   // quiet NaN
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        double* o = out + (long long)(q * 9) * ostride;
        o[0 * ostride] = c[0][q][0];                                   // (0,0)
        o[1 * ostride] = c[1][q][1];                                   // (0,1)
        o[2 * ostride] = 2.0 * c[1][q][2];                             // (0,2)
        o[3 * ostride] = c[0][q][1];                                   // (1,0)
        o[4 * ostride] = c[2][q][2] - c[0][q][2] - c[1][q][2];         // (1,1)
        o[5 * ostride] = nan;                                          // (1,2)
        o[6 * ostride] = 2.0 * c[0][q][2];                             // (2,0)
        o[7 * ostride] = nan;                                          // (2,1)
        o[8 * ostride] = nan;                                          // (2,2)
    }
}

// out[(((NT*2+ND)*2+NXi)*2+NXj)*2+NXk][npts]
template<class M, int MAXN>
TQ_HD void dir_ATrhoXiXjXk_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin, const int i, const int j,
                                 const int kk, double* out, const long long ostride) {
    // directions, as bit masks over (i, j, k): 1 = e_i, 2 = e_j, 4 = e_k, 3 = e_i + e_j, 5, 6, 7
    double c[8][4][4];   // [mask][(nt, nd)][order]
#pragma unroll 1
    for (int msk = 1; msk < 8; ++msk) {
        const J3<1, 1, 3> a = dir_jet<M, MAXN, 3>(p, nc, T0, rho0, xin, (msk & 1) ? i : -1, (msk & 2) ? j : -1, (msk & 4) ? kk : -1);
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int n = 0; n < 4; ++n) c[msk][q][n] = a.c[q >> 1].c[q & 1].c[n];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        double* o = out + (long long)(q * 8) * ostride;
        o[0 * ostride] = c[1][q][0];                                                                                         // 000
        o[1 * ostride] = c[4][q][1];                                                                                         // 001
        o[2 * ostride] = c[2][q][1];                                                                                         // 010
        o[3 * ostride] = c[6][q][2] - c[2][q][2] - c[4][q][2];                                                               // 011
        o[4 * ostride] = c[1][q][1];                                                                                         // 100
        o[5 * ostride] = c[5][q][2] - c[1][q][2] - c[4][q][2];                                                               // 101
        o[6 * ostride] = c[3][q][2] - c[1][q][2] - c[2][q][2];                                                               // 110
        o[7 * ostride] = c[7][q][3] - c[3][q][3] - c[5][q][3] - c[6][q][3] + c[1][q][3] + c[2][q][3] + c[4][q][3];           // 111
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


// ------------------------------------------------------------------------------------------------------------------
// get_ATrhoXi table by three SMALL jets instead of one rectangular J3<2,2,3> (36 coefficients per number: 40 KB of local memory per thread and
// 15 instruction-fetch stall cycles per issue for SAFT-VR-Mie + association, profiles/r01j).  The reference's 23 combinations need
// NT + ND <= 2, i.e. the (1/T, rho) blocks {(2,0)}, {(0,2)}, {(0,0),(1,0),(0,1),(1,1)} -- each with NXi <= 3:
//     part 0: J3<1,1,3> (16 coefficients) -> entries nt <= 1, nd <= 1      part 1: J3<2,0,3> (12) -> nt = 2, nd = 0      part 2: J3<0,2,3> (12) -> nt = 0, nd = 2
// Multiplication cost 90 + 60 + 60 against 360 for the rectangular jet, and every number fits the register file.  The entries of the table
// with NT + ND > 2, which no reference entry point exposes, are NaN (part 0 writes them).
// out[((i*3+nt)*3+nd)*4+nxi][npts]
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN, int CT, int CD>
TQ_HD void split_ATrhoXi_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin, double* out,
                               const long long ostride) {
    using XT = J3<0, 0, 3>;
    using TT = J3<CT, 0, 0>;
    using RT = J3<0, CD, 0>;
    const TT T = TT::reciprocal_scaled_variable(T0);
    RT rho;
    rho.c[0] = Tay<CD, Tay<0, double>>::scaled_variable(rho0);
#pragma unroll 1
    for (int i = 0; i < nc; ++i) {
        XT x[MAXN];
        for (int j = 0; j < nc; ++j) {
            x[j] = XT(xin[j]);
            if (j == i) x[j].c[0].c[0].c[1] = 1.0;
        }
        const J3<CT, CD, 3> a = embed<J3<CT, CD, 3>>(M::template alphar<MAXN>(p, T, rho, x));
        double* o = out + (long long)(i * 36) * ostride;
#pragma unroll
        for (int nx = 0; nx <= 3; ++nx) {
            const double fx = (nx == 2 ? 2.0 : (nx == 3 ? 6.0 : 1.0));
            if constexpr (CT == 1 && CD == 1) {
                const double nan = a.c[0].c[0].c[0] * 0.0 + (0.0 / 0.0);
#pragma unroll
                for (int nt = 0; nt <= 2; ++nt)
#pragma unroll
                    for (int nd = 0; nd <= 2; ++nd) {
                        if (nt <= 1 && nd <= 1) o[(long long)((nt * 3 + nd) * 4 + nx) * ostride] = fx * a.c[nt].c[nd].c[nx];
                        else if (nt + nd > 2) o[(long long)((nt * 3 + nd) * 4 + nx) * ostride] = nan;
                    }
            } else if constexpr (CT == 2) {
                o[(long long)((2 * 3 + 0) * 4 + nx) * ostride] = 2.0 * fx * a.c[2].c[0].c[nx];
            } else {
                o[(long long)((0 * 3 + 2) * 4 + nx) * ostride] = 2.0 * fx * a.c[0].c[2].c[nx];
            }
        }
    }
}

#if defined(__CUDACC__)
template<class M, int CT, int CD>
__global__ void __launch_bounds__(64) split_ATrhoXi_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in,
                                                           double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double x[M::kMaxC];
        load_x<M, M::kMaxC>(x, nc, in, k);
        split_ATrhoXi_point<M, M::kMaxC, CT, CD>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, out + k, npts);
    }
}
template<class M, int CT, int CD>
void split_launch_ATrhoXi(const typename M::Params& p, int nc, const DeviceBatch& in, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    split_ATrhoXi_kernel<M, CT, CD><<<grid_for(in.npts, block, sm_count() * 32), block, 0, st>>>(p, nc, in, out);
    TQ_CUDA(cudaGetLastError());
}
#endif

}  // namespace tq
