// Generic jet-sweep kernels, shared by every model family whose alphar(T, rho, x) is written once over jet types
// (cubic, PC-SAFT, SAFT-VR-Mie, ...).  A family supplies
//
//   struct Dev { using Params = <POD>;                                   // passed as a __grid_constant__ kernel parameter
//                template<int MAXN, class TT, class RT, class XT>
//                static __device__ auto alphar(const Params&, const TT& T, const RT& rho, const XT* x);
//                template<class XT> static __device__ auto R(const Params&, const XT* x); };
//
// and gets, one state point per thread:
//   * generic_launch_ar_table   -- Lambda^r_nm for all n<=NT, m<=ND in ONE sweep of a J3<NT,ND,0> jet
//                                  (reference: one nested-dual sweep per (n,m), derivs.hpp:150-238, 543-624)
//   * generic_launch_ATrhoXi    -- the whole get_ATrhoXi table per component in one J3<2,2,3> sweep (derivs.hpp:247-292)
//   * generic_launch_isochoric  -- Psi^r, its gradient and Hessian in the molar concentrations and ln(phi)
//                                  (derivs.hpp:1031-1124, 1208-1230)
#pragma once
#include "tq_common.hpp"
#include "tq_jet.cuh"

namespace tq {

// A family that declares `static constexpr bool kNoSmallVariants = true` gets ONE kernel variant per jet shape (arrays sized
// kMaxC, component count at run time) instead of the four specialised for 1, 2, 3 and kMaxC components.
template<class M, class = void> struct small_variants : std::true_type {};
template<class M> struct small_variants<M, std::void_t<decltype(M::kNoSmallVariants)>> : std::false_type {};

// ------------------------------------------------------------------------------------------------------------------
// Lambda^r_nm table
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN, int CT, int CD>
TQ_HD void generic_ar_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin,
                            const int NT, const int ND, double* out, const long long ostride, const bool rho_unit_seed = false) {
    using XT = J3<0, 0, 0>;
    XT x[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i)
        if (i < nc) x[i] = XT(xin[i]);
    using TT = J3<CT, 0, 0>;
    using RT = J3<0, CD, 0>;
    const TT T = TT::reciprocal_scaled_variable(T0);
    RT rho;
    rho.c[0] = rho_unit_seed ? Tay<CD, Tay<0, double>>::unit_variable(rho0) : Tay<CD, Tay<0, double>>::scaled_variable(rho0);
    const J3<CT, CD, 0> a = embed<J3<CT, CD, 0>>(M::template alphar<MAXN>(p, T, rho, x));
    double fn = 1.0;
#pragma unroll
    for (int n = 0; n <= CT; ++n) {
        if (n >= 2) fn *= double(n);
        double fm = 1.0;
#pragma unroll
        for (int m = 0; m <= CD; ++m) {
            if (m >= 2) fm *= double(m);
            if (n <= NT && m <= ND) out[(long long)(n * (ND + 1) + m) * ostride] = (fn * fm) * a.c[n].c[m].c[0];
        }
    }
}

#if defined(__CUDACC__)
template<class M, int MAXN>
__device__ __forceinline__ void load_x(double (&x)[MAXN], const int nc, const DeviceBatch& in, const long long k) {
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i)
        if (i < nc) x[i] = __ldg(in.x + (long long)i * in.x_stride + (in.x_bcast ? 0 : k));
}

template<class M, int MAXN, int CT, int CD>
__global__ void __launch_bounds__(128) generic_ar_table_kernel(const __grid_constant__ typename M::Params p, const int nc,
                                                               const DeviceBatch in, const int NT, const int ND,
                                                               double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double x[MAXN];
        load_x<M, MAXN>(x, nc, in, k);
        generic_ar_point<M, MAXN, CT, CD>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, NT, ND, out + k, npts, in.rho_unit_seed != 0);
    }
}

template<class M, int CT, int CD>
void generic_launch_ar_table_inst(const typename M::Params& p, int nc, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) {
    const int block = 128;
    const int grid = grid_for(in.npts, block, sm_count() * 32);
    if constexpr (small_variants<M>::value) {
        if (nc == 1) generic_ar_table_kernel<M, 1, CT, CD><<<grid, block, 0, st>>>(p, nc, in, NT, ND, out);
        else if (nc == 2) generic_ar_table_kernel<M, 2, CT, CD><<<grid, block, 0, st>>>(p, nc, in, NT, ND, out);
        else if (nc == 3) generic_ar_table_kernel<M, 3, CT, CD><<<grid, block, 0, st>>>(p, nc, in, NT, ND, out);
        else generic_ar_table_kernel<M, M::kMaxC, CT, CD><<<grid, block, 0, st>>>(p, nc, in, NT, ND, out);
    } else {
        generic_ar_table_kernel<M, M::kMaxC, CT, CD><<<grid, block, 0, st>>>(p, nc, in, NT, ND, out);
    }
    TQ_CUDA(cudaGetLastError());
}

// picks the smallest compiled jet shape that covers (NT, ND).  Split in two so that a heavy model family can instantiate the
// density-only shapes and the temperature / mixed shapes in separate translation units (parallel compilation).
template<class M>
void generic_launch_ar_table_rho(const typename M::Params& p, int nc, int ND, const DeviceBatch& in, double* out, cudaStream_t st) {
    if (ND <= 2) generic_launch_ar_table_inst<M, 0, 2>(p, nc, 0, ND, in, out, st);
    else if (ND <= 4) generic_launch_ar_table_inst<M, 0, 4>(p, nc, 0, ND, in, out, st);
    else generic_launch_ar_table_inst<M, 0, 6>(p, nc, 0, ND, in, out, st);
}
template<class M>
void generic_launch_ar_table_mixed(const typename M::Params& p, int nc, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) {
    if (ND == 0) {
        if (NT <= 2) generic_launch_ar_table_inst<M, 2, 0>(p, nc, NT, ND, in, out, st);
        else generic_launch_ar_table_inst<M, 4, 0>(p, nc, NT, ND, in, out, st);
    } else if (NT <= 2 && ND <= 2) {
        generic_launch_ar_table_inst<M, 2, 2>(p, nc, NT, ND, in, out, st);
    } else if (NT <= 2 && ND <= 4) {
        generic_launch_ar_table_inst<M, 2, 4>(p, nc, NT, ND, in, out, st);
    } else {
        throw InvalidArgument("mixed derivative table (" + std::to_string(NT) + "," + std::to_string(ND) + ") is larger than the compiled jet shapes (NT<=2 with ND<=4, or one of them zero)");
    }
}
inline void generic_check_ar_shape(int NT, int ND) {
    if (NT < 0 || NT > kMaxDerivT || ND < 0 || ND > kMaxDerivD) throw InvalidArgument("derivative order out of range");
}
template<class M>
void generic_launch_ar_table(const typename M::Params& p, int nc, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) {
    generic_check_ar_shape(NT, ND);
    if (in.npts == 0) return;
    if (NT == 0) generic_launch_ar_table_rho<M>(p, nc, ND, in, out, st);
    else generic_launch_ar_table_mixed<M>(p, nc, NT, ND, in, out, st);
}

#endif  // __CUDACC__

// ------------------------------------------------------------------------------------------------------------------
// get_ATrhoXi table: out[((i*3+nt)*3+nd)*4+nxi][npts], nt,nd in 0..2, nxi in 0..3
// ------------------------------------------------------------------------------------------------------------------

template<class M, int MAXN>
TQ_HD void generic_ATrhoXi_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin,
                                 double* out, const long long ostride) {
    using XT = J3<0, 0, 3>;
    using TT = J3<2, 0, 0>;
    using RT = J3<0, 2, 0>;
    const TT T = TT::reciprocal_scaled_variable(T0);
    RT rho;
    rho.c[0] = Tay<2, Tay<0, double>>::scaled_variable(rho0);
    for (int i = 0; i < nc; ++i) {
        XT x[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
        for (int j = 0; j < MAXN; ++j) {
            if (j < nc) {
                x[j] = XT(xin[j]);
                if (j == i) x[j].c[0].c[0].c[1] = 1.0;
            }
        }
        const J3<2, 2, 3> a = embed<J3<2, 2, 3>>(M::template alphar<MAXN>(p, T, rho, x));
#pragma unroll
        for (int nt = 0; nt <= 2; ++nt)
#pragma unroll
            for (int nd = 0; nd <= 2; ++nd)
#pragma unroll
                for (int nx = 0; nx <= 3; ++nx) {
                    const double f = (nt == 2 ? 2.0 : 1.0) * (nd == 2 ? 2.0 : 1.0) * (nx == 2 ? 2.0 : (nx == 3 ? 6.0 : 1.0));
                    out[(long long)(((i * 3 + nt) * 3 + nd) * 4 + nx) * ostride] = f * a.c[nt].c[nd].c[nx];
                }
    }
}

#if defined(__CUDACC__)
template<class M, int MAXN>
__global__ void __launch_bounds__(64) generic_ATrhoXi_kernel(const __grid_constant__ typename M::Params p, const int nc,
                                                             const DeviceBatch in, double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double x[MAXN];
        load_x<M, MAXN>(x, nc, in, k);
        generic_ATrhoXi_point<M, MAXN>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, out + k, npts);
    }
}

template<class M>
void generic_launch_ATrhoXi(const typename M::Params& p, int nc, const DeviceBatch& in, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    const int grid = grid_for(in.npts, block, sm_count() * 32);
    if constexpr (small_variants<M>::value) {
        if (nc == 1) generic_ATrhoXi_kernel<M, 1><<<grid, block, 0, st>>>(p, nc, in, out);
        else if (nc == 2) generic_ATrhoXi_kernel<M, 2><<<grid, block, 0, st>>>(p, nc, in, out);
        else if (nc == 3) generic_ATrhoXi_kernel<M, 3><<<grid, block, 0, st>>>(p, nc, in, out);
        else generic_ATrhoXi_kernel<M, M::kMaxC><<<grid, block, 0, st>>>(p, nc, in, out);
    } else {
        generic_ATrhoXi_kernel<M, M::kMaxC><<<grid, block, 0, st>>>(p, nc, in, out);
    }
    TQ_CUDA(cudaGetLastError());
}

#endif  // __CUDACC__

// ------------------------------------------------------------------------------------------------------------------
// Cross-composition derivatives (derivs.hpp:294-339, 423-465): x_i, x_j (, x_k) perturbed independently.
//   XiXj   table: out[((nt*2+nd)*3+ni)*3+nj][npts], nt, nd in 0..1, ni, nj in 0..2     (the reference exposes 15 of the 36)
//   XiXjXk table: out[(((nt*2+nd)*2+ni)*2+nj)*2+nk][npts], all orders 0..1              (the reference exposes 12 of the 32)
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN, int NX>
TQ_HD void generic_ATrhoXiXj_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin,
                                   const int i, const int j, double* out, const long long ostride) {
    using XT = J5<0, 0, NX, NX, 0>;
    using TT = J5<1, 0, 0, 0, 0>;
    using RT = J5<0, 1, 0, 0, 0>;
    const TT T = TT::reciprocal_scaled_variable(T0);
    RT rho;
    rho.c[0] = Tay<1, Tay<0, Tay<0, Tay<0, double>>>>::scaled_variable(rho0);
    XT x[MAXN];
    for (int k = 0; k < nc; ++k) {
        x[k] = XT(xin[k]);
        if (k == i) x[k].c[0].c[0].c[1].c[0].c[0] = 1.0;
        if (k == j) x[k].c[0].c[0].c[0].c[1].c[0] = 1.0;
    }
    const J5<1, 1, NX, NX, 0> a = embed<J5<1, 1, NX, NX, 0>>(M::template alphar<MAXN>(p, T, rho, x));
#pragma unroll
    for (int nt = 0; nt <= 1; ++nt)
#pragma unroll
        for (int nd = 0; nd <= 1; ++nd)
#pragma unroll
            for (int ni = 0; ni <= NX; ++ni)
#pragma unroll
                for (int nj = 0; nj <= NX; ++nj) {
                    const double f = (ni == 2 ? 2.0 : 1.0) * (nj == 2 ? 2.0 : 1.0);
                    out[(long long)(((nt * 2 + nd) * 3 + ni) * 3 + nj) * ostride] = f * a.c[nt].c[nd].c[ni].c[nj].c[0];
                }
}

template<class M, int MAXN>
TQ_HD void generic_ATrhoXiXjXk_point(const typename M::Params& p, const int nc, const double T0, const double rho0, const double* xin,
                                     const int i, const int j, const int kk, double* out, const long long ostride) {
    using XT = J5<0, 0, 1, 1, 1>;
    using TT = J5<1, 0, 0, 0, 0>;
    using RT = J5<0, 1, 0, 0, 0>;
    const TT T = TT::reciprocal_scaled_variable(T0);
    RT rho;
    rho.c[0] = Tay<1, Tay<0, Tay<0, Tay<0, double>>>>::scaled_variable(rho0);
    XT x[MAXN];
    for (int k = 0; k < nc; ++k) {
        x[k] = XT(xin[k]);
        if (k == i) x[k].c[0].c[0].c[1].c[0].c[0] = 1.0;
        if (k == j) x[k].c[0].c[0].c[0].c[1].c[0] = 1.0;
        if (k == kk) x[k].c[0].c[0].c[0].c[0].c[1] = 1.0;
    }
    const J5<1, 1, 1, 1, 1> a = embed<J5<1, 1, 1, 1, 1>>(M::template alphar<MAXN>(p, T, rho, x));
#pragma unroll
    for (int q = 0; q < 32; ++q)
        out[(long long)q * ostride] = a.c[(q >> 4) & 1].c[(q >> 3) & 1].c[(q >> 2) & 1].c[(q >> 1) & 1].c[q & 1];
}

#if defined(__CUDACC__)
template<class M>
__global__ void __launch_bounds__(64) generic_ATrhoXiXj_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in,
                                                               const int i, const int j, double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double x[M::kMaxC];
        load_x<M, M::kMaxC>(x, nc, in, k);
        generic_ATrhoXiXj_point<M, M::kMaxC, M::kCrossOrder>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, i, j, out + k, npts);
    }
}
template<class M>
__global__ void __launch_bounds__(64) generic_ATrhoXiXjXk_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in,
                                                                 const int i, const int j, const int kk, double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double x[M::kMaxC];
        load_x<M, M::kMaxC>(x, nc, in, k);
        generic_ATrhoXiXjXk_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), __ldg(in.rho + k), x, i, j, kk, out + k, npts);
    }
}
template<class M>
void generic_launch_ATrhoXiXj(const typename M::Params& p, int nc, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    generic_ATrhoXiXj_kernel<M><<<grid_for(in.npts, block, sm_count() * 32), block, 0, st>>>(p, nc, in, i, j, out);
    TQ_CUDA(cudaGetLastError());
}
template<class M>
void generic_launch_ATrhoXiXjXk(const typename M::Params& p, int nc, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    generic_ATrhoXiXjXk_kernel<M><<<grid_for(in.npts, block, sm_count() * 32), block, 0, st>>>(p, nc, in, i, j, k, out);
    TQ_CUDA(cudaGetLastError());
}
#endif  // __CUDACC__

// ------------------------------------------------------------------------------------------------------------------
// Isochoric family.  in.x holds the molar concentrations rhovec[nc][npts]; in.rho is unused.
// Psi^r(T, rhovec) = alphar(T, sum rhovec, rhovec / sum rhovec) * R(x) * T * sum rhovec      (derivs.hpp:1038-1042)
// Levels: 1 = unit perturbation of rho_i, 2 = unit perturbation of rho_j.
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN, int B, int C>
TQ_HD J3<0, B, C> psir_jet(const typename M::Params& p, int nc, double T0, const double* rv, int i, int j) {
    using XT = J3<0, B, C>;
    XT r[MAXN];
    XT rhotot(0.0);
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int k = 0; k < MAXN; ++k) {
        if (k < nc) {
            r[k] = XT(rv[k]);
            if (B >= 1 && k == i) r[k].c[0].c[B >= 1 ? 1 : 0].c[0] = 1.0;
            if (C >= 1 && k == j) r[k].c[0].c[0].c[C >= 1 ? 1 : 0] = 1.0;
            rhotot = rhotot + r[k];
        }
    }
    // mole fractions by DIVISION, as the reference forms them (derivs.hpp:1039): a reciprocal-multiply differs in the last
    // bit of x_i, which -ln(1 - b rho) amplifies by 1/(b rho) at low density
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int k = 0; k < MAXN; ++k)
        if (k < nc) r[k] = r[k] / rhotot;
    const J3<0, 0, 0> T(T0);
    const XT a = embed<XT>(M::template alphar<MAXN>(p, T, rhotot, r));
    return a * M::R(p, r) * T0 * rhotot;
}

template<class M, int MAXN>
TQ_HD void generic_isochoric_point(const typename M::Params& p, const int nc, const double T0, const double* rv, const IsochoricOut& o,
                                   const long long k, const long long npts) {
    double rhotot = 0.0;
#pragma unroll(MAXN <= 4 ? MAXN : 1)
    for (int i = 0; i < MAXN; ++i)
        if (i < nc) rhotot += rv[i];
    const bool want_h = o.hess != nullptr;
    double lnZ = 0.0, RT = 0.0;
    if (o.lnphi || o.phi) {
        // Z = 1 + Lambda_01 at x = rhovec/rhotot (derivs.hpp:1221-1225)
        using XT0 = J3<0, 0, 0>;
        XT0 x[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
        for (int i = 0; i < MAXN; ++i)
            if (i < nc) x[i] = XT0(rv[i] / rhotot);
        J3<0, 1, 0> rho;
        rho.c[0] = Tay<1, Tay<0, double>>::scaled_variable(rhotot);
        const J3<0, 1, 0> a = embed<J3<0, 1, 0>>(M::template alphar<MAXN>(p, XT0(T0), rho, x));
        lnZ = log(1.0 + a.c[0].c[1].c[0]);
        RT = base(M::R(p, x)) * T0;
    }
    for (int i = 0; i < nc; ++i) {
        if (want_h) {
            const J3<0, 2, 0> d = psir_jet<M, MAXN, 2, 0>(p, nc, T0, rv, i, -1);
            const double g = d.c[0].c[1].c[0];
            if (i == 0 && o.Psir) o.Psir[k] = d.c[0].c[0].c[0];
            if (o.grad) o.grad[(long long)i * npts + k] = g;
            if (o.lnphi) o.lnphi[(long long)i * npts + k] = g / RT - lnZ;
            if (o.phi) o.phi[(long long)i * npts + k] = exp(g / RT - lnZ);
            o.hess[(long long)(i * nc + i) * npts + k] = 2.0 * d.c[0].c[2].c[0];
            for (int j = i + 1; j < nc; ++j) {
                const J3<0, 1, 1> e = psir_jet<M, MAXN, 1, 1>(p, nc, T0, rv, i, j);
                const double h = e.c[0].c[1].c[1];
                o.hess[(long long)(i * nc + j) * npts + k] = h;
                o.hess[(long long)(j * nc + i) * npts + k] = h;
            }
        } else {
            const J3<0, 1, 0> d = psir_jet<M, MAXN, 1, 0>(p, nc, T0, rv, i, -1);
            const double g = d.c[0].c[1].c[0];
            if (i == 0 && o.Psir) o.Psir[k] = d.c[0].c[0].c[0];
            if (o.grad) o.grad[(long long)i * npts + k] = g;
            if (o.lnphi) o.lnphi[(long long)i * npts + k] = g / RT - lnZ;
            if (o.phi) o.phi[(long long)i * npts + k] = exp(g / RT - lnZ);
        }
    }
}

#if defined(__CUDACC__)
template<class M, int MAXN>
__global__ void __launch_bounds__(64) generic_isochoric_kernel(const __grid_constant__ typename M::Params p, const int nc,
                                                               const DeviceBatch in, const IsochoricOut o) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double rv[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
        for (int i = 0; i < MAXN; ++i)
            if (i < nc) rv[i] = __ldg(in.x + (long long)i * in.x_stride + k);
        generic_isochoric_point<M, MAXN>(p, nc, __ldg(in.T + k), rv, o, k, npts);
    }
}

template<class M>
void generic_launch_isochoric(const typename M::Params& p, int nc, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    const int grid = grid_for(in.npts, block, sm_count() * 32);
    if constexpr (small_variants<M>::value) {
        if (nc == 1) generic_isochoric_kernel<M, 1><<<grid, block, 0, st>>>(p, nc, in, o);
        else if (nc == 2) generic_isochoric_kernel<M, 2><<<grid, block, 0, st>>>(p, nc, in, o);
        else if (nc == 3) generic_isochoric_kernel<M, 3><<<grid, block, 0, st>>>(p, nc, in, o);
        else generic_isochoric_kernel<M, M::kMaxC><<<grid, block, 0, st>>>(p, nc, in, o);
    } else {
        generic_isochoric_kernel<M, M::kMaxC><<<grid, block, 0, st>>>(p, nc, in, o);
    }
    TQ_CUDA(cudaGetLastError());
}


#endif  // __CUDACC__

// ------------------------------------------------------------------------------------------------------------------
// Derived isochoric properties (IsoPropsOut): everything the reference composes from Psi^r, its gradient, Hessian-vector
// products, d/dT of the gradient and a few Lambda_nm (derivs.hpp:962-978, 1196-1201, 1264-1437, 1482-1494).
// Levels of the jet: 0 = T (UNIT seeded: these are derivatives w.r.t. T itself), 1 = rho_i, 2 = a direction v in rhovec space
// (v = rhovec gives the Hessian-vector product H rhovec in one sweep per component instead of Nc second-order sweeps).
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN, int A, int B, int C>
TQ_HD J3<A, B, C> psir_dir_jet(const typename M::Params& p, int nc, double T0, const double* rv, int i, const double* v) {
    using XT = J3<0, B, C>;
    XT r[MAXN];
    XT rhotot(0.0);
    for (int k = 0; k < nc; ++k) {
        r[k] = XT(rv[k]);
        if (B >= 1 && k == i) r[k].c[0].c[B >= 1 ? 1 : 0].c[0] = 1.0;
        if (C >= 1 && v) r[k].c[0].c[0].c[C >= 1 ? 1 : 0] = v[k];
        rhotot = rhotot + r[k];
    }
    for (int k = 0; k < nc; ++k) r[k] = r[k] / rhotot;
    using TT = J3<A, 0, 0>;
    const TT T = TT::unit_variable(T0);
    const J3<A, B, C> a = embed<J3<A, B, C>>(M::template alphar<MAXN>(p, T, rhotot, r));
    return (a * M::R(p, r)) * (T * rhotot);
}

template<class M, int MAXN>
TQ_HD void generic_isoprops_point(const typename M::Params& p, const int nc, const double T0, const double* rv, const IsoPropsOut& o,
                                  const long long k, const long long npts) {
    double rhotot = 0.0;
    for (int i = 0; i < nc; ++i) rhotot += rv[i];
    using X0 = J3<0, 0, 0>;
    X0 x[MAXN];
    double xd[MAXN];
    for (int i = 0; i < nc; ++i) { xd[i] = rv[i] / rhotot; x[i] = X0(xd[i]); }
    const double R = base(M::R(p, x));
    const double RT = R * T0;
    // Lambda_nm, n <= 1, m <= 2 at x = rhovec/rhotot
    double L[2][3];
    {
        const J3<1, 0, 0> T = J3<1, 0, 0>::reciprocal_scaled_variable(T0);
        J3<0, 2, 0> rho;
        rho.c[0] = Tay<2, Tay<0, double>>::scaled_variable(rhotot);
        const J3<1, 2, 0> a = embed<J3<1, 2, 0>>(M::template alphar<MAXN>(p, T, rho, x));
        for (int n = 0; n < 2; ++n)
            for (int m = 0; m < 3; ++m) L[n][m] = (m == 2 ? 2.0 : 1.0) * a.c[n].c[m].c[0];
    }
    const double Z = 1.0 + L[0][1];
    const double dZdrho = (L[0][1] + L[0][2]) / rhotot;
    const double dZdT_Z = L[1][1] / (-T0) / Z;
    if (o.pr) o.pr[k] = L[0][1] * rhotot * R * T0;
    if (o.splus) o.splus[k] = L[0][0] - L[1][0];
    double dPsirdT = 0.0, dot_Tgrad = 0.0;
    const bool want_hx = o.dlnphi_dx != nullptr;
    for (int i = 0; i < nc; ++i) {
        const J3<1, 1, 1> d = psir_dir_jet<M, MAXN, 1, 1, 1>(p, nc, T0, rv, i, rv);
        const double grad = d.c[0].c[1].c[0], Tgrad = d.c[1].c[1].c[0], Hrho = d.c[0].c[1].c[1];
        dPsirdT = d.c[1].c[0].c[0];
        dot_Tgrad += rv[i] * Tgrad;
        const long long q = (long long)i * npts + k;
        const double dpdrho_i = RT + Hrho;
        if (o.d2PsirdTdrhoi) o.d2PsirdTdrhoi[q] = Tgrad;
        if (o.dpdrhovec) o.dpdrhovec[q] = dpdrho_i;
        if (o.pmv) o.pmv[q] = (-rhotot * dpdrho_i) / (-(rhotot * rhotot) * RT * (1.0 + 2.0 * L[0][1] + L[0][2]));
        if (o.chempotVLE) o.chempotVLE[q] = grad + RT * (1.0 + log(rv[i]));
        if (o.dchempotdT) o.dchempotdT[q] = Tgrad + R * (1.0 + log(rv[i]));
        if (o.dlnphi_dT) o.dlnphi_dT[q] = 1.0 / RT * (Tgrad - 1.0 / T0 * grad) - dZdT_Z;
        const double dlr = 1.0 / RT * (Hrho / rhotot) - dZdrho / Z;
        if (o.dlnphi_drho) o.dlnphi_drho[q] = dlr;
        if (o.dlnphi_dv) o.dlnphi_dv[q] = dlr * (-(rhotot * rhotot));
    }
    if (o.dpdT) o.dpdT[k] = rhotot * R - dPsirdT + dot_Tgrad;
    if (want_hx) {
        // d ln phi_i / d x_j = rho/(RT) H_ij - (dZ/dx_j)/Z, dZ/dx_j = rho d2alphar/drho dx_j (derivs.hpp:1372-1393)
        const double f = rhotot / RT;
        for (int j = 0; j < nc; ++j) {
            using XT = J3<0, 0, 1>;
            XT xj[MAXN];
            for (int c = 0; c < nc; ++c) { xj[c] = XT(xd[c]); if (c == j) xj[c].c[0].c[0].c[1] = 1.0; }
            J3<0, 1, 0> rho;
            rho.c[0] = Tay<1, Tay<0, double>>::scaled_variable(rhotot);
            const J3<0, 1, 1> a = embed<J3<0, 1, 1>>(M::template alphar<MAXN>(p, J3<0, 0, 0>(T0), rho, xj));
            const double dZdx_Z = a.c[0].c[1].c[1] / Z;
            // column j of the Hessian: diagonal from a second-order sweep, the rest from mixed first-order sweeps
            for (int i = 0; i < nc; ++i) {
                double h;
                if (i == j) h = 2.0 * psir_jet<M, MAXN, 2, 0>(p, nc, T0, rv, i, -1).c[0].c[2].c[0];
                else h = psir_jet<M, MAXN, 1, 1>(p, nc, T0, rv, i, j).c[0].c[1].c[1];
                o.dlnphi_dx[(long long)(i * nc + j) * npts + k] = f * h - dZdx_Z;
            }
        }
    }
}

#if defined(__CUDACC__)
template<class M>
__global__ void __launch_bounds__(64) generic_isoprops_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in,
                                                              const IsoPropsOut o) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double rv[M::kMaxC];
        for (int i = 0; i < nc; ++i) rv[i] = __ldg(in.x + (long long)i * in.x_stride + k);
        generic_isoprops_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), rv, o, k, npts);
    }
}
template<class M>
void generic_launch_isoprops(const typename M::Params& p, int nc, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    generic_isoprops_kernel<M><<<grid_for(in.npts, block, sm_count() * 32), block, 0, st>>>(p, nc, in, o);
    TQ_CUDA(cudaGetLastError());
}
#endif  // __CUDACC__


// ------------------------------------------------------------------------------------------------------------------
// The same derived properties as THREE kernels with ONE jet shape each, for families whose model code is too large for NVVM to compile
// all seven instantiations into one kernel (SAFT-VR-Mie + association): each lives in its own translation unit.
//   part L  (J3<1,2,0>): Lambda_nm (n <= 1, m <= 2) -> pr, splus, and the six numbers into scratch L[6][npts]
//   part C  (J3<1,1,1>): per component gradient, d/dT of the gradient and the Hessian-vector product H rhovec -> every vector output, dp/dT
//   part X  (J3<0,1,1>): dZ/dx_j, combined with the Hessian the isochoric kernel wrote to scratch H[nc*nc][npts] -> d ln phi_i / d x_j
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN>
TQ_HD void generic_ip_L_point(const typename M::Params& p, const int nc, const double T0, const double* rv, const IsoPropsOut& o, double* Ls,
                              const long long k, const long long npts) {
    double rhotot = 0.0;
    for (int i = 0; i < nc; ++i) rhotot += rv[i];
    using X0 = J3<0, 0, 0>;
    X0 x[MAXN];
    for (int i = 0; i < nc; ++i) x[i] = X0(rv[i] / rhotot);
    const double R = base(M::R(p, x));
    const J3<1, 0, 0> T = J3<1, 0, 0>::reciprocal_scaled_variable(T0);
    J3<0, 2, 0> rho;
    rho.c[0] = Tay<2, Tay<0, double>>::scaled_variable(rhotot);
    const J3<1, 2, 0> a = embed<J3<1, 2, 0>>(M::template alphar<MAXN>(p, T, rho, x));
    double L[2][3];
    for (int n = 0; n < 2; ++n)
        for (int m = 0; m < 3; ++m) { L[n][m] = (m == 2 ? 2.0 : 1.0) * a.c[n].c[m].c[0]; Ls[(long long)(n * 3 + m) * npts + k] = L[n][m]; }
    if (o.pr) o.pr[k] = L[0][1] * rhotot * R * T0;
    if (o.splus) o.splus[k] = L[0][0] - L[1][0];
}

template<class M, int MAXN>
TQ_HD void generic_ip_C_point(const typename M::Params& p, const int nc, const double T0, const double* rv, const IsoPropsOut& o, const double* Ls,
                              const long long k, const long long npts) {
    double rhotot = 0.0;
    for (int i = 0; i < nc; ++i) rhotot += rv[i];
    using X0 = J3<0, 0, 0>;
    X0 x[MAXN];
    for (int i = 0; i < nc; ++i) x[i] = X0(rv[i] / rhotot);
    const double R = base(M::R(p, x));
    const double RT = R * T0;
    const double L01 = Ls[1 * npts + k], L02 = Ls[2 * npts + k], L11 = Ls[4 * npts + k];
    const double Z = 1.0 + L01;
    const double dZdrho = (L01 + L02) / rhotot;
    const double dZdT_Z = L11 / (-T0) / Z;
    double dPsirdT = 0.0, dot_Tgrad = 0.0;
    for (int i = 0; i < nc; ++i) {
        const J3<1, 1, 1> d = psir_dir_jet<M, MAXN, 1, 1, 1>(p, nc, T0, rv, i, rv);
        const double grad = d.c[0].c[1].c[0], Tgrad = d.c[1].c[1].c[0], Hrho = d.c[0].c[1].c[1];
        dPsirdT = d.c[1].c[0].c[0];
        dot_Tgrad += rv[i] * Tgrad;
        const long long q = (long long)i * npts + k;
        const double dpdrho_i = RT + Hrho;
        if (o.d2PsirdTdrhoi) o.d2PsirdTdrhoi[q] = Tgrad;
        if (o.dpdrhovec) o.dpdrhovec[q] = dpdrho_i;
        if (o.pmv) o.pmv[q] = (-rhotot * dpdrho_i) / (-(rhotot * rhotot) * RT * (1.0 + 2.0 * L01 + L02));
        if (o.chempotVLE) o.chempotVLE[q] = grad + RT * (1.0 + log(rv[i]));
        if (o.dchempotdT) o.dchempotdT[q] = Tgrad + R * (1.0 + log(rv[i]));
        if (o.dlnphi_dT) o.dlnphi_dT[q] = 1.0 / RT * (Tgrad - 1.0 / T0 * grad) - dZdT_Z;
        const double dlr = 1.0 / RT * (Hrho / rhotot) - dZdrho / Z;
        if (o.dlnphi_drho) o.dlnphi_drho[q] = dlr;
        if (o.dlnphi_dv) o.dlnphi_dv[q] = dlr * (-(rhotot * rhotot));
    }
    if (o.dpdT) o.dpdT[k] = rhotot * R - dPsirdT + dot_Tgrad;
}

template<class M, int MAXN>
TQ_HD void generic_ip_X_point(const typename M::Params& p, const int nc, const double T0, const double* rv, const IsoPropsOut& o, const double* Ls,
                              const double* H, const long long k, const long long npts) {
    double rhotot = 0.0;
    for (int i = 0; i < nc; ++i) rhotot += rv[i];
    using X0 = J3<0, 0, 0>;
    X0 x0[MAXN];
    for (int i = 0; i < nc; ++i) x0[i] = X0(rv[i] / rhotot);
    const double RT = base(M::R(p, x0)) * T0;
    const double Z = 1.0 + Ls[1 * npts + k];
    const double f = rhotot / RT;
    for (int j = 0; j < nc; ++j) {
        using XT = J3<0, 0, 1>;
        XT xj[MAXN];
        for (int c = 0; c < nc; ++c) { xj[c] = XT(rv[c] / rhotot); if (c == j) xj[c].c[0].c[0].c[1] = 1.0; }
        J3<0, 1, 0> rho;
        rho.c[0] = Tay<1, Tay<0, double>>::scaled_variable(rhotot);
        const J3<0, 1, 1> a = embed<J3<0, 1, 1>>(M::template alphar<MAXN>(p, J3<0, 0, 0>(T0), rho, xj));
        const double dZdx_Z = a.c[0].c[1].c[1] / Z;
        for (int i = 0; i < nc; ++i) o.dlnphi_dx[(long long)(i * nc + j) * npts + k] = f * H[(long long)(i * nc + j) * npts + k] - dZdx_Z;
    }
}

#if defined(__CUDACC__)
template<class M, int PART>
__global__ void __launch_bounds__(64) generic_ip_part_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in,
                                                             const IsoPropsOut o, double* __restrict__ Ls, const double* __restrict__ H) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double rv[M::kMaxC];
        for (int i = 0; i < nc; ++i) rv[i] = __ldg(in.x + (long long)i * in.x_stride + k);
        if constexpr (PART == 0) generic_ip_L_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), rv, o, Ls, k, npts);
        else if constexpr (PART == 1) generic_ip_C_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), rv, o, Ls, k, npts);
        else generic_ip_X_point<M, M::kMaxC>(p, nc, __ldg(in.T + k), rv, o, Ls, H, k, npts);
    }
}
template<class M, int PART>
void generic_launch_ip_part(const typename M::Params& p, int nc, const DeviceBatch& in, const IsoPropsOut& o, double* Ls, const double* H, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    generic_ip_part_kernel<M, PART><<<grid_for(in.npts, block, sm_count() * 32), block, 0, st>>>(p, nc, in, o, Ls, H);
    TQ_CUDA(cudaGetLastError());
}
#endif  // __CUDACC__

// ------------------------------------------------------------------------------------------------------------------
// get_Psir_sigma_derivs (derivs.hpp:1496-1509): Psi^r(T, rhovec + sigma v) to fourth order in sigma, one Tay<4> sweep
// (the reference uses autodiff::Real<4> the same way): out[k][npts] = d^k Psir / dsigma^k, k = 0..4
// ------------------------------------------------------------------------------------------------------------------
template<class M, int MAXN>
TQ_HD void generic_sigma_point(const typename M::Params& p, const int nc, const double T0, const double* rv, const double* v, double* out,
                               const long long ostride) {
    const J3<0, 0, 4> a = psir_dir_jet<M, MAXN, 0, 0, 4>(p, nc, T0, rv, -1, v);
    double f = 1.0;
#pragma unroll
    for (int k = 0; k <= 4; ++k) {
        if (k >= 2) f *= double(k);
        out[(long long)k * ostride] = f * a.c[0].c[0].c[k];
    }
}

#if defined(__CUDACC__)
template<class M, int MAXN>
__global__ void __launch_bounds__(64) generic_sigma_kernel(const __grid_constant__ typename M::Params p, const int nc, const DeviceBatch in,
                                                           const double* __restrict__ vdir, double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        double rv[MAXN], v[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
        for (int i = 0; i < MAXN; ++i)
            if (i < nc) { rv[i] = __ldg(in.x + (long long)i * in.x_stride + k); v[i] = __ldg(vdir + (long long)i * npts + k); }
        generic_sigma_point<M, MAXN>(p, nc, __ldg(in.T + k), rv, v, out + k, npts);
    }
}

template<class M>
void generic_launch_sigma(const typename M::Params& p, int nc, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    const int block = 64;
    const int grid = grid_for(in.npts, block, sm_count() * 32);
    if constexpr (small_variants<M>::value) {
        if (nc <= 2) generic_sigma_kernel<M, 2><<<grid, block, 0, st>>>(p, nc, in, v, out);
        else generic_sigma_kernel<M, M::kMaxC><<<grid, block, 0, st>>>(p, nc, in, v, out);
    } else {
        generic_sigma_kernel<M, M::kMaxC><<<grid, block, 0, st>>>(p, nc, in, v, out);
    }
    TQ_CUDA(cudaGetLastError());
}
#endif  // __CUDACC__

}  // namespace tq
