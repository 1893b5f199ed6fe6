// TEST INFRASTRUCTURE ONLY -- the "truth" leg of the parity oracle: the reference's OWN model templates
// (/root/reference/include/teqp/models/**) evaluated in IEEE binary128 (__float128, libquadmath) instead of double, so that a
// disagreement between libteqpb200 and the double-precision oracle (oracle/ref_capi.cpp) can be adjudicated: which side is closer
// to the exact value of the reference's formulas?  (VERDICT r01 "What's weak" #1.)
//
// Derivatives come from ONE sweep of the model in nested truncated-Taylor arithmetic Real<NT, Real<ND, Real<NX, __float128>>>
// (oracle/shim/autodiff_shim.hpp: Real<N,T> is templated on its coefficient type); rounding noise is ~1e-34 * (family scale), i.e.
// 18 orders below double, whatever the derivative order.  The drivers below restate what the reference's drivers define
// (/root/reference/include/teqp/derivs.hpp:150-238 get_Agenxy, :247-259 get_ATrhoXi, :543-598 get_Agen0n/get_Agenn0,
// :1031-1069 Psir and its Hessian), not how they compute it (one nested-dual sweep per entry).  Where the reference's own driver
// is templated on the scalar type (get_Ar0n<N>: `autodiff::Real<Nderiv, Scalar>`, derivs.hpp:548) it is ALSO called directly with
// Scalar = __float128 (truth_Ar0n_refdriver) and must agree with the nested-Taylor route to ~1e-30.
//
// Results are rounded to double on output: comparisons are made in units of eps(double) * scale, so the 1.1e-16 relative rounding
// of the truth itself is below what is asserted.
#include "teqp/derivs.hpp"
#include "teqp/models/vdW.hpp"
#include "teqp/models/cubics/simple_cubics.hpp"
#include "teqp/models/GERG/GERG.hpp"
#include "teqp/models/pcsaft.hpp"
#include "teqp/models/saftvrmie.hpp"
#include "teqp/models/saft/genericsaft.hpp"
#include "teqp/models/multifluid.hpp"

#include <atomic>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <variant>
#include <vector>

using Q = __float128;
using namespace teqp;
using autodiff::detail::Real;

namespace {

template<class J> using Arr = Eigen::Array<J, Eigen::Dynamic, 1>;

using PRModel = decltype(make_canonicalPR(nlohmann::json{}));
using CubicModel = decltype(make_generalizedcubic(nlohmann::json{}));
using PCSAFTModel = decltype(PCSAFT::PCSAFTfactory(nlohmann::json{}));
using MieModel = decltype(SAFTVRMie::SAFTVRMiefactory(nlohmann::json{}));
using GSAFTModel = saft::genericsaft::GenericSAFT;
using MFModel = decltype(multifluidfactory(nlohmann::json{}));
using AnyModel = std::variant<vdWEOS1, vdWEOS<double>, PRModel, PCSAFTModel, MieModel, GSAFTModel, GERG2004::GERG2004ResidualModel,
                              GERG2008::GERG2008ResidualModel, MFModel>;

struct TruthModel { std::unique_ptr<AnyModel> m; };

std::unique_ptr<AnyModel> build(const nlohmann::json& j) {
    const std::string kind = j.at("kind").get<std::string>();
    const auto& spec = j.at("model");
    if (kind == "vdW1") return std::make_unique<AnyModel>(vdWEOS1(spec.at("a"), spec.at("b")));
    if (kind == "vdW") return std::make_unique<AnyModel>(vdWEOS<double>(spec.at("Tcrit / K"), spec.at("pcrit / Pa")));
    if (kind == "PR") return std::make_unique<AnyModel>(make_canonicalPR(spec));
    if (kind == "SRK") return std::make_unique<AnyModel>(make_canonicalSRK(spec));
    if (kind == "cubic") return std::make_unique<AnyModel>(make_generalizedcubic(spec));
    if (kind == "PCSAFT") return std::make_unique<AnyModel>(PCSAFT::PCSAFTfactory(spec));
    if (kind == "SAFT-VR-Mie") return std::make_unique<AnyModel>(SAFTVRMie::SAFTVRMiefactory(spec));
    if (kind == "genericSAFT") return std::make_unique<AnyModel>(GSAFTModel(spec));
    if (kind == "GERG2004resid") return std::make_unique<AnyModel>(GERG2004::GERG2004ResidualModel(spec.at("names").get<std::vector<std::string>>()));
    if (kind == "GERG2008resid") return std::make_unique<AnyModel>(GERG2008::GERG2008ResidualModel(spec.at("names").get<std::vector<std::string>>()));
    if (kind == "multifluid") return std::make_unique<AnyModel>(multifluidfactory(spec));
    throw std::invalid_argument("truth oracle: don't understand \"kind\" of: " + kind);
}

void put_err(const std::string& msg, char* err, int errlen) {
    if (err && errlen > 0) { std::strncpy(err, msg.c_str(), errlen - 1); err[errlen - 1] = 0; }
}

template<typename F>
int parallel_points(long long npts, int nthreads, char* err, int errlen, F&& fn) {
    if (nthreads < 1) nthreads = 1;
    if ((long long)nthreads > npts) nthreads = (int)std::max<long long>(1, npts);
    std::vector<std::thread> th;
    std::atomic<int> failed{0};
    std::string msg;
    std::mutex mtx;
    std::atomic<long long> next{0};
    auto body = [&]() {
        try { for (long long k = next++; k < npts; k = next++) fn(k); }
        catch (const std::exception& e) { std::lock_guard<std::mutex> g(mtx); if (!failed) { failed = 1; msg = e.what(); } }
    };
    for (int t = 0; t < nthreads; ++t) th.emplace_back(body);
    for (auto& t : th) t.join();
    if (failed) { put_err(msg, err, errlen); return 9999; }
    return 0;
}

Q qfact(int n) { Q f = 1; for (int i = 2; i <= n; ++i) f *= i; return f; }

// ---- nested Taylor numbers: level 0 = 1/T, level 1 = rho, level 2 = one composition direction ---------------------------------
template<int A, int B, int C> using J3 = Real<A, Real<B, Real<C, Q>>>;
template<int A, int B, int C> J3<A, B, C> jconst(Q v) { J3<A, B, C> r; r.c[0].c[0].c[0] = v; return r; }

// table[(nt*(ND+1)+nd)*(NX+1)+nx] = (1/T)^nt rho^nd d^(nt+nd+nx) alphar / d(1/T)^nt drho^nd dx_i^nx   (ix < 0: no composition direction)
template<int NT, int ND, int NX, class Model>
void taylor_table(const Model& model, double T0, double rho0, const double* x, int nc, long long xstride, int ix, double* out, long long ostride) {
    using J = J3<NT, ND, NX>;
    const Q Tr0 = Q(1) / Q(T0);
    J Trecip = jconst<NT, ND, NX>(Tr0);
    if constexpr (NT >= 1) Trecip.c[1].c[0].c[0] = Tr0;            // scaled seed: (1/T)(s) = (1/T0)(1+s)
    J rho = jconst<NT, ND, NX>(Q(rho0));
    if constexpr (ND >= 1) rho.c[0].c[1].c[0] = Q(rho0);           // rho(r) = rho0 (1+r)
    Arr<J> z(nc);
    for (int i = 0; i < nc; ++i) {
        z[i] = jconst<NT, ND, NX>(Q(x[i * xstride]));
        if constexpr (NX >= 1) if (i == ix) z[i].c[0].c[0].c[1] = Q(1);   // unit seed: derivatives w.r.t. x_i itself
    }
    const J Tj = Q(1) / Trecip;
    const J a = model.alphar(Tj, rho, z);
    for (int nt = 0; nt <= NT; ++nt)
        for (int nd = 0; nd <= ND; ++nd)
            for (int nx = 0; nx <= NX; ++nx)
                out[(long long)((nt * (ND + 1) + nd) * (NX + 1) + nx) * ostride] = (double)(a.c[nt].c[nd].c[nx] * qfact(nt) * qfact(nd) * qfact(nx));
}

// Psi^r(T, rhovec) = alphar(T, rho, rhovec/rho) R(x) T rho  (derivs.hpp:1038-1042) with rho_i -> rho_i + s, rho_j -> rho_j + r
template<int A, int B, class Model>
Real<A, Real<B, Q>> psir_jet(const Model& model, double T0, const double* rv, int nc, long long stride, int i, int j) {
    using J = Real<A, Real<B, Q>>;
    Arr<J> r(nc);
    J rhotot; rhotot.c[0].c[0] = 0;
    for (int k = 0; k < nc; ++k) {
        J v; v.c[0].c[0] = Q(rv[k * stride]);
        if (A >= 1 && k == i) v.c[A >= 1 ? 1 : 0].c[0] = 1;
        if (B >= 1 && k == j) v.c[0].c[B >= 1 ? 1 : 0] = 1;
        r[k] = v;
        rhotot = rhotot + v;
    }
    Arr<J> z(nc);
    for (int k = 0; k < nc; ++k) z[k] = r[k] / rhotot;
    J Tj; Tj.c[0].c[0] = Q(T0);
    const J a = model.alphar(Tj, rhotot, z);
    return a * model.R(z) * Tj * rhotot;
}

}  // namespace

#define TRUTH_TRY try {
#define TRUTH_CATCH } catch (const std::exception& e) { put_err(e.what(), err, errlen); return 9999; }

extern "C" {

void* truth_build_model(const char* json, char* err, int errlen) {
    try {
        auto m = new TruthModel();
        m->m = build(nlohmann::json::parse(json));
        return m;
    } catch (const std::exception& e) { put_err(e.what(), err, errlen); return nullptr; }
}
void truth_free_model(void* h) { delete static_cast<TruthModel*>(h); }

// out[(NT+1)*(ND+1)][npts], Lambda_nm for n <= NT, m <= ND; compiled shapes: (0,6), (4,0), (2,4)
int truth_ar_table(void* h, int NT, int ND, long long npts, const double* T, const double* rho, const double* x, int ncomp, double* out,
                   int nthreads, char* err, int errlen) {
    TRUTH_TRY
    const AnyModel& am = *static_cast<TruthModel*>(h)->m;
    return std::visit([&](const auto& model) -> int {
        return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
            if (NT == 0 && ND == 6) taylor_table<0, 6, 0>(model, T[k], rho[k], x + k, ncomp, npts, -1, out + k, npts);
            else if (NT == 4 && ND == 0) taylor_table<4, 0, 0>(model, T[k], rho[k], x + k, ncomp, npts, -1, out + k, npts);
            else if (NT == 2 && ND == 4) taylor_table<2, 4, 0>(model, T[k], rho[k], x + k, ncomp, npts, -1, out + k, npts);
            else throw std::invalid_argument("truth_ar_table: shapes (0,6), (4,0), (2,4) only");
        });
    }, am);
    TRUTH_CATCH
}

// the reference's OWN driver get_Ar0n<N> (derivs.hpp:543-553) instantiated with Scalar = __float128: out[(nmax+1)][npts], nmax = 4 or 6
int truth_Ar0n_refdriver(void* h, int nmax, long long npts, const double* T, const double* rho, const double* x, int ncomp, double* out,
                         int nthreads, char* err, int errlen) {
    TRUTH_TRY
    const AnyModel& am = *static_cast<TruthModel*>(h)->m;
    return std::visit([&](const auto& model) -> int {
        using M = std::decay_t<decltype(model)>;
        using tdx = TDXDerivatives<M, Q, Arr<Q>>;
        return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
            Arr<Q> z(ncomp);
            for (int i = 0; i < ncomp; ++i) z[i] = Q(x[(long long)i * npts + k]);
            if (nmax == 4) { auto v = tdx::template get_Ar0n<4>(model, Q(T[k]), Q(rho[k]), z); for (int n = 0; n <= 4; ++n) out[(long long)n * npts + k] = (double)v[n]; }
            else if (nmax == 6) { auto v = tdx::template get_Ar0n<6>(model, Q(T[k]), Q(rho[k]), z); for (int n = 0; n <= 6; ++n) out[(long long)n * npts + k] = (double)v[n]; }
            else throw std::invalid_argument("truth_Ar0n_refdriver: nmax 4 or 6");
        });
    }, am);
    TRUTH_CATCH
}

// get_ATrhoXi table, out[((i*3+nt)*3+nd)*4+nx][npts] (same layout as libteqpb200's get_ATrhoXi_table_batch)
int truth_ATrhoXi_table(void* h, long long npts, const double* T, const double* rho, const double* x, int ncomp, double* out, int nthreads,
                        char* err, int errlen) {
    TRUTH_TRY
    const AnyModel& am = *static_cast<TruthModel*>(h)->m;
    return std::visit([&](const auto& model) -> int {
        return parallel_points(npts * ncomp, nthreads, err, errlen, [&](long long q) {
            const long long k = q / ncomp; const int i = (int)(q % ncomp);
            taylor_table<2, 2, 3>(model, T[k], rho[k], x + k, ncomp, npts, i, out + (long long)i * 36 * npts + k, npts);
        });
    }, am);
    TRUTH_CATCH
}

// Psi^r, gradient [ncomp][npts], Hessian [ncomp*ncomp][npts] (any may be null)
int truth_Psir_fgradHessian(void* h, long long npts, const double* T, const double* rhovec, int ncomp, double* f, double* g, double* H,
                            int nthreads, char* err, int errlen) {
    TRUTH_TRY
    const AnyModel& am = *static_cast<TruthModel*>(h)->m;
    return std::visit([&](const auto& model) -> int {
        const long long nwork = H ? npts * ncomp : npts * ncomp;
        return parallel_points(nwork, nthreads, err, errlen, [&](long long q) {
            const long long k = q / ncomp; const int i = (int)(q % ncomp);
            const auto d = psir_jet<2, 0>(model, T[k], rhovec + k, ncomp, npts, i, -1);
            if (f && i == 0) f[k] = (double)d.c[0].c[0];
            if (g) g[(long long)i * npts + k] = (double)d.c[1].c[0];
            if (H) {
                H[(long long)(i * ncomp + i) * npts + k] = (double)(2 * d.c[2].c[0]);
                for (int j = i + 1; j < ncomp; ++j) {
                    const auto e = psir_jet<1, 1>(model, T[k], rhovec + k, ncomp, npts, i, j);
                    const double v = (double)e.c[1].c[1];
                    H[(long long)(i * ncomp + j) * npts + k] = v;
                    H[(long long)(j * ncomp + i) * npts + k] = v;
                }
            }
        });
    }, am);
    TRUTH_CATCH
}

// alphar itself at quad precision for arbitrary (T, rho, x) given as double-double pairs hi + lo -- lets a test take __float128 central
// differences of the reference's alphar that do not go through ANY automatic differentiation (pins the AD rules of the shim)
int truth_alphar_dd(void* h, long long npts, const double* Thi, const double* Tlo, const double* rhohi, const double* rholo, const double* xhi,
                    const double* xlo, int ncomp, double* out_hi, double* out_lo, int nthreads, char* err, int errlen) {
    TRUTH_TRY
    const AnyModel& am = *static_cast<TruthModel*>(h)->m;
    return std::visit([&](const auto& model) -> int {
        return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
            Arr<Q> z(ncomp);
            for (int i = 0; i < ncomp; ++i) z[i] = Q(xhi[(long long)i * npts + k]) + Q(xlo[(long long)i * npts + k]);
            const Q a = model.alphar(Q(Thi[k]) + Q(Tlo[k]), Q(rhohi[k]) + Q(rholo[k]), z);
            out_hi[k] = (double)a;
            out_lo[k] = (double)(a - Q(out_hi[k]));
        });
    }, am);
    TRUTH_CATCH
}

}  // extern "C"
