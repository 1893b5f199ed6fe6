// TEST INFRASTRUCTURE ONLY -- the "reference" leg of the parity oracle.  Never linked into, imported by, or called from the
// product library (teqp_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// What this is: a flat C API, batched over state points with std::thread, in front of the reference's OWN type-erased
// interface teqp::cppinterface::AbstractModel (/root/reference/include/teqp/cpp/teqpcpp.hpp:94-191), whose implementation
// DerivativeAdapter<> (/root/reference/include/teqp/cpp/deriv_adapter.hpp:106-248) and model factories
// (/root/reference/interface/CPP/model_factory_{GERG,pcsaft,SAFTVRMie,genericsaft}.cpp) are compiled from the sources where
// they lie.  Every value returned by this file was computed by the reference's model templates and derivative drivers
// (/root/reference/include/teqp/derivs.hpp); the only restated piece is the forward-mode AD engine (oracle/shim/), because
// autodiff @ cb2d6e8b is absent from the reference tree (/root/reference/.SUBMODULES.json:44-50).
//
// The kind -> factory map below restates the subset of /root/reference/interface/CPP/teqp_impl_factory.cpp:54-102 that is in
// scope (that TU itself cannot be used: it references all 35 factories and the cmake-generated schema library).
#include "teqp/cpp/teqpcpp.hpp"
#include "teqp/cpp/deriv_adapter.hpp"
#include "teqp/models/vdW.hpp"
#include "teqp/models/cubics/simple_cubics.hpp"
#include "teqp/models/cubics/advancedmixing_cubics.hpp"
#include "teqp/models/GERG/GERG.hpp"
#include "teqp/models/pcsaft.hpp"
#include "teqp/models/multifluid.hpp"
#include "teqp/ideal_eosterms.hpp"
#include "teqp/algorithms/VLE_pure.hpp"

#include <thread>
#include <vector>
#include <cstring>
#include <string>
#include <functional>
#include <atomic>

namespace teqp { namespace cppinterface {
    std::unique_ptr<AbstractModel> make_SAFTVRMie(const nlohmann::json &);
    std::unique_ptr<AbstractModel> make_PCSAFT(const nlohmann::json &);
    std::unique_ptr<AbstractModel> make_genericSAFT(const nlohmann::json &);
    std::unique_ptr<AbstractModel> make_GERG2004resid(const nlohmann::json &);
    std::unique_ptr<AbstractModel> make_GERG2008resid(const nlohmann::json &);
    std::unique_ptr<AbstractModel> make_multifluid(const nlohmann::json &);   // interface/CPP/model_factory_multifluid.cpp
    std::unique_ptr<AbstractModel> make_GERG2004idealgas(const nlohmann::json &);   // interface/CPP/model_factory_GERG.cpp:19-24
    std::unique_ptr<AbstractModel> make_GERG2008idealgas(const nlohmann::json &);
    std::unique_ptr<AbstractModel> make_IdealHelmholtz(const nlohmann::json &);     // interface/CPP/model_factory_IdealHelmholtz.cpp
}}

using namespace teqp;
using namespace teqp::cppinterface;
using teqp::cppinterface::adapter::make_owned;

namespace {

struct RefModel {
    std::unique_ptr<AbstractModel> am;
    std::string kind;
};

std::unique_ptr<AbstractModel> build(const nlohmann::json& j, std::string& kind) {
    kind = j.at("kind").get<std::string>();
    const auto& spec = j.at("model");
    if (kind == "vdW1") return make_owned(vdWEOS1(spec.at("a"), spec.at("b")));
    if (kind == "vdW") return make_owned(vdWEOS<double>(spec.at("Tcrit / K"), spec.at("pcrit / Pa")));
    if (kind == "PR") return make_owned(make_canonicalPR(spec));
    if (kind == "SRK") return make_owned(make_canonicalSRK(spec));
    if (kind == "cubic") return make_owned(make_generalizedcubic(spec));
    if (kind == "PCSAFT") return make_PCSAFT(spec);
    if (kind == "SAFT-VR-Mie") return make_SAFTVRMie(spec);
    if (kind == "genericSAFT") return make_genericSAFT(spec);
    if (kind == "GERG2004resid") return make_GERG2004resid(spec);
    if (kind == "GERG2008resid") return make_GERG2008resid(spec);
    if (kind == "multifluid") return make_multifluid(spec);
    if (kind == "GERG2004idealgas") return make_GERG2004idealgas(spec);
    if (kind == "GERG2008idealgas") return make_GERG2008idealgas(spec);
    if (kind == "IdealHelmholtz") return make_IdealHelmholtz(spec);
    if (kind == "QCPRAasen") return make_owned(QuantumCorrectedPR(spec));   // interface/CPP/teqp_impl_factory.cpp:60
    if (kind == "advancedPRaEres") return make_owned(make_AdvancedPRaEres(spec));   // interface/CPP/teqp_impl_factory.cpp:59
    if (kind == "RKPRCismondi2005") return make_owned(RKPRCismondi2005(spec));   // interface/CPP/teqp_impl_factory.cpp:61
    throw std::invalid_argument("Don't understand \"kind\" of: " + kind);
}

}  // namespace
// C++ entry for interop/test_b200_model.cpp: the reference's AbstractModel for a spec (what teqp::cppinterface::make_model returns for it)
std::unique_ptr<teqp::cppinterface::AbstractModel> teqp_ref_oracle_make_model(const nlohmann::json& j) { std::string kind; return build(j, kind); }
namespace {

void put_err(const std::string& msg, char* err, int errlen) {
    if (err && errlen > 0) { std::strncpy(err, msg.c_str(), errlen - 1); err[errlen - 1] = 0; }
}

// static contiguous partition of [0,npts) over nthreads; fn(k) evaluates point k.  First exception wins.
template<typename F>
int parallel_points(long long npts, int nthreads, char* err, int errlen, F&& fn) {
    if (nthreads < 1) nthreads = 1;
    if ((long long)nthreads > npts) nthreads = (int)std::max<long long>(1, npts);
    std::vector<std::thread> th;
    std::atomic<int> failed{0};
    std::string msg;
    std::mutex mtx;
    auto body = [&](long long lo, long long hi) {
        try { for (long long k = lo; k < hi; ++k) fn(k); }
        catch (const std::exception& e) { std::lock_guard<std::mutex> g(mtx); if (!failed) { failed = 1; msg = e.what(); } }
    };
    const long long chunk = (npts + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
        long long lo = t * chunk, hi = std::min(npts, lo + chunk);
        if (lo >= hi) break;
        if (nthreads == 1) body(lo, hi); else th.emplace_back(body, lo, hi);
    }
    for (auto& t : th) t.join();
    if (failed) { put_err(msg, err, errlen); return 9999; }
    return 0;
}

inline void gather(const double* soa, long long npts, long long k, int n, Eigen::ArrayXd& z) {
    for (int i = 0; i < n; ++i) z[i] = soa[(long long)i * npts + k];
}

} // namespace

#define REF_TRY try {
#define REF_CATCH } catch (const std::exception& e) { put_err(e.what(), err, errlen); return 9999; }

extern "C" {

void* ref_build_model(const char* json, char* err, int errlen) {
    try {
        auto m = new RefModel();
        m->am = build(nlohmann::json::parse(json), m->kind);
        return m;
    } catch (const std::exception& e) { put_err(e.what(), err, errlen); return nullptr; }
}
void ref_free_model(void* h) { delete static_cast<RefModel*>(h); }

int ref_get_R(void* h, const double* z, int ncomp, double* out, char* err, int errlen) {
    REF_TRY
    Eigen::ArrayXd zz = Eigen::Map<const Eigen::ArrayXd>(z, ncomp);
    *out = static_cast<RefModel*>(h)->am->get_R(zz);
    return 0;
    REF_CATCH
}

// out[(nmax+1)][npts]: AbstractModel::get_Ar0{nmax}n (teqpcpp.hpp:111-113) per point
int ref_get_Ar0n(void* h, int nmax, long long npts, const double* T, const double* rho, const double* x, int ncomp,
                 double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    if (nmax < 1 || nmax > 6) { put_err("nmax must be 1..6", err, errlen); return 1; }
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        Eigen::ArrayXd v;
        switch (nmax) {
            case 1: v = am.get_Ar01n(T[k], rho[k], z); break;
            case 2: v = am.get_Ar02n(T[k], rho[k], z); break;
            case 3: v = am.get_Ar03n(T[k], rho[k], z); break;
            case 4: v = am.get_Ar04n(T[k], rho[k], z); break;
            case 5: v = am.get_Ar05n(T[k], rho[k], z); break;
            default: v = am.get_Ar06n(T[k], rho[k], z); break;
        }
        for (int n = 0; n <= nmax; ++n) out[(long long)n * npts + k] = v[n];
    });
}

// out[(nmax+1)][npts]: AbstractModel::get_Ar{nmax}0n (teqpcpp.hpp:115-117)
int ref_get_Arn0(void* h, int nmax, long long npts, const double* T, const double* rho, const double* x, int ncomp,
                 double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    if (nmax < 1 || nmax > 4) { put_err("nmax must be 1..4", err, errlen); return 1; }
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        Eigen::ArrayXd v;
        switch (nmax) {
            case 1: v = am.get_Ar10n(T[k], rho[k], z); break;
            case 2: v = am.get_Ar20n(T[k], rho[k], z); break;
            case 3: v = am.get_Ar30n(T[k], rho[k], z); break;
            default: v = am.get_Ar40n(T[k], rho[k], z); break;
        }
        for (int n = 0; n <= nmax; ++n) out[(long long)n * npts + k] = v[n];
    });
}

// compile-time-dispatched single derivative get_Ar{NT}{ND} (teqpcpp.hpp:107-109), NT in 0..2, ND in 0..4
int ref_get_ArNM(void* h, int NT, int ND, long long npts, const double* T, const double* rho, const double* x, int ncomp,
                 double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    using Fn = double (AbstractModel::*)(const double, const double, const REArrayd&) const;
    Fn fn = nullptr;
#define X(i,j) if (NT == i && ND == j) fn = &AbstractModel::get_Ar ## i ## j;
    ARXY_args
#undef X
    if (!fn) { put_err("unsupported (NT,ND) for get_ArNM", err, errlen); return 1; }
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        out[k] = (am.*fn)(T[k], rho[k], z);
    });
}

// runtime-dispatched get_Arxy (teqpcpp.hpp:104; derivs.hpp:627-682), what the reference C ABI calls (teqpc.cpp:105-117)
int ref_get_Arxy(void* h, int NT, int ND, long long npts, const double* T, const double* rho, const double* x, int ncomp,
                 double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        out[k] = am.get_Arxy(NT, ND, T[k], rho[k], z);
    });
}

int ref_get_ATrhoXi(void* h, int NT, int ND, int i, int NXi, long long npts, const double* T, const double* rho, const double* x,
                    int ncomp, double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        out[k] = am.get_ATrhoXi(T[k], NT, rho[k], ND, z, i, NXi);
    });
}
int ref_get_ATrhoXiXj(void* h, int NT, int ND, int i, int NXi, int j, int NXj, long long npts, const double* T, const double* rho,
                      const double* x, int ncomp, double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        out[k] = am.get_ATrhoXiXj(T[k], NT, rho[k], ND, z, i, NXi, j, NXj);
    });
}
int ref_get_ATrhoXiXjXk(void* h, int NT, int ND, int i, int NXi, int j, int NXj, int kk, int NXk, long long npts, const double* T,
                        const double* rho, const double* x, int ncomp, double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        out[k] = am.get_ATrhoXiXjXk(T[k], NT, rho[k], ND, z, i, NXi, j, NXj, kk, NXk);
    });
}

// Isochoric family (teqpcpp.hpp:52-75; deriv_adapter.hpp:226-241).  rhovec is SoA [ncomp][npts].
// what: 0 get_pr, 1 get_splus, 2 get_dpdT_constrhovec                                   -> out[npts]
int ref_isochoric_scalar(void* h, int what, long long npts, const double* T, const double* rhovec, int ncomp, double* out,
                         int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd r; if (r.size() != ncomp) r.resize(ncomp);
        gather(rhovec, npts, k, ncomp, r);
        out[k] = (what == 0) ? am.get_pr(T[k], r) : (what == 1) ? am.get_splus(T[k], r) : am.get_dpdT_constrhovec(T[k], r);
    });
}
// what: 0 build_Psir_gradient_autodiff, 1 get_chempotVLE_autodiff, 2 get_dchempotdT_autodiff, 3 get_fugacity_coefficients,
//       4 get_partial_molar_volumes, 5 build_d2PsirdTdrhoi_autodiff, 6 get_dpdrhovec_constT -> out[ncomp][npts]
int ref_isochoric_array(void* h, int what, long long npts, const double* T, const double* rhovec, int ncomp, double* out,
                        int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    if (what < 0 || what > 6) { put_err("bad selector", err, errlen); return 1; }
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd r; if (r.size() != ncomp) r.resize(ncomp);
        gather(rhovec, npts, k, ncomp, r);
        Eigen::ArrayXd v;
        switch (what) {
            case 0: v = am.build_Psir_gradient_autodiff(T[k], r); break;
            case 1: v = am.get_chempotVLE_autodiff(T[k], r); break;
            case 2: v = am.get_dchempotdT_autodiff(T[k], r); break;
            case 3: v = am.get_fugacity_coefficients(T[k], r); break;
            case 4: v = am.get_partial_molar_volumes(T[k], r); break;
            case 5: v = am.build_d2PsirdTdrhoi_autodiff(T[k], r); break;
            default: v = am.get_dpdrhovec_constT(T[k], r); break;
        }
        for (int i = 0; i < ncomp; ++i) out[(long long)i * npts + k] = v[i];
    });
}
// build_Psir_fgradHessian_autodiff: f[npts], g[ncomp][npts], H[ncomp*ncomp][npts] (row-major index i*ncomp+j). Any may be null.
// psi_total != 0 adds the ideal-gas diagonal, i.e. build_Psi_Hessian_autodiff (derivs.hpp:1076-1084), to H only.
int ref_Psir_fgradHessian(void* h, int psi_total, long long npts, const double* T, const double* rhovec, int ncomp, double* f,
                          double* g, double* H, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd r; if (r.size() != ncomp) r.resize(ncomp);
        gather(rhovec, npts, k, ncomp, r);
        auto [fv, gv, Hv] = am.build_Psir_fgradHessian_autodiff(T[k], r);
        if (f) f[k] = fv;
        if (g) for (int i = 0; i < ncomp; ++i) g[(long long)i * npts + k] = gv[i];
        if (H) {
            if (psi_total) { Eigen::ArrayXXd Ht = am.build_Psi_Hessian_autodiff(T[k], r); for (int i = 0; i < ncomp; ++i) for (int j = 0; j < ncomp; ++j) H[(long long)(i * ncomp + j) * npts + k] = Ht(i, j); }
            else for (int i = 0; i < ncomp; ++i) for (int j = 0; j < ncomp; ++j) H[(long long)(i * ncomp + j) * npts + k] = Hv(i, j);
        }
    });
}

// ln(phi) and its derivatives are template-only in the reference (derivs.hpp:1219-1230, 1306-1393; not in AbstractModel), so
// they need the concrete model type: recovered with get_model_cref<> (deriv_adapter.hpp:279-296) for the kinds listed.
// what: 0 ln phi [ncomp], 1 d ln phi/dT|rhovec [ncomp], 2 d ln phi/drho|T,x [ncomp], 3 d ln phi_i/dx_j|T,rho [ncomp*ncomp], 4 d ln phi/dv|T,x
} // extern "C"
template<typename Model>
void lnphi_point(const Model& model, int what, double T, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) {
    using iso = IsochoricDerivatives<const Model&, double, Eigen::ArrayXd>;
    switch (what) {
        case 0: v = iso::get_ln_fugacity_coefficients(model, T, r); break;
        case 1: v = iso::get_d_ln_fugacity_coefficients_dT_constrhovec(model, T, r); break;
        case 2: v = iso::get_d_ln_fugacity_coefficients_drho_constTmolefracs(model, T, r); break;
        case 4: v = iso::get_d_ln_fugacity_coefficients_dv_constTmolefracs(model, T, r); break;
        default: {
            Eigen::ArrayXXd M = iso::get_d_ln_fugacity_coefficients_dmolefracs_constTrho(model, T, r);
            const int n = (int)r.size(); v.resize(n * n);
            for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) v[i * n + j] = M(i, j);
        }
    }
}
extern "C" {
int ref_lnphi_family(void* h, int what, long long npts, const double* T, const double* rhovec, int ncomp, double* out,
                     int nthreads, char* err, int errlen) {
    REF_TRY
    auto* rm = static_cast<RefModel*>(h);
    const AbstractModel* am = rm->am.get();
    const int nout = (what == 3) ? ncomp * ncomp : ncomp;
    std::function<void(double, const Eigen::ArrayXd&, Eigen::ArrayXd&)> one;
    using namespace teqp::cppinterface::adapter;
    if (rm->kind == "GERG2008resid") { const auto& m = get_model_cref<GERG2008::GERG2008ResidualModel>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "GERG2004resid") { const auto& m = get_model_cref<GERG2004::GERG2004ResidualModel>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "PR" || rm->kind == "SRK" || rm->kind == "cubic") { using M = decltype(make_canonicalPR(nlohmann::json{})); const auto& m = get_model_cref<M>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "PCSAFT") { using M = decltype(PCSAFT::PCSAFTfactory(nlohmann::json{})); const auto& m = get_model_cref<M>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "multifluid") { using M = decltype(multifluidfactory(nlohmann::json{})); const auto& m = get_model_cref<M>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "GERG2008idealgas") { const auto& m = get_model_cref<GERG2008::GERG2008IdealGasModel>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "GERG2004idealgas") { const auto& m = get_model_cref<GERG2004::GERG2004IdealGasModel>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "IdealHelmholtz") { const auto& m = get_model_cref<IdealHelmholtz>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "QCPRAasen") { const auto& m = get_model_cref<QuantumCorrectedPR>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "RKPRCismondi2005") { const auto& m = get_model_cref<RKPRCismondi2005>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else if (rm->kind == "advancedPRaEres") { const auto& m = get_model_cref<advancedPRaEres_t>(am); one = [&m, what](double T_, const Eigen::ArrayXd& r, Eigen::ArrayXd& v) { lnphi_point(m, what, T_, r, v); }; }
    else { put_err("ln(phi) family: kind not wired in the oracle: " + rm->kind, err, errlen); return 200; }
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd r, v; if (r.size() != ncomp) r.resize(ncomp);
        gather(rhovec, npts, k, ncomp, r);
        one(T[k], r, v);
        for (int i = 0; i < nout; ++i) out[(long long)i * npts + k] = v[i];
    });
    REF_CATCH
}

int ref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }


// AbstractModel::get_dmBnvirdTm (teqpcpp.hpp:131; VirialDerivatives::get_dmBnvirdTm_runtime, derivs.hpp:808-841) per temperature;
// Nvir > 4 with NT = 0 goes through get_Bnvir (derivs.hpp:715-752)
int ref_get_dmBnvirdTm(void* h, int Nvir, int NTd, long long npts, const double* T, const double* x, int ncomp, double* out, int nthreads,
                       char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        if (Nvir > 4 && NTd == 0) out[k] = am.get_Bnvir(Nvir, T[k], z)[Nvir];
        else out[k] = am.get_dmBnvirdTm(Nvir, NTd, T[k], z);
    });
}
// AbstractModel::get_Psir_sigma_derivs (teqpcpp.hpp:160; derivs.hpp:1496-1509): out[5][npts]
int ref_Psir_sigma_derivs(void* h, long long npts, const double* T, const double* rhovec, const double* v, int ncomp, double* out, int nthreads,
                          char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd r, d; if (r.size() != ncomp) { r.resize(ncomp); d.resize(ncomp); }
        gather(rhovec, npts, k, ncomp, r); gather(v, npts, k, ncomp, d);
        const Eigen::ArrayXd o = am.get_Psir_sigma_derivs(T[k], r, d);
        for (int i = 0; i < 5; ++i) out[(long long)i * npts + k] = o[i];
    });
}
// AbstractModel::get_deriv_mat2 (teqpcpp.hpp:164; DerivativeHolderSquare<2>, derivs.hpp:1512-1536): the six entries the reference fills,
// out[6][npts] = (0,0) (0,1) (0,2) (1,0) (2,0) (1,1)
int ref_deriv_mat2(void* h, long long npts, const double* T, const double* rho, const double* x, int ncomp, double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        const auto m = am.get_deriv_mat2(T[k], rho[k], z);
        const double v[6] = {m(0, 0), m(0, 1), m(0, 2), m(1, 0), m(2, 0), m(1, 1)};
        for (int i = 0; i < 6; ++i) out[(long long)i * npts + k] = v[i];
    });
}
// AbstractModel::get_reducing_temperature / get_reducing_density (teqpcpp.hpp:124-125), one composition
int ref_get_reducing(void* h, int which, const double* z, int ncomp, double* out, char* err, int errlen) {
    REF_TRY
    Eigen::ArrayXd zz = Eigen::Map<const Eigen::ArrayXd>(z, ncomp);
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    *out = which == 0 ? am.get_reducing_temperature(zz) : am.get_reducing_density(zz);
    return 0;
    REF_CATCH
}
// AbstractModel::get_AtaudeltaXi / XiXj / XiXjXk (teqpcpp.hpp:139-141; derivs.hpp:341-414); idx/N hold (i, NXi[, j, NXj[, k, NXk]])
int ref_get_Ataudelta(void* h, int which, int Ntau, int Ndelta, const int* idx, long long npts, const double* tau, const double* delta, const double* x,
                      int ncomp, double* out, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        thread_local Eigen::ArrayXd z; if (z.size() != ncomp) z.resize(ncomp);
        gather(x, npts, k, ncomp, z);
        if (which == 1) out[k] = am.get_AtaudeltaXi(tau[k], Ntau, delta[k], Ndelta, z, idx[0], idx[1]);
        else if (which == 2) out[k] = am.get_AtaudeltaXiXj(tau[k], Ntau, delta[k], Ndelta, z, idx[0], idx[1], idx[2], idx[3]);
        else out[k] = am.get_AtaudeltaXiXjXk(tau[k], Ntau, delta[k], Ndelta, z, idx[0], idx[1], idx[2], idx[3], idx[4], idx[5]);
    });
}
// teqp::pure_VLE_T and teqp::dpsatdT_pure (algorithms/VLE_pure.hpp:138-171; what AbstractModel::pure_VLE_T / dpsatdT_pure forward to,
// AbstractModel_implementation.cpp) per isotherm; z = one composition for all points; psat from the vapour density
int ref_pure_VLE_T(void* h, long long npts, const double* T, const double* rhoL0, const double* rhoV0, int maxiter, const double* z, int ncomp,
                   double* rhoL, double* rhoV, double* psat, double* dpsatdT, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    const Eigen::ArrayXd zz = Eigen::Map<const Eigen::ArrayXd>(z, ncomp);
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        const auto r = teqp::pure_VLE_T(am, T[k], rhoL0[k], rhoV0[k], maxiter, zz);
        rhoL[k] = r[0]; rhoV[k] = r[1];
        if (psat) psat[k] = r[1] * am.get_R(zz) * T[k] * (1.0 + am.get_Ar01(T[k], r[1], zz));
        if (dpsatdT) dpsatdT[k] = teqp::dpsatdT_pure(am, T[k], r[0], r[1], zz);
    });
}
// teqp::mix_VLE_Tx (algorithms/VLE.hpp:76-192).  ORACLE PORT, not the reference's own function: VLE.hpp cannot be compiled here (it includes
// critical_tracing.hpp -> boost::numeric::odeint and the multicomplex library, both missing blobs), and the reference holds no stored numbers for this
// solver ("parity unpinned" for the loop itself).  The restatement follows VLE.hpp:84-191 step by step on the reference's OWN AbstractModel
// (build_Psir_fgradHessian_autodiff, get_R) and Eigen's colPivHouseholderQr; the tests additionally check the equilibrium conditions at the solution.
int ref_mix_VLE_Tx(void* h, long long npts, const double* T, const double* rvL0, const double* rvV0, const double* xspec, double atol, double reltol,
                   double axtol, double relxtol, int maxiter, double* rvL, double* rvV, double* code, int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        const int N = 2;
        Eigen::MatrixXd J(2 * N, 2 * N), r(2 * N, 1), x(2 * N, 1);
        Eigen::ArrayXd xs(2); xs << xspec[k], xspec[npts + k];
        x(0) = rvL0[k]; x(1) = rvL0[npts + k]; x(2) = rvV0[k]; x(3) = rvV0[npts + k];
        const double RT = am.get_R(xs) * T[k];
        int rc = 0;                                                                                   // VLE_return_code::unset
        for (int iter = 0; iter < maxiter; ++iter) {
            const Eigen::ArrayXd rhovecL = x.col(0).array().head(N), rhovecV = x.col(0).array().tail(N);
            auto [PsirL, gL, HL] = am.build_Psir_fgradHessian_autodiff(T[k], rhovecL);               // VLE.hpp:96-97
            auto [PsirV, gV, HV] = am.build_Psir_fgradHessian_autodiff(T[k], rhovecV);
            const double rhoL = rhovecL.sum(), rhoV = rhovecV.sum();
            const double pL = rhoL * RT - PsirL + (rhovecL * gL).sum(), pV = rhoV * RT - PsirV + (rhovecV * gV).sum();   // :100-101
            const Eigen::ArrayXd dpL = RT + (HL.matrix() * rhovecL.matrix()).array(), dpV = RT + (HV.matrix() * rhovecV.matrix()).array();
            const bool nz0 = rhovecL(0) > 0 && rhovecV(0) > 0, nz1 = rhovecL(1) > 0 && rhovecV(1) > 0;                  // :105-106
            r(0) = nz0 ? gL(0) + RT * log(rhovecL(0)) - (gV(0) + RT * log(rhovecV(0))) : gL(0) - gV(0);
            r(1) = nz1 ? gL(1) + RT * log(rhovecL(1)) - (gV(1) + RT * log(rhovecV(1))) : gL(1) - gV(1);
            r(2) = pL - pV;
            r(3) = rhovecL(0) / rhovecL.sum() - xs(0);
            J.setZero();
            J(0, 0) = HL(0, 0) + (nz0 ? RT / rhovecL(0) : 0); J(0, 1) = HL(0, 1); J(1, 0) = HL(1, 0); J(1, 1) = HL(1, 1) + (nz1 ? RT / rhovecL(1) : 0);
            J(0, 2) = -(HV(0, 0) + (nz0 ? RT / rhovecV(0) : 0)); J(0, 3) = -HV(0, 1); J(1, 2) = -HV(1, 0); J(1, 3) = -(HV(1, 1) + (nz1 ? RT / rhovecV(1) : 0));
            J(2, 0) = dpL(0); J(2, 1) = dpL(1); J(2, 2) = -dpV(0); J(2, 3) = -dpV(1);
            J(3, 0) = (rhoL - rhovecL(0)) / (rhoL * rhoL); J(3, 1) = -rhovecL(0) / (rhoL * rhoL);
            Eigen::ArrayXd dx = J.colPivHouseholderQr().solve(-r);                                     // :141
            if ((!dx.isFinite()).all()) { rc = 5; break; }
            if ((x.array() + dx < 0).any()) { Eigen::ArrayXd dxmax = -x.array(); dx *= (dx / dxmax).minCoeff() / 2; }   // :148-156
            for (int i = 0; i < 2; ++i) if (xs[i] == 0) { dx[i] = 0; dx[i + 2] = 0; }
            x.array() += dx;
            if ((dx.abs() < (axtol + relxtol * x.array().abs())).all()) { rc = 1; break; }
            if ((r.array().abs() < (atol + reltol * r.array().abs())).all()) { rc = 2; break; }
            if (((x.array() - dx).abs() < std::numeric_limits<double>::min()).all()) { rc = 1; break; }
            if (iter == maxiter - 1) rc = 4;
        }
        rvL[k] = x(0); rvL[npts + k] = x(1); rvV[k] = x(2); rvV[npts + k] = x(3);
        if (code) code[k] = rc;
    });
}
// teqp::mix_VLE_Tp, Newton branch (algorithms/VLE.hpp:196-272 residuals / Jacobian, :351-367 loop).  ORACLE PORT for the same reason as ref_mix_VLE_Tx.
int ref_mix_VLE_Tp(void* h, long long npts, const double* T, const double* pg, const double* rvL0, const double* rvV0, int maxiter, double* rvL, double* rvV,
                   int nthreads, char* err, int errlen) {
    const AbstractModel& am = *static_cast<RefModel*>(h)->am;
    return parallel_points(npts, nthreads, err, errlen, [&](long long k) {
        const int N = 2;
        Eigen::MatrixXd J(2 * N, 2 * N);
        Eigen::VectorXd r(2 * N), x(2 * N);
        x(0) = rvL0[k]; x(1) = rvL0[npts + k]; x(2) = rvV0[k]; x(3) = rvV0[npts + k];
        const double p = pg[k];
        for (int iter = 0; iter < maxiter; ++iter) {
            const Eigen::ArrayXd rhovecL = x.array().head(N), rhovecV = x.array().tail(N);
            const double RT = am.get_R((rhovecL / rhovecL.sum()).eval()) * T[k];                       // :209
            auto [PsirL, gL, HL] = am.build_Psir_fgradHessian_autodiff(T[k], rhovecL);
            auto [PsirV, gV, HV] = am.build_Psir_fgradHessian_autodiff(T[k], rhovecV);
            const double rhoL = rhovecL.sum(), rhoV = rhovecV.sum();
            const double pL = rhoL * RT - PsirL + (rhovecL * gL).sum(), pV = rhoV * RT - PsirV + (rhovecV * gV).sum();
            const Eigen::ArrayXd dpL = RT + (HL.matrix() * rhovecL.matrix()).array(), dpV = RT + (HV.matrix() * rhovecV.matrix()).array();
            const bool nz0 = rhovecL(0) > 0 && rhovecV(0) > 0, nz1 = rhovecL(1) > 0 && rhovecV(1) > 0;
            r(0) = nz0 ? gL(0) + RT * log(rhovecL(0)) - (gV(0) + RT * log(rhovecV(0))) : gL(0) - gV(0);
            r(1) = nz1 ? gL(1) + RT * log(rhovecL(1)) - (gV(1) + RT * log(rhovecV(1))) : gL(1) - gV(1);
            r(2) = (pV - p) / p; r(3) = (pL - p) / p;                                                  // :231-232
            J.setZero();
            J(0, 0) = HL(0, 0) + (nz0 ? RT / rhovecL(0) : 0); J(0, 1) = HL(0, 1); J(1, 0) = HL(1, 0); J(1, 1) = HL(1, 1) + (nz1 ? RT / rhovecL(1) : 0);
            J(0, 2) = -(HV(0, 0) + (nz0 ? RT / rhovecV(0) : 0)); J(0, 3) = -HV(0, 1); J(1, 2) = -HV(1, 0); J(1, 3) = -(HV(1, 1) + (nz1 ? RT / rhovecV(1) : 0));
            J(2, 2) = dpV(0) / p; J(2, 3) = dpV(1) / p; J(3, 0) = dpL(0) / p; J(3, 1) = dpL(1) / p;     // :261-270
            Eigen::ArrayXd dx = J.colPivHouseholderQr().solve(-r);
            if ((x.array() + dx < 0).any()) { Eigen::ArrayXd dxmax = -x.array(); dx *= (dx / dxmax).minCoeff() / 2; }
            x.array() += dx;
        }
        rvL[k] = x(0); rvL[npts + k] = x(1); rvV[k] = x(2); rvV[npts + k] = x(3);
    });
}
} // extern "C"
