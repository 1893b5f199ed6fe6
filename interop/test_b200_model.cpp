// C++ test of the drop-in boundary (needs a GPU): a B200Model registered the way a teqp plug-in is
// (add_model_pointer_factory_function, /root/reference/include/teqp/cpp/teqpcpp.hpp:214-221; interface/CPP/teqp_impl_factory.cpp:104-147)
// is driven through `const teqp::cppinterface::AbstractModel&` side by side with the reference's own adapter for the same model
// specification (oracle/_ref: DerivativeAdapter over the reference's templates), and every virtual the hot path owns is compared.
// Tolerance: the parity rule of tests/parity.py (1e-12 relative + 2e-14 of the family scale), loosened only where stated.
#include <cmath>
#include <cstdio>
#include <functional>
#include <map>
#include <string>

#include "b200_model.hpp"
#include "teqp/exceptions.hpp"
#include "teqp/algorithms/VLE_pure.hpp"   // the reference's own teqp::pure_VLE_T / dpsatdT_pure (what AbstractModel::pure_VLE_T forwards to)

std::unique_ptr<teqp::cppinterface::AbstractModel> teqp_ref_oracle_make_model(const nlohmann::json& j);   // oracle/ref_capi.cpp

using teqp::cppinterface::AbstractModel;
using teqp::cppinterface::ModelPointerFactoryFunction;

namespace {
// the registry of teqp_impl_factory.cpp:104-147, restated: kind -> factory; make_model_local (teqp: make_model) looks the kind up and passes spec["model"]
std::map<std::string, ModelPointerFactoryFunction> registry;
void register_factory(const std::string& key, ModelPointerFactoryFunction& func) {
    if (registry.count(key)) throw teqp::InvalidArgument("key is already present, overwriting is not currently allowed");
    registry[key] = func;
}
std::unique_ptr<AbstractModel> make_model_local(const nlohmann::json& j) { return registry.at(j.at("kind"))(j.at("model")); }

int failures = 0, checks = 0;
void close(const char* what, double got, double ref, double scale, double rtol = 1e-12, double ftol = 2e-14) {
    ++checks;
    const double tol = rtol * std::fabs(ref) + ftol * scale;
    if (!(std::fabs(got - ref) <= tol) && !(std::isnan(got) && std::isnan(ref))) {
        ++failures;
        std::printf("FAIL %-44s got %.17g ref %.17g |d| %.3e tol %.3e\n", what, got, ref, std::fabs(got - ref), tol);
    }
}
template<class A, class B> void close_arr(const char* what, const A& got, const B& ref, double rtol = 1e-12, double ftol = 2e-14) {
    if (got.size() != ref.size()) { ++failures; std::printf("FAIL %s: size %ld vs %ld\n", what, (long)got.size(), (long)ref.size()); return; }
    double scale = 0;
    for (Eigen::Index i = 0; i < ref.size(); ++i) scale = std::max(scale, std::fabs(ref.data()[i]));
    for (Eigen::Index i = 0; i < ref.size(); ++i) close(what, got.data()[i], ref.data()[i], scale, rtol, ftol);
}

void compare(const char* label, const nlohmann::json& spec, double T, double rho, const Eigen::ArrayXd& z, bool multifluid_like, bool has_isoprops) {
    std::printf("== %s\n", label);
    const auto gpu = make_model_local({{"kind", "B200"}, {"model", spec}});
    const auto ref = teqp_ref_oracle_make_model(spec);
    const AbstractModel& a = *gpu;
    const AbstractModel& b = *ref;
    close("get_R", a.get_R(z), b.get_R(z), 0);
    const auto r06 = b.get_Ar06n(T, rho, z);
    double fam = 0, f = 1;
    for (int k = 0; k <= 6; ++k) { if (k > 1) f *= k; fam = std::max(fam, std::fabs(r06[k]) / f); }
    close_arr("get_Ar04n", a.get_Ar04n(T, rho, z), b.get_Ar04n(T, rho, z));
    close_arr("get_Ar06n", a.get_Ar06n(T, rho, z), r06, 1e-12, 2e-14 * 720);
    close_arr("get_Ar20n", a.get_Ar20n(T, rho, z), b.get_Ar20n(T, rho, z));
    close("get_Ar00", a.get_Ar00(T, rho, z), b.get_Ar00(T, rho, z), fam);
    close("get_Ar01", a.get_Ar01(T, rho, z), b.get_Ar01(T, rho, z), fam);
    close("get_Ar11", a.get_Ar11(T, rho, z), b.get_Ar11(T, rho, z), fam);
    close("get_Ar12", a.get_Ar12(T, rho, z), b.get_Ar12(T, rho, z), 2 * fam);
    close("get_Ar21", a.get_Ar21(T, rho, z), b.get_Ar21(T, rho, z), 2 * fam);
    close("get_Arxy(1,2)", a.get_Arxy(1, 2, T, rho, z), b.get_Arxy(1, 2, T, rho, z), 2 * fam);
    close("get_Arxy(0,3)", a.get_Arxy(0, 3, T, rho, z), b.get_Arxy(0, 3, T, rho, z), 6 * fam);
    {
        const auto m2 = a.get_deriv_mat2(T, rho, z), r2 = b.get_deriv_mat2(T, rho, z);
        const int idx[6][2] = {{0, 0}, {0, 1}, {0, 2}, {1, 0}, {2, 0}, {1, 1}};   // the entries DerivativeHolderSquare<2> fills (derivs.hpp:1524-1534)
        for (auto& ij : idx) close("get_deriv_mat2", m2(ij[0], ij[1]), r2(ij[0], ij[1]), 2 * fam);
    }
    if (z.size() > 1) {
        close("get_ATrhoXi(1,1,x1^2)", a.get_ATrhoXi(T, 1, rho, 1, z, 1, 2), b.get_ATrhoXi(T, 1, rho, 1, z, 1, 2), fam, 1e-11, 1e-13);
        close("get_ATrhoXi(0,0,x0^1)", a.get_ATrhoXi(T, 0, rho, 0, z, 0, 1), b.get_ATrhoXi(T, 0, rho, 0, z, 0, 1), fam, 1e-11, 1e-13);
    }
    for (int n = 2; n <= 4; ++n)
        for (int m = 0; m <= 2; ++m) {
            const double r = b.get_dmBnvirdTm(n, m, T, z);
            close("get_dmBnvirdTm", a.get_dmBnvirdTm(n, m, T, z), r, std::fabs(b.get_dmBnvirdTm(n, 0, T, z)) / std::pow(T, m), 1e-11, 1e-12);
        }
    close("get_B2vir", a.get_B2vir(T, z), b.get_B2vir(T, z), 0, 1e-11);
    if (z.size() == 2) close("get_B12vir", a.get_B12vir(T, z), b.get_B12vir(T, z), std::fabs(b.get_B2vir(T, z)), 1e-10, 1e-12);
    if (multifluid_like) {
        close("get_reducing_temperature", a.get_reducing_temperature(z), b.get_reducing_temperature(z), 0);
        close("get_reducing_density", a.get_reducing_density(z), b.get_reducing_density(z), 0);
    }
    // isochoric family
    const Eigen::ArrayXd rv = z * rho;
    const double RT = b.get_R(z) * T;
    close_arr("build_Psir_gradient_autodiff", a.build_Psir_gradient_autodiff(T, rv), b.build_Psir_gradient_autodiff(T, rv));
    close_arr("get_fugacity_coefficients", a.get_fugacity_coefficients(T, rv), b.get_fugacity_coefficients(T, rv), 5e-12);
    close_arr("build_Psir_Hessian_autodiff", a.build_Psir_Hessian_autodiff(T, rv), b.build_Psir_Hessian_autodiff(T, rv), 1e-12, 1e-13);
    close_arr("build_Psi_Hessian_autodiff", a.build_Psi_Hessian_autodiff(T, rv), b.build_Psi_Hessian_autodiff(T, rv), 1e-12, 1e-13);
    {
        const auto [f1, g1, H1] = a.build_Psir_fgradHessian_autodiff(T, rv);
        const auto [f2, g2, H2] = b.build_Psir_fgradHessian_autodiff(T, rv);
        close("fgradHessian f", f1, f2, RT * rho * fam);
        close_arr("fgradHessian g", g1, g2);
        close_arr("fgradHessian H", H1, H2, 1e-12, 1e-13);
    }
    {
        Eigen::ArrayXd v = rv;
        for (Eigen::Index i = 0; i < v.size(); ++i) v[i] *= 1.0 + 0.5 * std::cos((double)i);
        const auto s1 = a.get_Psir_sigma_derivs(T, rv, v), s2 = b.get_Psir_sigma_derivs(T, rv, v);
        double sc = 0, ff = 1;
        for (int k = 0; k < 5; ++k) { if (k > 1) ff *= k; sc = std::max(sc, std::fabs(s2[k]) / ff); }
        ff = 1;
        for (int k = 0; k < 5; ++k) { if (k > 1) ff *= k; close("get_Psir_sigma_derivs", s1[k], s2[k], sc * ff, 1e-12, 1e-13); }
    }
    if (has_isoprops) {
        close("get_pr", a.get_pr(T, rv), b.get_pr(T, rv), RT * rho * fam);
        close("get_splus", a.get_splus(T, rv), b.get_splus(T, rv), fam);
        close("get_dpdT_constrhovec", a.get_dpdT_constrhovec(T, rv), b.get_dpdT_constrhovec(T, rv), b.get_R(z) * rho);
        close_arr("get_partial_molar_volumes", a.get_partial_molar_volumes(T, rv), b.get_partial_molar_volumes(T, rv), 1e-11);
        close_arr("get_chempotVLE_autodiff", a.get_chempotVLE_autodiff(T, rv), b.get_chempotVLE_autodiff(T, rv), 1e-12, 1e-13);
        close_arr("get_dchempotdT_autodiff", a.get_dchempotdT_autodiff(T, rv), b.get_dchempotdT_autodiff(T, rv), 1e-12, 1e-13);
        close_arr("build_d2PsirdTdrhoi_autodiff", a.build_d2PsirdTdrhoi_autodiff(T, rv), b.build_d2PsirdTdrhoi_autodiff(T, rv), 1e-12, 1e-13);
        close_arr("get_dpdrhovec_constT", a.get_dpdrhovec_constT(T, rv), b.get_dpdrhovec_constT(T, rv), 1e-12, 1e-13);
    }
    // error behaviour: wrong composition length throws a teqp exception on both sides
    bool threw = false;
    try { Eigen::ArrayXd bad(z.size() + 1); bad.setConstant(0.1); a.get_Ar01(T, rho, bad); } catch (const std::exception&) { threw = true; }
    ++checks; if (!threw) { ++failures; std::printf("FAIL wrong-size molefrac did not throw\n"); }
}
}  // namespace

int main() {
    ModelPointerFactoryFunction f = teqp_b200::factory;
    register_factory("B200", f);   // what add_model_pointer_factory_function does in a teqp build
    try {
        Eigen::ArrayXd z1(1); z1 << 1.0;
        Eigen::ArrayXd z2(2); z2 << 0.4, 0.6;
        Eigen::ArrayXd z3(3); z3 << 0.2, 0.7, 0.1;
        compare("PR methane+ethane", nlohmann::json::parse(R"({"kind":"PR","model":{"Tcrit / K":[190.564,305.322],"pcrit / Pa":[4599200,4872200],"acentric":[0.011,0.099]}})"),
                300.0, 3000.0, z2, false, true);
        compare("PC-SAFT methane+ethane", nlohmann::json::parse(R"({"kind":"PCSAFT","model":{"names":["Methane","Ethane"]}})"), 250.0, 4000.0, z2, false, true);
        compare("GERG-2008 nitrogen+methane+ethane", nlohmann::json::parse(R"({"kind":"GERG2008resid","model":{"names":["nitrogen","methane","ethane"]}})"),
                280.0, 6000.0, z3, true, true);
        compare("SAFT-VR-Mie methane", nlohmann::json::parse(R"({"kind":"SAFT-VR-Mie","model":{"names":["Methane"]}})"), 200.0, 5000.0, z1, false, false);
        // batched extension == the scalar virtual
        {
            const auto gpu = make_model_local({{"kind", "B200"}, {"model", nlohmann::json::parse(R"({"kind":"GERG2008resid","model":{"names":["nitrogen","methane","ethane"]}})")}});
            const auto& m = dynamic_cast<const teqp_b200::B200Model&>(*gpu);
            const int n = 1000;
            Eigen::ArrayXd T = Eigen::ArrayXd::LinSpaced(n, 250, 450), rho = Eigen::ArrayXd::LinSpaced(n, 10, 20000);
            Eigen::ArrayXXd x(n, 3);
            for (int k = 0; k < n; ++k) { x(k, 0) = 0.1 + 0.0002 * k; x(k, 2) = 0.1; x(k, 1) = 1 - x(k, 0) - x(k, 2); }
            const auto out = m.get_Ar0n_batch(4, T, rho, x);
            for (int k : {0, 17, 999}) {
                const Eigen::ArrayXd zk = x.row(k).transpose();
                const auto one = gpu->get_Ar04n(T[k], rho[k], zk);
                for (int j = 0; j < 5; ++j) { ++checks; if (out(k, j) != one[j]) { ++failures; std::printf("FAIL batch != scalar at %d,%d\n", k, j); } }
            }
        }
        // batched saturation solver == the reference's pure_VLE_T driven by the reference's own adapter, isotherm by isotherm
        {
            const auto spec = nlohmann::json::parse(R"({"kind":"PR","model":{"Tcrit / K":[190.564],"pcrit / Pa":[4599200],"acentric":[0.011]}})");
            const auto gpu = make_model_local({{"kind", "B200"}, {"model", spec}});
            const auto ref = teqp_ref_oracle_make_model(spec);
            const auto& m = dynamic_cast<const teqp_b200::B200Model&>(*gpu);
            Eigen::ArrayXd T(4), rL0(4), rV0(4);
            T << 120.0, 140.0, 160.0, 180.0;
            rL0 << 25500.0, 23500.0, 21000.0, 17000.0;     // rough Peng-Robinson saturation densities of methane
            rV0 << 170.0, 570.0, 1500.0, 3800.0;
            const auto out = m.pure_VLE_T_batch(T, rL0, rV0, 10);
            for (int k = 0; k < 4; ++k) {
                const auto r = teqp::pure_VLE_T(*ref, T[k], rL0[k], rV0[k], 10);
                const double dp = teqp::dpsatdT_pure(*ref, T[k], r[0], r[1]);
                ++checks; if (!(std::abs(out(k, 0) / r[0] - 1) < 1e-10 && std::abs(out(k, 1) / r[1] - 1) < 1e-10 && r[0] > 1.2 * r[1])) { ++failures; std::printf("FAIL pure_VLE_T_batch at %g K: %.15g %.15g vs %.15g %.15g\n", T[k], out(k, 0), out(k, 1), r[0], r[1]); }
                ++checks; if (!(std::abs(out(k, 3) / dp - 1) < 1e-9)) { ++failures; std::printf("FAIL dpsatdT at %g K: %.15g vs %.15g\n", T[k], out(k, 3), dp); }
            }
        }
    } catch (const std::exception& e) { std::printf("EXCEPTION %s\n", e.what()); return 2; }
    std::printf("%d checks, %d failures\n", checks, failures);
    return failures ? 1 : 0;
}
