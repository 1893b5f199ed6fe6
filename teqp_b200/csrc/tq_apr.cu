// "advancedPRaEres": JSON -> parameter block (make_AdvancedPRaEres, /root/reference/include/teqp/models/cubics/advancedmixing_cubics.hpp:220-250;
// options :32-47; alpha functions build_alpha_functions, simple_cubics.hpp:97-133) and the generic jet-sweep launches.
#include <cmath>
#include <cstring>

#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_apr.cuh"

namespace tq {
void apr_launch_isoprops(const AprParams&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);                   // tq_ip_apr.cu
void apr_launch_ATrhoXiXj(const AprParams&, const DeviceBatch&, int, int, double*, cudaStream_t);                   // tq_xx_apr.cu
void apr_launch_ATrhoXiXjXk(const AprParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void apr_launch_sigma(const AprParams&, const DeviceBatch&, const double*, double*, cudaStream_t);

namespace {
class AprModel final : public Model {
public:
    AprParams p{};
    int ncomp() const override { return p.nc; }
    const char* family() const override { return "advancedPRaEres"; }
    double R(const double*, int) const override { return p.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ar_table<AprDev>(p, p.nc, NT, ND, in, out, st); }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { launch_ar_table(NT, ND, in, out, st); }   // in.rho_unit_seed is set by the caller
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ATrhoXi<AprDev>(p, p.nc, in, out, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { generic_launch_isochoric<AprDev>(p, p.nc, in, o, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override { apr_launch_isoprops(p, in, o, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { apr_launch_ATrhoXiXj(p, in, i, j, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { apr_launch_ATrhoXiXjXk(p, in, i, j, k, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { apr_launch_sigma(p, in, v, out, st); }
};

double pr78_m(double w) {   // simple_cubics.hpp:118-124
    return (w < 0.491) ? 0.37464 + 1.54226 * w - 0.26992 * (w * w) : 0.379642 + 1.48503 * w - 0.164423 * (w * w) + 0.016666 * (w * w * w);
}
void wilson_matrix(const Json& j, size_t N, const char* what, double* out) {
    if (!j.is_array() || j.size() != N) throw InvalidArgument(std::string("Wilson matrix \"") + what + "\" needs to be square, the same size as the number of components [" + std::to_string(N) + "]");
    for (size_t i = 0; i < N; ++i) {
        if (!j[i].is_array() || j[i].size() != N) throw InvalidArgument("provided matrix is not square");
        for (size_t k = 0; k < N; ++k) out[i * N + k] = j[i][k].as_double();
    }
}
}  // namespace

std::unique_ptr<Model> make_apr_model(const Json& spec) {
    auto mdl = std::make_unique<AprModel>();
    AprParams& p = mdl->p;
    std::memset(&p, 0, sizeof(p));
    const std::vector<double> Tc = spec.at("Tcrit / K").as_doubles(), pc = spec.at("pcrit / Pa").as_doubles();
    const size_t N = Tc.size();
    if (pc.size() != N) throw InvalidArgument("advancedPRaEres: Tcrit / K and pcrit / Pa must have one entry per component");
    const Json& ja = spec.at("alphas");
    if (ja.size() != N) throw InvalidArgument("alpha [size " + std::to_string(ja.size()) + "] must be the same size as components [size " + std::to_string(N) + "]");
    // the reference's factory hands a 2 x 2 zero lmat to the model, whose check_lmat (:91-102) then refuses every other component count
    if (N != 2) throw InvalidArgument("lmat needs to be a square matrix the same size as the number of components [" + std::to_string(N) + "]");
    p.nc = (int)N;
    for (size_t i = 0; i < N; ++i) {
        const std::string at = ja[i].at("type").as_string();
        if (at == "Twu" || at == "Mathias-Copeman") {
            const std::vector<double> c = ja[i].at("c").as_doubles();
            if (c.size() != 3) throw InvalidArgument("coefficients c for " + at + " alpha function must have length 3");
            p.alpha_kind[i] = (at == "Twu") ? 1 : 2;
            p.c0[i] = c[0]; p.c1[i] = c[1]; p.c2[i] = c[2];
        } else if (at == "PR78") {
            p.alpha_kind[i] = 0;
            p.c0[i] = pr78_m(ja[i].at("acentric").as_double());
        } else {
            throw InvalidArgument("alpha type is not understood: " + at);
        }
    }
    const Json& jo = spec.at("options");
    const Json& jb = jo.at("brule");
    if (jb.is_null()) throw InvalidArgument("Mixing rule for b is invalid");
    const std::string brule = jb.as_string();
    if (brule != "Linear" && brule != "Quadratic") throw InvalidArgument("Mixing rule for b is invalid");
    p.linear_b = (brule == "Linear");
    const double s = jo.at("s").as_double();
    p.CEoS = jo.at("CEoS").as_double();
    p.R = jo.value("R / J/mol/K", 8.31446261815324);
    const double OmegaA = 0.45723552892138218938, OmegaB = 0.077796073903888455972;
    for (size_t i = 0; i < N; ++i) {
        p.Tc[i] = Tc[i];
        p.ai[i] = OmegaA * ((p.R * Tc[i]) * (p.R * Tc[i])) / pc[i];
        p.bi[i] = OmegaB * p.R * Tc[i] / pc[i];
        p.v[i] = OmegaB * 8.31446261815324 * Tc[i] / pc[i];   // the static get_bi (:113-117): CODATA R whatever "R / J/mol/K" says
    }
    for (size_t i = 0; i < N; ++i)
        for (size_t j = 0; j < N; ++j) {
            p.bij[i * N + j] = (1 - 0.0) * std::pow((std::pow(p.bi[i], 1.0 / s) + std::pow(p.bi[j], 1.0 / s)) / 2.0, s);
            p.vratio[i * N + j] = p.v[j] / p.v[i];
        }
    const Json& jw = spec.at("aresmodel");
    const std::string wt = jw.at("type").as_string();
    if (wt != "Wilson") throw InvalidArgument("bad type of ares model: " + wt);
    wilson_matrix(jw.at("m"), N, "m", p.m);
    wilson_matrix(jw.at("n"), N, "n", p.n);
    return mdl;
}

}  // namespace tq
