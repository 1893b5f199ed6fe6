// "RKPRCismondi2005": JSON -> parameter block (RKPRCismondi2005::RKPRCismondi2005, build_ac, build_bc,
// /root/reference/include/teqp/models/cubics/simple_cubics.hpp:603-626) and the generic jet-sweep launches.
#include <cmath>
#include <cstring>

#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_rkpr.cuh"

namespace tq {
void rkpr_launch_isoprops(const RkprParams&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);                   // tq_ip_rkpr.cu
void rkpr_launch_ATrhoXiXj(const RkprParams&, const DeviceBatch&, int, int, double*, cudaStream_t);                   // tq_xx_rkpr.cu
void rkpr_launch_ATrhoXiXjXk(const RkprParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void rkpr_launch_sigma(const RkprParams&, const DeviceBatch&, const double*, double*, cudaStream_t);

namespace {
class RkprModel final : public Model {
public:
    RkprParams p{};
    int ncomp() const override { return p.nc; }
    const char* family() const override { return "RKPR"; }
    double R(const double*, int) const override { return p.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ar_table<RkprDev>(p, p.nc, NT, ND, in, out, st); }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { launch_ar_table(NT, ND, in, out, st); }   // in.rho_unit_seed is set by the caller
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ATrhoXi<RkprDev>(p, p.nc, in, out, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { generic_launch_isochoric<RkprDev>(p, p.nc, in, o, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override { rkpr_launch_isoprops(p, in, o, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { rkpr_launch_ATrhoXiXj(p, in, i, j, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { rkpr_launch_ATrhoXiXjXk(p, in, i, j, k, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { rkpr_launch_sigma(p, in, v, out, st); }
};

void square(const Json& j, size_t N, const char* what, double* one_minus) {
    // as for QCPRAasen (tq_qcpr.cu): the reference indexes a 0 x 0 matrix out of bounds when kmat / lmat is null or [], so an N x N matrix is required
    if (!j.is_array() || j.size() != N) throw InvalidArgument(std::string(what) + " needs to be a square matrix the same size as the number of components [" + std::to_string(N) + "]");
    for (size_t i = 0; i < N; ++i) {
        if (!j[i].is_array() || j[i].size() != N) throw InvalidArgument("provided matrix is not square");
        for (size_t k = 0; k < N; ++k) one_minus[i * N + k] = 1.0 - j[i][k].as_double();
    }
}
}  // namespace

std::unique_ptr<Model> make_rkpr_model(const Json& spec) {
    auto m = std::make_unique<RkprModel>();
    RkprParams& p = m->p;
    std::memset(&p, 0, sizeof(p));
    const std::vector<double> d1 = spec.at("delta_1").as_doubles(), Tc = spec.at("Tcrit / K").as_doubles(), pc = spec.at("pcrit / Pa").as_doubles(), k = spec.at("k").as_doubles();
    const size_t N = d1.size();
    if (N == 0 || N > (size_t)kRkprMaxC) throw InvalidArgument("RKPRCismondi2005 supports 1.." + std::to_string(kRkprMaxC) + " components on the device path");
    if (Tc.size() != N || pc.size() != N || k.size() != N) throw InvalidArgument("RKPRCismondi2005: delta_1, Tcrit / K, pcrit / Pa and k must have one entry per component");
    p.nc = (int)N;
    p.R = spec.value("R / J/mol/K", 8.31446261815324);   // constants::R_CODATA2017
    for (size_t i = 0; i < N; ++i) {
        // get_yd1, build_ac, build_bc (:605-626), same operation order
        const double y = 1 + std::cbrt(2 * (1 + d1[i])) + std::cbrt(4 / (1 + d1[i])), dd = (1 + d1[i] * d1[i]) / (1 + d1[i]);
        const double Omega_a = (3 * y * y + 3 * y * dd + dd * dd + dd - 1.0) / std::pow(3.0 * y + dd - 1.0, 2);
        const double Omega_b = 1.0 / (3.0 * y + dd - 1.0);
        p.Tc[i] = Tc[i];
        p.ac[i] = Omega_a * std::pow(p.R * Tc[i], 2) / pc[i];
        p.bc[i] = Omega_b * p.R * Tc[i] / pc[i];
        p.k[i] = k[i];
        p.delta1[i] = d1[i];
    }
    square(spec.at("kmat"), N, "kmat", p.one_minus_k);
    square(spec.at("lmat"), N, "lmat", p.one_minus_l);
    return m;
}

}  // namespace tq
