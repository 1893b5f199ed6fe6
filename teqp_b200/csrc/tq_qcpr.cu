// "QCPRAasen": JSON -> parameter block (QuantumCorrectedPR::QuantumCorrectedPR, /root/reference/include/teqp/models/cubics/simple_cubics.hpp:447-475)
// and the generic jet-sweep launches.
#include <cmath>
#include <cstring>

#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_qcpr.cuh"

namespace tq {
void qcpr_launch_isoprops(const QcprParams&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);                   // tq_ip_qcpr.cu
void qcpr_launch_ATrhoXiXj(const QcprParams&, const DeviceBatch&, int, int, double*, cudaStream_t);                   // tq_xx_qcpr.cu
void qcpr_launch_ATrhoXiXjXk(const QcprParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void qcpr_launch_sigma(const QcprParams&, const DeviceBatch&, const double*, double*, cudaStream_t);

namespace {
class QcprModel final : public Model {
public:
    QcprParams p{};
    int ncomp() const override { return p.nc; }
    const char* family() const override { return "QCPR"; }
    double R(const double*, int) const override { return p.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ar_table<QcprDev>(p, p.nc, NT, ND, in, out, st); }
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ATrhoXi<QcprDev>(p, p.nc, in, out, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { generic_launch_isochoric<QcprDev>(p, p.nc, in, o, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override { qcpr_launch_isoprops(p, in, o, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { qcpr_launch_ATrhoXiXj(p, in, i, j, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { qcpr_launch_ATrhoXiXjXk(p, in, i, j, k, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { qcpr_launch_sigma(p, in, v, out, st); }
    // no virial table: the volume translation forms 1/rho, which has no rho -> 0 jet
};

std::vector<double> vec(const Json& j, const char* key, size_t n) {
    std::vector<double> v = j.at(key).as_doubles();
    if (n && v.size() != n) throw InvalidArgument(std::string("QCPRAasen: array \"") + key + "\" must have one entry per component");
    return v;
}
void square(const Json& j, size_t N, const char* what, double* one_minus) {
    // build_square_matrix, /root/reference/include/teqp/json_tools.hpp:45-71, turns null or [] into a 0 x 0 matrix, which the reference's
    // QuantumCorrectedPR then indexes out of bounds (simple_cubics.hpp:472, 507-516 have no size check): an N x N matrix is required here
    if (!j.is_array() || j.size() != N) throw InvalidArgument(std::string(what) + " needs to be a square matrix the same size as the number of components [" + std::to_string(N) + "]");
    for (size_t i = 0; i < N; ++i) {
        if (!j[i].is_array() || j[i].size() != N) throw InvalidArgument("provided matrix is not square");
        for (size_t k = 0; k < N; ++k) one_minus[i * N + k] = 1.0 - j[i][k].as_double();
    }
}
}  // namespace

std::unique_ptr<Model> make_qcpr_model(const Json& spec) {
    auto m = std::make_unique<QcprModel>();
    QcprParams& p = m->p;
    std::memset(&p, 0, sizeof(p));
    const std::vector<double> Tc = vec(spec, "Tcrit / K", 0);
    const size_t N = Tc.size();
    if (N == 0 || N > (size_t)kQcprMaxC) throw InvalidArgument("QCPRAasen supports 1.." + std::to_string(kQcprMaxC) + " components on the device path");
    const std::vector<double> pc = vec(spec, "pcrit / Pa", N), L = vec(spec, "Ls", 0), M = vec(spec, "Ms", 0), Nn = vec(spec, "Ns", 0), As = vec(spec, "As", N),
                              Bs = vec(spec, "Bs", N), cs = vec(spec, "cs / m^3/mol", N);
    if (L.size() != M.size() || M.size() != Nn.size()) throw InvalidArgument("L,M,N must all be the same length");
    if (L.size() != N) throw InvalidArgument("QCPRAasen: Ls, Ms, Ns must have one entry per component");
    p.nc = (int)N;
    p.R = spec.value("R / J/mol/K", 8.31446261815324);   // constants::R_CODATA2017
    for (size_t i = 0; i < N; ++i) {
        p.Tc[i] = Tc[i];
        p.a0[i] = 0.45723552892138218938 * ((Tc[i] * p.R) * (Tc[i] * p.R)) / pc[i];   // OmegaA * POW2(Tc Ru) / pc (:499-500)
        p.b0[i] = 0.07780 * Tc[i] * p.R / pc[i];                                        // :489
        const double q = 1.0 + As[i] / (Tc[i] + Bs[i]);
        p.bden[i] = q * q * q;
        p.L[i] = L[i]; p.M[i] = M[i]; p.N[i] = Nn[i]; p.A[i] = As[i]; p.B[i] = Bs[i]; p.c[i] = cs[i];
    }
    square(spec.at("kmat"), N, "kmat", p.one_minus_k);
    square(spec.at("lmat"), N, "lmat", p.one_minus_l);
    return m;
}

}  // namespace tq
