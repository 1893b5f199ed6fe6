// See b200_model.hpp.  Every method is a thin call into the C ABI of libteqpb200 (include/teqpb200.h); errors come back as return codes and
// are re-thrown as the reference's exception types (teqp/exceptions.hpp), so callers see what they see from the reference's adapter.
#include "b200_model.hpp"

#include <typeindex>

#include "teqp/exceptions.hpp"
#include "../include/teqpb200.h"

namespace teqp_b200 {

namespace {
constexpr int kErr = 2000;
[[noreturn]] void raise(int code, const char* msg) {
    if (code == 200) throw teqp::NotImplementedError(msg);
    if (code == 1) throw teqp::InvalidArgument(msg);
    if (code == 100) throw teqp::IterationFailure(msg);
    throw teqp::InvalidArgument(std::string(msg));   // the reference's C layer flattens everything else to "an exception" as well
}
#define TQ_CALL(expr) do { char err_[kErr] = ""; const int rc_ = (expr); if (rc_ != 0) raise(rc_, err_); } while (0)
#define ERRARGS err_, kErr
}  // namespace

B200Model::B200Model(const nlohmann::json& spec) {
    const std::string s = spec.dump();
    char err_[kErr] = "";
    if (build_model(s.c_str(), &uuid_, ERRARGS) != 0) throw teqp::InvalidArgument(std::string("libteqpb200: ") + err_);
    TQ_CALL(teqpb200_get_ncomp(uuid_, &ncomp_, ERRARGS));
}
B200Model::~B200Model() { char e[64]; if (uuid_ >= 0) free_model(uuid_, e, 64); }

const std::type_index& B200Model::get_type_index() const { static const std::type_index ti(typeid(B200Model)); return ti; }

double B200Model::get_R(const EArrayd& z) const { double R = 0; TQ_CALL(teqpb200_get_R(uuid_, z.data(), (int)z.size(), &R, ERRARGS)); return R; }

double B200Model::get_Arxy(const int NT, const int ND, const double T, const double rho, const EArrayd& z) const {
    double v = 0;
    TQ_CALL(::get_Arxy(uuid_, NT, ND, T, rho, z.data(), (int)z.size(), &v, ERRARGS));
    return v;
}

// one sweep gives the whole (NT+1) x (ND+1) table, row-major
Eigen::ArrayXd B200Model::ar_table(int NT, int ND, double T, double rho, const REArrayd& z) const {
    const EArrayd zc = z;   // contiguous copy of the Ref
    Eigen::ArrayXd out((NT + 1) * (ND + 1));
    if (NT == 0) TQ_CALL(::get_Ar0n_batch(uuid_, ND, 1, &T, &rho, zc.data(), (int)zc.size(), out.data(), 0, ERRARGS));
    else if (ND == 0) TQ_CALL(::get_Arn0_batch(uuid_, NT, 1, &T, &rho, zc.data(), (int)zc.size(), out.data(), 0, ERRARGS));
    else {
        std::vector<int> nt, nd;
        for (int i = 0; i <= NT; ++i) for (int j = 0; j <= ND; ++j) { nt.push_back(i); nd.push_back(j); }
        TQ_CALL(::get_Arxy_batch(uuid_, (int)nt.size(), nt.data(), nd.data(), 1, &T, &rho, zc.data(), (int)zc.size(), out.data(), 0, ERRARGS));
    }
    return out;
}

#define X(i, j) double B200Model::get_Ar##i##j(const double T, const double rho, const REArrayd& z) const { \
        const EArrayd zc = z; double v = 0; const int nt = i, nd = j; \
        TQ_CALL(::get_Arxy_batch(uuid_, 1, &nt, &nd, 1, &T, &rho, zc.data(), (int)zc.size(), &v, 0, ERRARGS)); return v; }
ARXY_args
#undef X
#define X(i) EArrayd B200Model::get_Ar0##i##n(const double T, const double rho, const REArrayd& z) const { return ar_table(0, i, T, rho, z); }
AR0N_args
#undef X
#define X(i) EArrayd B200Model::get_Ar##i##0n(const double T, const double rho, const REArrayd& z) const { return ar_table(i, 0, T, rho, z); }
ARN0_args
#undef X

// extended-precision probes of the reference's test-suite (boost::multiprecision, deriv_adapter.hpp:148-171): not part of the GPU path
double B200Model::get_Ar01ep(const double, const double, const EArrayd&) const { throw teqp::NotImplementedError("get_Ar01ep: no extended-precision path on the GPU"); }
double B200Model::get_Ar02ep(const double, const double, const EArrayd&) const { throw teqp::NotImplementedError("get_Ar02ep: no extended-precision path on the GPU"); }
double B200Model::get_Ar03ep(const double, const double, const EArrayd&) const { throw teqp::NotImplementedError("get_Ar03ep: no extended-precision path on the GPU"); }

double B200Model::get_reducing_density(const EArrayd& z) const { double v = 0; TQ_CALL(teqpb200_get_reducing(uuid_, 1, z.data(), (int)z.size(), &v, ERRARGS)); return v; }
double B200Model::get_reducing_temperature(const EArrayd& z) const { double v = 0; TQ_CALL(teqpb200_get_reducing(uuid_, 0, z.data(), (int)z.size(), &v, ERRARGS)); return v; }

double B200Model::get_dmBnvirdTm(const int Nderiv, const int NTderiv, const double T, const EArrayd& z) const {
    double v = 0;
    TQ_CALL(::get_dmBnvirdTm(uuid_, Nderiv, NTderiv, T, z.data(), (int)z.size(), &v, ERRARGS));
    return v;
}
double B200Model::get_B2vir(const double T, const EArrayd& z) const { return get_dmBnvirdTm(2, 0, T, z); }
std::map<int, double> B200Model::get_Bnvir(const int Nderiv, const double T, const EArrayd& z) const {
    std::map<int, double> o;
    for (int n = 2; n <= Nderiv; ++n) o[n] = get_dmBnvirdTm(n, 0, T, z);   // VirialDerivatives::get_Bnvir (derivs.hpp:715-736) returns B_2..B_N
    return o;
}
double B200Model::get_B12vir(const double T, const EArrayd& z) const {
    // derivs.hpp:840-853: B12 = (B2(z) - z0^2 B2(pure 0) - z1^2 B2(pure 1)) / (2 z0 z1), binary mixtures only
    if (z.size() != 2) throw teqp::InvalidArgument("length of mole fraction vector must be 2 in get_B12vir");
    EArrayd a(2), b(2); a << 1.0, 0.0; b << 0.0, 1.0;
    return (get_B2vir(T, z) - z[0] * z[0] * get_B2vir(T, a) - z[1] * z[1] * get_B2vir(T, b)) / (2 * z[0] * z[1]);
}

double B200Model::get_ATrhoXi(const double T, const int NT, const double rho, int ND, const EArrayd& z, const int i, const int NXi) const {
    double v = 0; TQ_CALL(::get_ATrhoXi(uuid_, T, NT, rho, ND, z.data(), (int)z.size(), i, NXi, &v, ERRARGS)); return v;
}
double B200Model::get_ATrhoXiXj(const double T, const int NT, const double rho, int ND, const EArrayd& z, const int i, const int NXi, const int j, const int NXj) const {
    double v = 0; TQ_CALL(::get_ATrhoXiXj(uuid_, T, NT, rho, ND, z.data(), (int)z.size(), i, NXi, j, NXj, &v, ERRARGS)); return v;
}
double B200Model::get_ATrhoXiXjXk(const double T, const int NT, const double rho, int ND, const EArrayd& z, const int i, const int NXi, const int j, const int NXj,
                                  const int k, const int NXk) const {
    double v = 0; TQ_CALL(::get_ATrhoXiXjXk(uuid_, T, NT, rho, ND, z.data(), (int)z.size(), i, NXi, j, NXj, k, NXk, &v, ERRARGS)); return v;
}
double B200Model::get_AtaudeltaXi(const double tau, const int Ntau, const double delta, int Ndelta, const EArrayd& z, const int i, const int NXi) const {
    double v = 0; TQ_CALL(::get_AtaudeltaXi(uuid_, tau, Ntau, delta, Ndelta, z.data(), (int)z.size(), i, NXi, &v, ERRARGS)); return v;
}
double B200Model::get_AtaudeltaXiXj(const double tau, const int Ntau, const double delta, int Ndelta, const EArrayd& z, const int i, const int NXi, const int j,
                                    const int NXj) const {
    double v = 0; TQ_CALL(::get_AtaudeltaXiXj(uuid_, tau, Ntau, delta, Ndelta, z.data(), (int)z.size(), i, NXi, j, NXj, &v, ERRARGS)); return v;
}
double B200Model::get_AtaudeltaXiXjXk(const double tau, const int Ntau, const double delta, int Ndelta, const EArrayd& z, const int i, const int NXi, const int j,
                                      const int NXj, const int k, const int NXk) const {
    double v = 0; TQ_CALL(::get_AtaudeltaXiXjXk(uuid_, tau, Ntau, delta, Ndelta, z.data(), (int)z.size(), i, NXi, j, NXj, k, NXk, &v, ERRARGS)); return v;
}

// ---- isochoric family: one call of isochoric_properties_batch with npts = 1, only the requested output non-NULL ---------------
struct B200Model::IsoOne {
    double *pr = nullptr, *splus = nullptr, *dpdT = nullptr, *d2PsirdTdrhoi = nullptr, *dpdrhovec = nullptr, *pmv = nullptr, *chempotVLE = nullptr,
           *dchempotdT = nullptr, *dlnphi_dT = nullptr, *dlnphi_drho = nullptr, *dlnphi_dv = nullptr, *dlnphi_dx = nullptr;
};
void B200Model::iso(double T, const EArrayd& rv, IsoOne& o) const {
    TQ_CALL(::isochoric_properties_batch(uuid_, 1, &T, rv.data(), (int)rv.size(), o.pr, o.splus, o.dpdT, o.d2PsirdTdrhoi, o.dpdrhovec, o.pmv, o.chempotVLE,
                                         o.dchempotdT, o.dlnphi_dT, o.dlnphi_drho, o.dlnphi_dv, o.dlnphi_dx, 0, ERRARGS));
}
double B200Model::get_pr(const double T, const EArrayd& rv) const { double v = 0; IsoOne o; o.pr = &v; iso(T, rv, o); return v; }
double B200Model::get_splus(const double T, const EArrayd& rv) const { double v = 0; IsoOne o; o.splus = &v; iso(T, rv, o); return v; }
double B200Model::get_dpdT_constrhovec(const double T, const EArrayd& rv) const { double v = 0; IsoOne o; o.dpdT = &v; iso(T, rv, o); return v; }
EArrayd B200Model::build_Psir_gradient_autodiff(const double T, const EArrayd& rv) const {
    EArrayd g(rv.size());
    TQ_CALL(::build_Psir_fgradHessian_batch(uuid_, 1, &T, rv.data(), (int)rv.size(), nullptr, g.data(), nullptr, 0, ERRARGS));
    return g;
}
EArrayd B200Model::get_chempotVLE_autodiff(const double T, const EArrayd& rv) const { EArrayd v(rv.size()); IsoOne o; o.chempotVLE = v.data(); iso(T, rv, o); return v; }
EArrayd B200Model::get_dchempotdT_autodiff(const double T, const EArrayd& rv) const { EArrayd v(rv.size()); IsoOne o; o.dchempotdT = v.data(); iso(T, rv, o); return v; }
EArrayd B200Model::get_fugacity_coefficients(const double T, const EArrayd& rv) const {
    EArrayd v(rv.size());
    TQ_CALL(::get_fugacity_coefficients_batch(uuid_, 1, &T, rv.data(), (int)rv.size(), v.data(), 0, ERRARGS));
    return v;
}
EArrayd B200Model::get_partial_molar_volumes(const double T, const EArrayd& rv) const { EArrayd v(rv.size()); IsoOne o; o.pmv = v.data(); iso(T, rv, o); return v; }
EArrayd B200Model::build_d2PsirdTdrhoi_autodiff(const double T, const EArrayd& rv) const { EArrayd v(rv.size()); IsoOne o; o.d2PsirdTdrhoi = v.data(); iso(T, rv, o); return v; }
EArrayd B200Model::get_dpdrhovec_constT(const double T, const EArrayd& rv) const { EArrayd v(rv.size()); IsoOne o; o.dpdrhovec = v.data(); iso(T, rv, o); return v; }
EMatrixd B200Model::build_Psir_Hessian_autodiff(const double T, const EArrayd& rv) const {
    const int n = (int)rv.size();
    EMatrixd H(n, n);   // symmetric: the row-major [i*n+j] layout of the C ABI equals the column-major Eigen layout
    TQ_CALL(::build_Psir_fgradHessian_batch(uuid_, 1, &T, rv.data(), n, nullptr, nullptr, H.data(), 0, ERRARGS));
    return H;
}
EMatrixd B200Model::build_Psi_Hessian_autodiff(const double T, const EArrayd& rv) const {
    // derivs.hpp:1076-1084: the residual Hessian plus the ideal-gas diagonal R(x) T / rho_i
    EMatrixd H = build_Psir_Hessian_autodiff(T, rv);
    const EArrayd x = rv / rv.sum();
    const double R = get_R(x);
    for (Eigen::Index i = 0; i < rv.size(); ++i) H(i, i) += R * T / rv[i];
    return H;
}
std::tuple<double, Eigen::ArrayXd, Eigen::MatrixXd> B200Model::build_Psir_fgradHessian_autodiff(const double T, const EArrayd& rv) const {
    const int n = (int)rv.size();
    double f = 0; Eigen::ArrayXd g(n); Eigen::MatrixXd H(n, n);
    TQ_CALL(::build_Psir_fgradHessian_batch(uuid_, 1, &T, rv.data(), n, &f, g.data(), H.data(), 0, ERRARGS));
    return std::make_tuple(f, g, H);
}
Eigen::ArrayXd B200Model::get_Psir_sigma_derivs(const double T, const EArrayd& rv, const EArrayd& v) const {
    Eigen::ArrayXd o(5);
    TQ_CALL(::get_Psir_sigma_derivs_batch(uuid_, 1, &T, rv.data(), v.data(), (int)rv.size(), o.data(), 0, ERRARGS));
    return o;
}
EArray33d B200Model::get_deriv_mat2(const double T, double rho, const EArrayd& z) const {
    // DerivativeHolderSquare<2> (derivs.hpp:1512-1536) fills (0,0..2), (1..2,0) and (1,1); the full 3x3 block comes from one sweep here
    const Eigen::ArrayXd t = ar_table(2, 2, T, rho, z);
    EArray33d m;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m(i, j) = t[i * 3 + j];
    return m;
}

// ---- batched extensions ---------------------------------------------------------------------------------------------------------
namespace {
struct Xarg { const double* p; int flags; };
Xarg xarg(const REMatrixd& z, Eigen::Index npts, int ncomp, EMatrixd& keep) {
    if (z.size() == ncomp && npts != 1) { keep = z; return {keep.data(), TEQPB200_BROADCAST_X}; }
    if (z.rows() != npts || z.cols() != ncomp) throw teqp::InvalidArgument("molefrac must be [npts x Ncomp] (or one composition of size Ncomp)");
    keep = z;   // column-major [npts x Ncomp] == SoA rows [Ncomp][npts]
    return {keep.data(), 0};
}
}  // namespace
EMatrixd B200Model::get_Ar0n_batch(int nmax, const REArrayd& T, const REArrayd& rho, const REMatrixd& z) const {
    const EArrayd Tc = T, rc = rho; EMatrixd keep; const Xarg x = xarg(z, Tc.size(), ncomp_, keep);
    EMatrixd out(Tc.size(), nmax + 1);
    TQ_CALL(::get_Ar0n_batch(uuid_, nmax, Tc.size(), Tc.data(), rc.data(), x.p, ncomp_, out.data(), x.flags, ERRARGS));
    return out;
}
EMatrixd B200Model::get_Arn0_batch(int nmax, const REArrayd& T, const REArrayd& rho, const REMatrixd& z) const {
    const EArrayd Tc = T, rc = rho; EMatrixd keep; const Xarg x = xarg(z, Tc.size(), ncomp_, keep);
    EMatrixd out(Tc.size(), nmax + 1);
    TQ_CALL(::get_Arn0_batch(uuid_, nmax, Tc.size(), Tc.data(), rc.data(), x.p, ncomp_, out.data(), x.flags, ERRARGS));
    return out;
}
EMatrixd B200Model::get_Arxy_batch(const std::vector<std::pair<int, int>>& pairs, const REArrayd& T, const REArrayd& rho, const REMatrixd& z) const {
    const EArrayd Tc = T, rc = rho; EMatrixd keep; const Xarg x = xarg(z, Tc.size(), ncomp_, keep);
    std::vector<int> nt, nd;
    for (auto& p : pairs) { nt.push_back(p.first); nd.push_back(p.second); }
    EMatrixd out(Tc.size(), (Eigen::Index)pairs.size());
    TQ_CALL(::get_Arxy_batch(uuid_, (int)pairs.size(), nt.data(), nd.data(), Tc.size(), Tc.data(), rc.data(), x.p, ncomp_, out.data(), x.flags, ERRARGS));
    return out;
}
EMatrixd B200Model::get_ln_fugacity_coefficients_batch(const REArrayd& T, const REMatrixd& rhovec) const {
    const EArrayd Tc = T; const EMatrixd rv = rhovec;
    if (rv.rows() != Tc.size() || rv.cols() != ncomp_) throw teqp::InvalidArgument("rhovec must be [npts x Ncomp]");
    EMatrixd out(Tc.size(), ncomp_);
    TQ_CALL(::get_ln_fugacity_coefficients_batch(uuid_, Tc.size(), Tc.data(), rv.data(), ncomp_, out.data(), 0, ERRARGS));
    return out;
}

EMatrixd B200Model::pure_VLE_T_batch(const REArrayd& T, const REArrayd& rhoL0, const REArrayd& rhoV0, int maxiter, const std::optional<EArrayd>& molefracs) const {
    const EArrayd Tc = T, L0 = rhoL0, V0 = rhoV0;
    if (L0.size() != Tc.size() || V0.size() != Tc.size()) throw teqp::InvalidArgument("T, rhoL0 and rhoV0 must have the same length");
    EArrayd z = molefracs ? molefracs.value() : EArrayd::Ones(1).eval();
    if (z.size() != ncomp_) throw teqp::InvalidArgument("molefracs must have one entry per component");
    EMatrixd out(Tc.size(), 4);   // column-major: four contiguous columns
    TQ_CALL(::pure_VLE_T_batch(uuid_, Tc.size(), Tc.data(), L0.data(), V0.data(), maxiter, z.data(), ncomp_, out.data(), out.data() + Tc.size(), out.data() + 2 * Tc.size(),
                               out.data() + 3 * Tc.size(), 0, ERRARGS));
    return out;
}

std::unique_ptr<teqp::cppinterface::AbstractModel> factory(const nlohmann::json& j) { return std::make_unique<B200Model>(j); }

}  // namespace teqp_b200
