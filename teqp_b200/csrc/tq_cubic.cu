// Cubic EOS: JSON -> device parameter block, and the batched kernels.
//
// JSON ingestion mirrors make_canonicalPR / make_canonicalSRK / make_generalizedcubic
// (/root/reference/include/teqp/models/cubics/simple_cubics.hpp:258-340, 343-437) and the vdW factories
// (/root/reference/interface/CPP/teqp_impl_factory.cpp:56-57; /root/reference/include/teqp/models/vdW.hpp:10-128).
//
// The Lambda^r_nm table kernel does NOT run the model through jet arithmetic in rho: for a cubic the density dependence is two
// logarithms whose scaled derivatives are closed forms,
//     rho^m d^m/drho^m [-ln(1-y)]      = (m-1)! (y/(1-y))^m                      y = b rho
//     rho^m d^m/drho^m [ ln(1+D y)]    = (-1)^(m-1) (m-1)! (D y/(1+D y))^m
// and the temperature dependence sits entirely in q(T) = a(T,x)/(R T), carried as a small jet in 1/T.  This keeps the kernel at
// ~100 FP64 instructions per point, which is what it takes to stay near the HBM roofline on B200 (56 B/point, FP64:HBM ~ 3 flop/B).
#include <cstring>

#include "tq_common.hpp"
#include "tq_cubic.cuh"
#include "tq_generic.cuh"

namespace tq {
void cubic_launch_isoprops(const CubicParams&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);
void cubic_launch_virial(const CubicParams&, int NT, int ND, const DeviceBatch&, double*, cudaStream_t);   // tq_ex_cubic.cu
void cubic_launch_sigma(const CubicParams&, const DeviceBatch&, const double*, double*, cudaStream_t);
void cubic_launch_ATrhoXiXj(const CubicParams&, const DeviceBatch&, int, int, double*, cudaStream_t);
void cubic_launch_ATrhoXiXjXk(const CubicParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);

namespace {

template<int NT, int MAXN>
__global__ void __launch_bounds__(256) cubic_ar_table_kernel(const __grid_constant__ CubicParams p, const DeviceBatch in, const int ND,
                                                             double* __restrict__ out) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        const double T = __ldg(in.T + k);
        const double rho = __ldg(in.rho + k);
        double x[MAXN];
#pragma unroll(MAXN <= 4 ? MAXN : 1)
        for (int i = 0; i < MAXN; ++i)
            if (i < p.nc) x[i] = __ldg(in.x + (long long)i * in.x_stride + (in.x_bcast ? 0 : k));
        cubic_table_point<NT, MAXN>(p, T, rho, x, ND, out + k, npts);
    }
}

class CubicModel final : public Model {
public:
    CubicParams p{};
    std::string meta_json;
    int ncomp() const override { return p.nc; }
    const char* family() const override { return p.is_vdw ? "vdW" : "cubic"; }
    double R(const double*, int) const override { return p.R; }
    std::string meta() const override { return meta_json; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }

    template<int NT> void launch_nt(int ND, const DeviceBatch& in, double* out, cudaStream_t st) const {
        const int block = 256;
        const int grid = grid_for(in.npts, block, sm_count() * 16);
        switch (p.nc) {
            case 1: cubic_ar_table_kernel<NT, 1><<<grid, block, 0, st>>>(p, in, ND, out); break;
            case 2: cubic_ar_table_kernel<NT, 2><<<grid, block, 0, st>>>(p, in, ND, out); break;
            case 3: cubic_ar_table_kernel<NT, 3><<<grid, block, 0, st>>>(p, in, ND, out); break;
            case 4: cubic_ar_table_kernel<NT, 4><<<grid, block, 0, st>>>(p, in, ND, out); break;
            default: cubic_ar_table_kernel<NT, kCubicMaxC><<<grid, block, 0, st>>>(p, in, ND, out); break;
        }
        TQ_CUDA(cudaGetLastError());
    }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { cubic_launch_virial(p, NT, ND, in, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { cubic_launch_sigma(p, in, v, out, st); }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override {
        if (NT < 0 || NT > kMaxDerivT || ND < 0 || ND > kMaxDerivD) throw InvalidArgument("derivative order out of range");
        if (in.npts == 0) return;
        switch (NT) {
            case 0: launch_nt<0>(ND, in, out, st); break;
            case 1: launch_nt<1>(ND, in, out, st); break;
            case 2: launch_nt<2>(ND, in, out, st); break;
            case 3: launch_nt<3>(ND, in, out, st); break;
            default: launch_nt<4>(ND, in, out, st); break;
        }
    }
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override {
        generic_launch_ATrhoXi<CubicDev>(p, p.nc, in, out, st);
    }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { cubic_launch_ATrhoXiXj(p, in, i, j, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { cubic_launch_ATrhoXiXjXk(p, in, i, j, k, out, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override { cubic_launch_isoprops(p, in, o, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override {
        generic_launch_isochoric<CubicDev>(p, p.nc, in, o, st);
    }
};

std::vector<std::vector<double>> square_matrix(const Json& j, size_t N, const char* what) {
    // reference: build_square_matrix, /root/reference/include/teqp/json_tools.hpp:45-71 (null or [] -> zeros)
    std::vector<std::vector<double>> m(N, std::vector<double>(N, 0.0));
    if (j.is_null() || (j.is_array() && j.size() == 0)) return m;
    if (!j.is_array()) throw InvalidArgument(std::string("Unable to convert this ") + what + " to a NxN matrix of doubles:" + j.dump());
    if (j.size() != N) throw InvalidArgument(std::string(what) + " needs to be a square matrix the same size as the number of components [" + std::to_string(N) + "]");
    for (size_t i = 0; i < N; ++i) {
        if (!j[i].is_array() || j[i].size() != N) throw InvalidArgument("provided matrix is not square");
        for (size_t k = 0; k < N; ++k) m[i][k] = j[i][k].as_double();
    }
    return m;
}

double pr_m(double w) {
    return (w < 0.491) ? 0.37464 + 1.54226 * w - 0.26992 * (w * w)
                       : 0.379642 + 1.48503 * w - 0.164423 * (w * w) + 0.016666 * (w * w * w);
}

}  // namespace

std::unique_ptr<Model> make_cubic_model(const std::string& kind, const Json& spec) {
    auto m = std::make_unique<CubicModel>();
    CubicParams& p = m->p;
    const std::vector<double> Tc = spec.at("Tcrit / K").as_doubles();
    const std::vector<double> pc = spec.at("pcrit / Pa").as_doubles();
    const std::vector<double> w = spec.at("acentric").as_doubles();
    const size_t N = Tc.size();
    if (pc.size() != N || w.size() != N) throw InvalidArgument("Tcrit / K, pcrit / Pa and acentric must have the same length");
    if (N == 0) throw InvalidArgument("at least one component is required");
    if (N > (size_t)kCubicMaxC) throw InvalidArgument("cubic models support at most " + std::to_string(kCubicMaxC) + " components on the device path");

    std::string type = kind;
    if (kind == "cubic") type = spec.at("type").as_string();
    double OmegaA, OmegaB;
    std::string kname;
    if (type == "PR") {
        p.Delta1 = 1 + std::sqrt(2.0); p.Delta2 = 1 - std::sqrt(2.0);
        OmegaA = 0.45723552892138218938; OmegaB = 0.077796073903888455972;
        kname = "Peng-Robinson";
    } else if (type == "SRK") {
        p.Delta1 = 1; p.Delta2 = 0;
        OmegaA = 1.0 / (9.0 * (std::cbrt(2) - 1)); OmegaB = (std::cbrt(2) - 1) / 3;
        kname = "Soave-Redlich-Kwong";
    } else {
        throw InvalidArgument("Generic cubic EOS are not yet supported (open an issue on github if you want this)");
    }
    p.R = 8.31446261815324;  // CODATA 2017, /root/reference/include/teqp/constants.hpp:10
    if (kind == "cubic" && spec.contains("R / J/mol/K")) p.R = spec.at("R / J/mol/K").as_double();
    p.nc = (int)N;
    p.is_vdw = 0;
    for (size_t i = 0; i < N; ++i) {
        p.ai[i] = OmegaA * ((p.R * Tc[i]) * (p.R * Tc[i])) / pc[i];
        p.bi[i] = OmegaB * p.R * Tc[i] / pc[i];
        p.Tci[i] = Tc[i];
        p.alpha_kind[i] = ALPHA_BASIC;
        p.c1[i] = p.c2[i] = 0.0;
        p.c0[i] = (type == "PR") ? pr_m(w[i]) : 0.48 + 1.574 * w[i] - 0.176 * w[i] * w[i];
    }
    if (kind == "cubic" && spec.contains("alpha")) {
        const Json& ja = spec.at("alpha");
        if (!ja.is_array()) throw InvalidArgument("alpha must be array of objects");
        if (ja.size() != N) throw InvalidArgument("alpha [size " + std::to_string(ja.size()) + "] must be the same size as components [size " + std::to_string(N) + "]");
        for (size_t i = 0; i < N; ++i) {
            const std::string at = ja[i].at("type").as_string();
            if (at == "Twu" || at == "Mathias-Copeman") {
                auto c = ja[i].at("c").as_doubles();
                if (c.size() != 3) throw InvalidArgument("coefficients c for " + at + " alpha function must have length 3");
                p.alpha_kind[i] = (at == "Twu") ? ALPHA_TWU : ALPHA_MC;
                p.c0[i] = c[0]; p.c1[i] = c[1]; p.c2[i] = c[2];
            } else if (at == "PR78") {
                p.alpha_kind[i] = ALPHA_BASIC;
                p.c0[i] = pr_m(ja[i].at("acentric").as_double());
            } else {
                throw InvalidArgument("alpha type is not understood: " + at);
            }
        }
    }
    auto kmat = square_matrix(spec.contains("kmat") ? spec.at("kmat") : Json(), N, "kmat");
    for (size_t i = 0; i < N; ++i)
        for (size_t j = 0; j < N; ++j) p.one_minus_k[i * N + j] = 1 - kmat[i][j];
    char buf[512];
    std::snprintf(buf, sizeof buf, "{\"Delta1\":%.17g,\"Delta2\":%.17g,\"OmegaA\":%.17g,\"OmegaB\":%.17g,\"kind\":\"%s\",\"R / J/mol/K\":%.17g}",
                  p.Delta1, p.Delta2, OmegaA, OmegaB, kname.c_str(), p.R);
    m->meta_json = buf;
    return m;
}

std::unique_ptr<Model> make_vdw_model(const std::string& kind, const Json& spec) {
    auto m = std::make_unique<CubicModel>();
    CubicParams& p = m->p;
    p.is_vdw = 1;
    p.Delta1 = p.Delta2 = 0;
    p.R = 1.380649e-23 * 6.02214076e23;  // k_B N_A, vdW.hpp:21
    if (kind == "vdW1") {
        p.is_vdw = 2;
        p.nc = 1;
        p.ai[0] = spec.at("a").as_double();
        p.bi[0] = spec.at("b").as_double();
        p.Tci[0] = 1.0;
        p.alpha_kind[0] = ALPHA_NONE;
        p.one_minus_k[0] = 1.0;
        p.c0[0] = p.c1[0] = p.c2[0] = 0;
        return m;
    }
    const std::vector<double> Tc = spec.at("Tcrit / K").as_doubles();
    const std::vector<double> pc = spec.at("pcrit / Pa").as_doubles();
    if (Tc.size() != pc.size()) throw InvalidArgument("Sizes of Tc_K " + std::to_string(Tc.size()) + " and pc_Pa" + std::to_string(pc.size()) + " do not agree");
    const size_t N = Tc.size();
    if (N == 0 || N > (size_t)kCubicMaxC) throw InvalidArgument("vdW supports 1.." + std::to_string(kCubicMaxC) + " components on the device path");
    p.nc = (int)N;
    for (size_t i = 0; i < N; ++i) {
        p.ai[i] = 27.0 / 64.0 * std::pow(p.R * Tc[i], 2) / pc[i];
        p.bi[i] = 1.0 / 8.0 * p.R * Tc[i] / pc[i];
        p.Tci[i] = Tc[i];
        p.alpha_kind[i] = ALPHA_NONE;
        p.c0[i] = p.c1[i] = p.c2[i] = 0;
        for (size_t j = 0; j < N; ++j) p.one_minus_k[i * N + j] = 1.0;
    }
    return m;
}

}  // namespace tq
