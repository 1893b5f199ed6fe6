// Shared host-side declarations of libteqpb200: error types (numeric codes follow the reference's exception hierarchy,
// /root/reference/include/teqp/exceptions.hpp:17-64, and the C ABI's 30 / 9999 codes, /root/reference/interface/C/teqpc.cpp:45-73),
// CUDA error checking, and the abstract batched model that every kernel family implements.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "tq_json.hpp"

namespace tq {

struct Error : std::exception {
    int code;
    std::string msg;
    Error(int c, std::string m) : code(c), msg(std::move(m)) {}
    const char* what() const noexcept override { return msg.c_str(); }
};
struct InvalidArgument : Error { explicit InvalidArgument(const std::string& m) : Error(1, m) {} };
struct IterationFailure : Error { explicit IterationFailure(const std::string& m) : Error(100, m) {} };
struct InvalidValue : Error { explicit InvalidValue(const std::string& m) : Error(101, m) {} };
struct NotImplementedError : Error { explicit NotImplementedError(const std::string& m) : Error(200, m) {} };
struct DeviceError : Error { explicit DeviceError(const std::string& m) : Error(500, m) {} };

#define TQ_CUDA(call)                                                                                              \
    do {                                                                                                           \
        cudaError_t e__ = (call);                                                                                  \
        if (e__ != cudaSuccess)                                                                                    \
            throw tq::DeviceError(std::string("CUDA error: ") + cudaGetErrorString(e__) + " at " __FILE__ ":" +    \
                                  std::to_string(__LINE__) + " in " #call);                                        \
    } while (0)

constexpr int kMaxDerivT = 4;   // highest order in 1/T any entry point asks for
constexpr int kMaxDerivD = 6;   // highest order in rho (get_Ar06n)

// A batch of state points resident on ONE device, structure-of-arrays:
//   T[npts], rho[npts], x[ncomp][npts]  (x_stride = npts) -- or x[ncomp] broadcast to all points (x_stride = 1, x_bcast = 1).
struct DeviceBatch {
    long long npts = 0;
    const double* T = nullptr;
    const double* rho = nullptr;
    const double* x = nullptr;
    long long x_stride = 0;  // distance between component rows
    int x_bcast = 0;         // 1: one composition x[ncomp] for every point (x_stride = 1)
    int rho_unit_seed = 0;   // 1: the density jet is rho(s) = rho0 + s (UNSCALED derivatives d^m/drho^m, needed at rho0 = 0 for the virial
                             //    coefficients, derivs.hpp:693-867) instead of rho0 (1 + s); honoured by the generic jet kernels only
};

// What an isochoric-family launch should write; null pointers are skipped.  rhovec is x*rho formed in-kernel exactly as
// the reference's callers do; all outputs SoA with row stride npts.
struct IsochoricOut {
    double* Psir = nullptr;        // [npts]
    double* grad = nullptr;        // [ncomp][npts]   dPsir/drho_i
    double* hess = nullptr;        // [ncomp*ncomp][npts] row-major (i*ncomp+j)
    double* lnphi = nullptr;       // [ncomp][npts]
    double* phi = nullptr;         // [ncomp][npts]
};

// Derived isochoric properties (derivs.hpp:962-978, 1196-1201, 1264-1437, 1482-1494); null pointers are skipped, SoA, row stride npts.
struct IsoPropsOut {
    double* pr = nullptr;              // [npts]        get_pr
    double* splus = nullptr;           // [npts]        get_splus
    double* dpdT = nullptr;            // [npts]        get_dpdT_constrhovec
    double* d2PsirdTdrhoi = nullptr;   // [ncomp][npts] build_d2PsirdTdrhoi_autodiff
    double* dpdrhovec = nullptr;       // [ncomp][npts] get_dpdrhovec_constT
    double* pmv = nullptr;             // [ncomp][npts] get_partial_molar_volumes
    double* chempotVLE = nullptr;      // [ncomp][npts] get_chempotVLE_autodiff
    double* dchempotdT = nullptr;      // [ncomp][npts] get_dchempotdT_autodiff
    double* dlnphi_dT = nullptr;       // [ncomp][npts] get_d_ln_fugacity_coefficients_dT_constrhovec
    double* dlnphi_drho = nullptr;     // [ncomp][npts] get_d_ln_fugacity_coefficients_drho_constTmolefracs
    double* dlnphi_dv = nullptr;       // [ncomp][npts] get_d_ln_fugacity_coefficients_dv_constTmolefracs
    double* dlnphi_dx = nullptr;       // [ncomp*ncomp][npts] get_d_ln_fugacity_coefficients_dmolefracs_constTrho, (i*ncomp+j) = d ln phi_i / d x_j
};

class Model {
public:
    virtual ~Model() = default;
    virtual int ncomp() const = 0;
    virtual const char* family() const = 0;
    // gas constant R(x) (host; /root/reference/include/teqp/cpp/teqpcpp.hpp:102 get_R)
    virtual double R(const double* x, int n) const = 0;

    // Lambda^r_{nm} for every 0<=n<=NT, 0<=m<=ND in one sweep: out[(n*(ND+1)+m)][npts]
    // (reference: derivs.hpp:150-238 get_Agenxy, :543-624 get_Ar0n / get_Arn0)
    virtual void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const = 0;
    // false: launch_ar_table(NT, ND, ...) would first have to generate / load a kernel (which allocates); asked by the no-allocation entry point
    virtual bool ar_table_ready(int NT, int ND) const { (void)NT; (void)ND; return true; }

    // composition-derivative table (derivs.hpp:247-292): for each component i and each (NT,ND,NXi) with NT+ND<=2, NXi<=3:
    // out[((i*3+nt)*3+nd)*4+nxi][npts] = (1/T)^nt rho^nd d^(nt+nd+nxi) alphar / d(1/T)^nt drho^nd dx_i^nxi
    virtual void launch_ATrhoXi_table(const DeviceBatch&, double*, cudaStream_t) const {
        throw NotImplementedError(std::string("get_ATrhoXi table is not implemented for model family ") + family());
    }
    // cross-composition tables (derivs.hpp:294-339, 423-465), layouts in tq_generic.cuh
    virtual void launch_ATrhoXiXj_table(const DeviceBatch&, int, int, double*, cudaStream_t) const {
        throw NotImplementedError(std::string("get_ATrhoXiXj is not implemented for model family ") + family());
    }
    virtual void launch_ATrhoXiXjXk_table(const DeviceBatch&, int, int, int, double*, cudaStream_t) const {
        throw NotImplementedError(std::string("get_ATrhoXiXjXk is not implemented for model family ") + family());
    }
    // isochoric family (derivs.hpp:1031-1124, 1208-1230)
    virtual void launch_isochoric(const DeviceBatch&, const IsochoricOut&, cudaStream_t) const {
        throw NotImplementedError(std::string("isochoric derivatives are not implemented for model family ") + family());
    }
    virtual void launch_isoprops(const DeviceBatch&, const IsoPropsOut&, cudaStream_t) const {
        throw NotImplementedError(std::string("derived isochoric properties are not implemented for model family ") + family());
    }
    // virial table (VirialDerivatives, derivs.hpp:693-867): out[(n*(ND+1)+m)][npts] = (1/T)^n d^(n+m) alphar / d(1/T)^n drho^m at the batch's
    // rho (= 0), i.e. the Lambda table with UNSCALED density derivatives; `in.rho_unit_seed` is set by the caller
    virtual void launch_virial_table(int, int, const DeviceBatch&, double*, cudaStream_t) const {
        throw NotImplementedError(std::string("virial coefficients are not implemented for model family ") + family());
    }
    // get_AtaudeltaXi / XiXj / XiXjXk tables (derivs.hpp:341-414): in.T holds 1/tau, in.rho holds delta; which = 1: out as launch_ATrhoXi_table,
    // 2: as launch_ATrhoXiXj_table (i, j), 3: as launch_ATrhoXiXjXk_table (i, j, k).  Only models with alphar_taudelta (multifluid).
    virtual void launch_Ataudelta_table(int, const DeviceBatch&, int, int, int, double*, cudaStream_t) const {
        throw NotImplementedError("Cannot take derivatives of a class that doesn't define the alphar_taudelta method");
    }
    // IsochoricDerivatives::get_Psir_sigma_derivs (derivs.hpp:1496-1509): d^k Psir(T, rhovec + sigma v) / dsigma^k, k = 0..4;
    // in.x = rhovec rows, v[ncomp][npts], out[5][npts]
    virtual void launch_sigma_derivs(const DeviceBatch&, const double*, double*, cudaStream_t) const {
        throw NotImplementedError(std::string("get_Psir_sigma_derivs is not implemented for model family ") + family());
    }
    // reducing functions (teqpcpp.hpp get_reducing_density / get_reducing_temperature); host, single composition
    virtual bool reducing(const double*, int, double*, double*) const { return false; }
    virtual std::string meta() const { return "{}"; }
    // families whose kernels are generated per model (tq_gerg_jit.hpp): the generated translation unit + its constant pool, and the
    // compiled cubin (no device needed: used by the CPU test tier and by tools/ to inspect the SASS)
    virtual std::string jit_source(int, int, std::vector<double>*) const { return std::string(); }
    virtual std::vector<char> jit_cubin(int, int) const { return {}; }
    // the POD parameter block the kernels receive (test harnesses replay the per-point device functions on it)
    virtual const void* params_blob(size_t* nbytes) const = 0;
};

// factories (one translation unit per model family, like /root/reference/interface/CPP/model_factory_*.cpp)
std::unique_ptr<Model> make_cubic_model(const std::string& kind, const Json& spec);
std::unique_ptr<Model> make_vdw_model(const std::string& kind, const Json& spec);
std::unique_ptr<Model> make_gerg_model(const std::string& kind, const Json& spec);
std::unique_ptr<Model> make_pcsaft_model(const Json& spec);
std::unique_ptr<Model> make_saftvrmie_model(const Json& spec);
std::unique_ptr<Model> make_genericsaft_model(const Json& spec);
std::unique_ptr<Model> make_multifluid_model(const Json& spec);
std::unique_ptr<Model> make_model(const Json& j);

inline int grid_for(long long npts, int block, int max_blocks) {
    long long g = (npts + block - 1) / block;
    if (g > max_blocks) g = max_blocks;
    if (g < 1) g = 1;
    return static_cast<int>(g);
}
int sm_count();  // SMs of the current device (cached)

}  // namespace tq
