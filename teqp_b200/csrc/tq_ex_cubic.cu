// virial tables (unit-seeded density jets at rho = 0, derivs.hpp:693-867) and get_Psir_sigma_derivs (derivs.hpp:1496-1509) for the cubic family
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_cubic.cuh"
namespace tq {
void cubic_launch_virial(const CubicParams& p, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table<CubicDev>(p, p.nc, NT, ND, in, out, st); }
void cubic_launch_sigma(const CubicParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<CubicDev>(p, p.nc, in, v, out, st); }
}
