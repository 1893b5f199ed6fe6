// GERG / multifluid over jets: virial tables (unit-seeded density jets at rho = 0; the factorised kernels work in delta (1 + s) and cannot
// be evaluated there) and get_Psir_sigma_derivs
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_jet.cuh"
namespace tq {
void gergjet_launch_virial(const GergProgram& p, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table<GergJetDev>(p, p.nc, NT, ND, in, out, st); }
void gergjet_launch_isochoric(const GergProgram& p, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) { generic_launch_isochoric<GergJetDev>(p, p.nc, in, o, st); }
void gergjet_launch_sigma(const GergProgram& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<GergJetDev>(p, p.nc, in, v, out, st); }
}
