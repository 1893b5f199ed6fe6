// get_Psir_sigma_derivs (derivs.hpp:1496-1509) for the PC-SAFT family (its virial tables use the ar-table kernels with the unit-seed flag)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_sigma(const PcsaftParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<PcsaftDev>(p, p.nc, in, v, out, st); }
}
