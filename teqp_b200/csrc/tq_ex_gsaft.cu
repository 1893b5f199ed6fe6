// get_Psir_sigma_derivs (derivs.hpp:1496-1509) for SAFT-VR-Mie / genericSAFT (virial tables: the ar-table kernels with the unit-seed flag)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_sigma(const GsaftParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<GsaftDev>(p, p.mie.nc, in, v, out, st); }
}
