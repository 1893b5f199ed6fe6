// SAFT-VR-Mie / genericSAFT kernels, part 1 (see tq_gsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_ar_rho(const GsaftParams& p, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table_rho<GsaftDev>(p, p.mie.nc, ND, in, out, st); }
}
