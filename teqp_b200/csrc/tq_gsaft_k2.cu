// SAFT-VR-Mie / genericSAFT kernels, part 2 (see tq_gsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_ar_mixed(const GsaftParams& p, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table_mixed<GsaftDev>(p, p.mie.nc, NT, ND, in, out, st); }
}
