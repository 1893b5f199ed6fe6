// SAFT-VR-Mie / genericSAFT kernels, part 3 (see tq_gsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_ATrhoXi(const GsaftParams& p, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ATrhoXi<GsaftDev>(p, p.mie.nc, in, out, st); }
}
