// SAFT-VR-Mie / genericSAFT kernels, part 4 (see tq_gsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_isochoric(const GsaftParams& p, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) { generic_launch_isochoric<GsaftDev>(p, p.mie.nc, in, o, st); }
}
