// PC-SAFT kernels, part 4 (see tq_pcsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_isochoric(const PcsaftParams& p, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) { generic_launch_isochoric<PcsaftDev>(p, p.nc, in, o, st); }
}
