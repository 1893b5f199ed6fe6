// PC-SAFT kernels, part 3 (see tq_pcsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_ATrhoXi(const PcsaftParams& p, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ATrhoXi<PcsaftDev>(p, p.nc, in, out, st); }
}
