// PC-SAFT kernels, part 1 (see tq_pcsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_ar_rho(const PcsaftParams& p, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table_rho<PcsaftDev>(p, p.nc, ND, in, out, st); }
}
