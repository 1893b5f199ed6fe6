// PC-SAFT kernels, part 2 (see tq_pcsaft.cu)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_ar_mixed(const PcsaftParams& p, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table_mixed<PcsaftDev>(p, p.nc, NT, ND, in, out, st); }
}
