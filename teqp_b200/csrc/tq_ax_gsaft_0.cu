// get_ATrhoXi table for SAFT-VR-Mie / genericSAFT, part 0 of three small-jet kernels: J3<1,1,3> (tq_generic_dir.cuh)
#include "tq_common.hpp"
#include "tq_generic_dir.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_ATrhoXi_part0(const GsaftParams& p, const DeviceBatch& in, double* out, cudaStream_t st) { split_launch_ATrhoXi<GsaftDev, 1, 1>(p, p.mie.nc, in, out, st); }
}
