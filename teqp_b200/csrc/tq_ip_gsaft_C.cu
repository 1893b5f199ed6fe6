// derived isochoric properties for SAFT-VR-Mie / genericSAFT, part C of three single-jet-shape kernels (see tq_generic.cuh: NVVM does not
// finish compiling the seven-instantiation kernel for this family; one shape per translation unit does)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_ip_part1(const GsaftParams& p, const DeviceBatch& in, const IsoPropsOut& o, double* Ls, const double* H, cudaStream_t st) {
    generic_launch_ip_part<GsaftDev, 1>(p, p.mie.nc, in, o, Ls, H, st);
}
}
