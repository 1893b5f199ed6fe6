// cross-composition derivative tables for the gsaft family by directional jets (tq_generic_dir.cuh), part 3: get_ATrhoXiXjXk
#include "tq_common.hpp"
#include "tq_generic_dir.cuh"
#include "tq_assoc.cuh"
namespace tq {
void gsaft_launch_cross3(const GsaftParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { dir_launch_cross<GsaftDev, 3>(p, p.mie.nc, in, i, j, k, out, st); }
}
