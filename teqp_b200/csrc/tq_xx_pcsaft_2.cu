// cross-composition derivative tables for the pcsaft family by directional jets (tq_generic_dir.cuh), part 2: get_ATrhoXiXj
#include "tq_common.hpp"
#include "tq_generic_dir.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_cross2(const PcsaftParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { dir_launch_cross<PcsaftDev, 2>(p, p.nc, in, i, j, k, out, st); }
}
