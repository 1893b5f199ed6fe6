// cross-composition derivative tables (get_ATrhoXiXj, get_ATrhoXiXjXk) for the cubic family: five-level jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_cubic.cuh"
namespace tq {
void cubic_launch_ATrhoXiXj(const CubicParams& p, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) { generic_launch_ATrhoXiXj<CubicDev>(p, p.nc, in, i, j, out, st); }
void cubic_launch_ATrhoXiXjXk(const CubicParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { generic_launch_ATrhoXiXjXk<CubicDev>(p, p.nc, in, i, j, k, out, st); }
}
