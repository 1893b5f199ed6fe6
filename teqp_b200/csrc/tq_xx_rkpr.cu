// cross-composition derivative tables (five-level jet sweeps)and get_Psir_sigma_derivs for the RK-PR family
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_rkpr.cuh"
namespace tq {
void rkpr_launch_ATrhoXiXj(const RkprParams& p, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) { generic_launch_ATrhoXiXj<RkprDev>(p, p.nc, in, i, j, out, st); }
void rkpr_launch_ATrhoXiXjXk(const RkprParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { generic_launch_ATrhoXiXjXk<RkprDev>(p, p.nc, in, i, j, k, out, st); }
void rkpr_launch_sigma(const RkprParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<RkprDev>(p, p.nc, in, v, out, st); }
}
