// cross-composition derivative tables (five-level jet sweeps) and get_Psir_sigma_derivs for the ideal-gas family, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_ideal.cuh"
namespace tq {
void ideal_launch_ATrhoXiXj(const IdealParams& p, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) { generic_launch_ATrhoXiXj<IdealDev>(p, p.nc, in, i, j, out, st); }
void ideal_launch_ATrhoXiXjXk(const IdealParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { generic_launch_ATrhoXiXjXk<IdealDev>(p, p.nc, in, i, j, k, out, st); }
void ideal_launch_sigma(const IdealParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<IdealDev>(p, p.nc, in, v, out, st); }
}
