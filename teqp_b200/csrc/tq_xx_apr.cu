// cross-composition derivative tables (five-level jet sweeps) and get_Psir_sigma_derivs for the advancedPRaEres family, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_apr.cuh"
namespace tq {
void apr_launch_ATrhoXiXj(const AprParams& p, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) { generic_launch_ATrhoXiXj<AprDev>(p, p.nc, in, i, j, out, st); }
void apr_launch_ATrhoXiXjXk(const AprParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { generic_launch_ATrhoXiXjXk<AprDev>(p, p.nc, in, i, j, k, out, st); }
void apr_launch_sigma(const AprParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<AprDev>(p, p.nc, in, v, out, st); }
}
