// cross-composition derivative tables (five-level jet sweeps) and get_Psir_sigma_derivs for the QCPR family, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_qcpr.cuh"
namespace tq {
void qcpr_launch_ATrhoXiXj(const QcprParams& p, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) { generic_launch_ATrhoXiXj<QcprDev>(p, p.nc, in, i, j, out, st); }
void qcpr_launch_ATrhoXiXjXk(const QcprParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { generic_launch_ATrhoXiXjXk<QcprDev>(p, p.nc, in, i, j, k, out, st); }
void qcpr_launch_sigma(const QcprParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<QcprDev>(p, p.nc, in, v, out, st); }
}
