// derived isochoric properties (IsoPropsOut) for the QCPR family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_qcpr.cuh"
namespace tq {
void qcpr_launch_isoprops(const QcprParams& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<QcprDev>(p, p.nc, in, o, st); }
}
