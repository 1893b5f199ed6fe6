// derived isochoric properties (IsoPropsOut) for the RK-PR family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_rkpr.cuh"
namespace tq {
void rkpr_launch_isoprops(const RkprParams& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<RkprDev>(p, p.nc, in, o, st); }
}
