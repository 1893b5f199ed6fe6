// derived isochoric properties (IsoPropsOut) for the advancedPRaEres family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_apr.cuh"
namespace tq {
void apr_launch_isoprops(const AprParams& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<AprDev>(p, p.nc, in, o, st); }
}
