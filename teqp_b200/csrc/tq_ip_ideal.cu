// derived isochoric properties (IsoPropsOut) for the ideal-gas family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_ideal.cuh"
namespace tq {
void ideal_launch_isoprops(const IdealParams& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<IdealDev>(p, p.nc, in, o, st); }
}
