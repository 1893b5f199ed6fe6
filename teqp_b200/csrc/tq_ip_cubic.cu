// derived isochoric properties (IsoPropsOut) for the cubic family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_cubic.cuh"
namespace tq {
void cubic_launch_isoprops(const CubicParams& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<CubicDev>(p, p.nc, in, o, st); }
}
