// derived isochoric properties (IsoPropsOut) for the pcsaft family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"
namespace tq {
void pcsaft_launch_isoprops(const PcsaftParams& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<PcsaftDev>(p, p.nc, in, o, st); }
}
