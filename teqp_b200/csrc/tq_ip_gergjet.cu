// derived isochoric properties (IsoPropsOut) for the gergjet family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_jet.cuh"
namespace tq {
void gergjet_launch_isoprops(const GergProgram& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) { generic_launch_isoprops<GergJetDev>(p, p.nc, in, o, st); }
}
