// derived isochoric properties (IsoPropsOut) for the gergjet family: generic jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_jet.cuh"
#include "tq_stack.hpp"
namespace tq {
void gergjet_launch_isoprops(const GergProgram& p, const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) {
    ensure_stack({TQ_K(generic_isoprops_kernel<GergJetDev>)});
    generic_launch_isoprops<GergJetDev>(p, p.nc, in, o, st);
}
}
