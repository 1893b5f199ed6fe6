// GERG-2004/2008 over jets: the generic get_ATrhoXi sweep (see tq_gerg_jet.cuh)
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_iso.cuh"
#include "tq_gerg_jet.cuh"
#include "tq_stack.hpp"
namespace tq {
void gerg_launch_ATrhoXi(const GergProgram& P, const DeviceBatch& in, double* out, cudaStream_t st) {
    ensure_stack({TQ_K(generic_ATrhoXi_kernel<GergJetDev, GergJetDev::kMaxC>)});
    generic_launch_ATrhoXi<GergJetDev>(P, P.nc, in, out, st);
}
}
