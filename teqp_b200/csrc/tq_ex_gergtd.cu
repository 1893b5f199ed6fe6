// get_AtaudeltaXi / XiXj / XiXjXk (derivs.hpp:341-414) for the JSON-driven multifluid models: the generic composition-derivative sweeps
// over GergTauDeltaDev (alphar as a function of tau, delta, x; tq_gerg_jet.cuh), driven with T := 1/tau and rho := delta
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_jet.cuh"
#include "tq_stack.hpp"
namespace tq {
void gergtd_launch(const GergProgram& p, int which, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) {
    ensure_stack({TQ_K(generic_ATrhoXi_kernel<GergTauDeltaDev, GergTauDeltaDev::kMaxC>), TQ_K(generic_ATrhoXiXj_kernel<GergTauDeltaDev>), TQ_K(generic_ATrhoXiXjXk_kernel<GergTauDeltaDev>)});
    if (which == 1) generic_launch_ATrhoXi<GergTauDeltaDev>(p, p.nc, in, out, st);
    else if (which == 2) generic_launch_ATrhoXiXj<GergTauDeltaDev>(p, p.nc, in, i, j, out, st);
    else generic_launch_ATrhoXiXjXk<GergTauDeltaDev>(p, p.nc, in, i, j, k, out, st);
}
}
