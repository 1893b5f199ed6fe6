// cross-composition derivative tables (get_ATrhoXiXj, get_ATrhoXiXjXk) for the gergjet family: five-level jet sweeps, see tq_generic.cuh
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_jet.cuh"
#include "tq_stack.hpp"
namespace tq {
void gergjet_launch_ATrhoXiXj(const GergProgram& p, const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) {
    ensure_stack({TQ_K(generic_ATrhoXiXj_kernel<GergJetDev>)});
    generic_launch_ATrhoXiXj<GergJetDev>(p, p.nc, in, i, j, out, st);
}
void gergjet_launch_ATrhoXiXjXk(const GergProgram& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) {
    ensure_stack({TQ_K(generic_ATrhoXiXjXk_kernel<GergJetDev>)});
    generic_launch_ATrhoXiXjXk<GergJetDev>(p, p.nc, in, i, j, k, out, st);
}
}
