// GERG / multifluid over jets: virial tables (unit-seeded density jets at rho = 0; the factorised kernels work in delta (1 + s) and cannot
// be evaluated there) and get_Psir_sigma_derivs
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_gerg_jet.cuh"
#include "tq_stack.hpp"
namespace tq {
void gergjet_launch_virial(const GergProgram& p, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) {
    ensure_stack({TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 0, 2>), TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 0, 4>), TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 0, 6>), TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 2, 0>), TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 4, 0>), TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 2, 2>), TQ_K(generic_ar_table_kernel<GergJetDev, GergJetDev::kMaxC, 2, 4>)});
    generic_launch_ar_table<GergJetDev>(p, p.nc, NT, ND, in, out, st);
}
void gergjet_launch_isochoric(const GergProgram& p, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) {
    ensure_stack({TQ_K(generic_isochoric_kernel<GergJetDev, GergJetDev::kMaxC>)});
    generic_launch_isochoric<GergJetDev>(p, p.nc, in, o, st);
}
void gergjet_launch_sigma(const GergProgram& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) {
    ensure_stack({TQ_K(generic_sigma_kernel<GergJetDev, GergJetDev::kMaxC>)});
    generic_launch_sigma<GergJetDev>(p, p.nc, in, v, out, st);
}
}
