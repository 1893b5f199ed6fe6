// genericSAFT with a PC-SAFT non-polar term: Lambda tables, density shapes
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_ar_rho(const PcAssocParams& p, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table_rho<PcAssocDev>(p, p.pc.nc, ND, in, out, st); }
}
