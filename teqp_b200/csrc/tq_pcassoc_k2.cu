// genericSAFT with a PC-SAFT non-polar term: Lambda tables, temperature / mixed shapes
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_ar_mixed(const PcAssocParams& p, int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) { generic_launch_ar_table_mixed<PcAssocDev>(p, p.pc.nc, NT, ND, in, out, st); }
}
