// genericSAFT with a PC-SAFT non-polar term: get_Psir_sigma_derivs
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_sigma(const PcAssocParams& p, const DeviceBatch& in, const double* v, double* out, cudaStream_t st) { generic_launch_sigma<PcAssocDev>(p, p.pc.nc, in, v, out, st); }
}
