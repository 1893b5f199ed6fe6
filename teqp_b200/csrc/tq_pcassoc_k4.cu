// genericSAFT with a PC-SAFT non-polar term: isochoric family
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_isochoric(const PcAssocParams& p, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) { generic_launch_isochoric<PcAssocDev>(p, p.pc.nc, in, o, st); }
}
