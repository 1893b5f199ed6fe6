// genericSAFT with a PC-SAFT non-polar term: derived isochoric properties, single-jet-shape part X
#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_ip_part2(const PcAssocParams& p, const DeviceBatch& in, const IsoPropsOut& o, double* Ls, const double* H, cudaStream_t st) { generic_launch_ip_part<PcAssocDev, 2>(p, p.pc.nc, in, o, Ls, H, st); }
}
