// genericSAFT with a PC-SAFT non-polar term: get_ATrhoXi table, small-jet part 1 (tq_generic_dir.cuh)
#include "tq_common.hpp"
#include "tq_generic_dir.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_ATrhoXi_part1(const PcAssocParams& p, const DeviceBatch& in, double* out, cudaStream_t st) { split_launch_ATrhoXi<PcAssocDev, 2,0>(p, p.pc.nc, in, out, st); }
}
