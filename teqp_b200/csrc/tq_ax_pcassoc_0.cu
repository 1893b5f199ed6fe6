// genericSAFT with a PC-SAFT non-polar term: get_ATrhoXi table, small-jet part 0 (tq_generic_dir.cuh)
#include "tq_common.hpp"
#include "tq_generic_dir.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_ATrhoXi_part0(const PcAssocParams& p, const DeviceBatch& in, double* out, cudaStream_t st) { split_launch_ATrhoXi<PcAssocDev, 1,1>(p, p.pc.nc, in, out, st); }
}
