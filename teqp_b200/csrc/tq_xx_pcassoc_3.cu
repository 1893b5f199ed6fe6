// genericSAFT with a PC-SAFT non-polar term: get_ATrhoXiXjXk by directional jets
#include "tq_common.hpp"
#include "tq_generic_dir.cuh"
#include "tq_pcassoc.cuh"
namespace tq {
void pcassoc_launch_cross3(const PcAssocParams& p, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) { dir_launch_cross<PcAssocDev, 3>(p, p.pc.nc, in, i, j, k, out, st); }
}
