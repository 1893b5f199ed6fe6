// GERG-2004/2008 isochoric family on the device: one state point per thread, per-thread working set in a column of the CTA's
// shared staging area (see tq_gerg_iso.cuh for the mathematics and the reference lines it replaces).
#include "tq_gerg_iso.cuh"

namespace tq {

namespace {

template<int WS>
__global__ void __launch_bounds__(WS) gerg_isoprops_kernel(const __grid_constant__ GergProgram P, const double Rgas, const DeviceBatch in,
                                                           const IsoPropsOut ip) {
    extern __shared__ double Wsh[];
    double* W = Wsh + threadIdx.x;
    const long long k = (long long)blockIdx.x * WS + threadIdx.x;
    if (k >= in.npts) return;
    Exp2K K = Exp2K::load(in.npts >> 62);
    K.load_thread((long long)(((unsigned)__double2hiint(__ldg(in.T + k)) >> 31) << 3));
    for (int i = 0; i < P.nc; ++i) W[i * WS] = __ldg(in.x + (long long)i * in.x_stride + k);
    const IsochoricOut none{};
    gerg_iso_point<WS, true>(P, K, Rgas, __ldg(in.T + k), W, none, k, in.npts, &ip);
}

template<int WS, bool WANT_H>
__global__ void __launch_bounds__(WS) gerg_isochoric_kernel(const __grid_constant__ GergProgram P, const double Rgas, const DeviceBatch in,
                                                            const IsochoricOut o) {
    extern __shared__ double Wsh[];  // [rows][WS]; a thread touches only its own column -> no barriers
    double* W = Wsh + threadIdx.x;
    const long long k = (long long)blockIdx.x * WS + threadIdx.x;
    if (k >= in.npts) return;
    Exp2K K = Exp2K::load(in.npts >> 62);
    K.load_thread((long long)(((unsigned)__double2hiint(__ldg(in.T + k)) >> 31) << 3));
    for (int i = 0; i < P.nc; ++i) W[i * WS] = __ldg(in.x + (long long)i * in.x_stride + k);
    gerg_iso_point<WS, WANT_H>(P, K, Rgas, __ldg(in.T + k), W, o, k, in.npts);
}

constexpr size_t kSmemPerSM = 227 * 1024;

template<int WS, bool WANT_H>
void launch_ws(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st, size_t smem) {
    auto kern = gerg_isochoric_kernel<WS, WANT_H>;
    if (smem > 48 * 1024) TQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = grid_for(in.npts, WS, 1 << 30);
    kern<<<grid, WS, smem, st>>>(P, Rgas, in, o);
    TQ_CUDA(cudaGetLastError());
}

template<bool WANT_H>
void launch_h(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) {
    const size_t rows = (size_t)gerg_iso_rows(P.nc, P.nw, WANT_H);
    // CTA size: the one that keeps the most warps resident per SM given the staging area (ties: the smaller CTA)
    int best = 0, best_warps = -1;
    for (int ws : {32, 64, 128}) {
        const size_t smem = rows * ws * sizeof(double) + 1024;  // + per-CTA reservation
        if (smem > kSmemPerSM) continue;
        const int ctas = (int)std::min<size_t>(32, kSmemPerSM / smem);
        const int warps = std::min(64, ctas * (ws / 32));
        if (warps > best_warps) { best_warps = warps; best = ws; }
    }
    if (!best) throw InvalidArgument("mixture too large for the isochoric kernel's staging area");
    const size_t smem = rows * best * sizeof(double);
    if (best == 32) launch_ws<32, WANT_H>(P, Rgas, in, o, st, smem);
    else if (best == 64) launch_ws<64, WANT_H>(P, Rgas, in, o, st, smem);
    else launch_ws<128, WANT_H>(P, Rgas, in, o, st, smem);
}

template<int WS>
void launch_ip(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsoPropsOut& ip, cudaStream_t st, size_t smem) {
    auto kern = gerg_isoprops_kernel<WS>;
    if (smem > 48 * 1024) TQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid_for(in.npts, WS, 1 << 30), WS, smem, st>>>(P, Rgas, in, ip);
    TQ_CUDA(cudaGetLastError());
}

}  // namespace

void gerg_launch_isoprops(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsoPropsOut& ip, cudaStream_t st) {
    if (in.npts == 0) return;
    const size_t rows = (size_t)gerg_iso_rows(P.nc, P.nw, true);
    const size_t smem = rows * 32 * sizeof(double);
    if (smem + 1024 > kSmemPerSM) throw InvalidArgument("mixture too large for the isochoric kernel's staging area");
    launch_ip<32>(P, Rgas, in, ip, st, smem);
}

void gerg_launch_isochoric(const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) {
    if (in.npts == 0) return;
    if (o.hess) launch_h<true>(P, Rgas, in, o, st);
    else launch_h<false>(P, Rgas, in, o, st);
}

}  // namespace tq
