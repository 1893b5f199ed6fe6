// GERG-2004/2008 isochoric family on the device: one state point per thread, per-thread working set in a column of the CTA's
// shared staging area (see tq_gerg_iso.cuh for the mathematics and the reference lines it replaces).
#include "tq_gerg_iso.cuh"
#include "tq_gerg_iso_coop.cuh"

#include <cstdlib>

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

// ---- lane-cooperative kernels (tq_gerg_iso_coop.cuh) ----------------------------------------------------------------------------
namespace {

constexpr int kCoopMaxWarps = 16;   // 512 threads: 128 registers per thread

// persistent CTA: the per-model tables go to shared memory once, then every warp walks its share of the batch, 32 / L points at a time
template<int L, bool WANT_H, bool IP>
__global__ void __launch_bounds__(32 * kCoopMaxWarps) gerg_iso_coop_kernel(const double* __restrict__ blob_g, const int blob_words, const int rows,
                                                                             const double Rgas, const DeviceBatch in, const IsochoricOut o,
                                                                             const IsoPropsOut ip) {
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < blob_words; i += blockDim.x) sm[i] = __ldg(blob_g + i);
    __syncthreads();
    constexpr int PPW = 32 / L;
    const int warp = threadIdx.x >> 5, l32 = threadIdx.x & 31;
    const int grp = l32 / L, lane = l32 % L;
    const int wpc = blockDim.x >> 5;
    double* W = sm + blob_words + (size_t)(warp * PPW + grp) * rows;   // rows = the point stride (gerg_coop_pstride)
    const int nc = reinterpret_cast<const GergIsoHdr*>(sm)->nc;
    const long long stride = (long long)gridDim.x * wpc * PPW;
    for (long long base = ((long long)blockIdx.x * wpc + warp) * PPW; base < in.npts; base += stride) {
        const long long k0 = base + grp;
        const bool valid = k0 < in.npts;
        const long long k = valid ? k0 : in.npts - 1;
        for (int i = lane; i < nc; i += L) W[i] = __ldg(in.x + (long long)i * in.x_stride + k);
        __syncwarp();
        gerg_iso_coop_point<L, WANT_H>(sm, Rgas, __ldg(in.T + k), W, lane, valid, o, k, in.npts, IP ? &ip : nullptr);
        __syncwarp();
    }
}

template<int L, bool WANT_H, bool IP>
bool coop_launch_l(const double* blob_d, const GergIsoHdr& h, double Rgas, const DeviceBatch& in, const IsochoricOut& o, const IsoPropsOut& ip,
                   cudaStream_t st) {
    constexpr int PPW = 32 / L;
    const int rows = gerg_coop_pstride(gerg_coop_rows(h.nc, h.nw, h.nT, h.ngroup, WANT_H), L);
    const size_t blob_bytes = (size_t)h.nwords * 8, per_warp = (size_t)rows * PPW * 8;
    if (blob_bytes + 2 * per_warp + 1024 > kSmemPerSM) return false;
    int nwarps = (int)std::min<size_t>(kCoopMaxWarps, (kSmemPerSM - 1024 - blob_bytes) / per_warp);
    if (const char* e = std::getenv("TEQPB200_ISO_COOP_WARPS")) nwarps = std::max(1, std::min(nwarps, std::atoi(e)));
    const size_t smem = blob_bytes + per_warp * nwarps;
    auto kern = gerg_iso_coop_kernel<L, WANT_H, IP>;
    if (smem > 48 * 1024) TQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long per_cta = (long long)nwarps * PPW;
    const int grid = (int)std::min<long long>(sm_count(), (in.npts + per_cta - 1) / per_cta);
    kern<<<grid, 32 * nwarps, smem, st>>>(blob_d, h.nwords, rows, Rgas, in, o, ip);
    TQ_CUDA(cudaGetLastError());
    return true;
}

template<bool WANT_H, bool IP>
bool coop_launch(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o,
                 const IsoPropsOut& ip, cudaStream_t st) {
    static const int L = [] {
        const char* e = std::getenv("TEQPB200_ISO_COOP_L");
        const int v = e ? std::atoi(e) : 8;
        return (v == 0 || v == 4 || v == 16) ? v : 8;
    }();
    if (L == 0) return false;   // TEQPB200_ISO_COOP_L=0: the thread-per-point kernels (A/B measurements)
    int dev = 0;
    TQ_CUDA(cudaGetDevice(&dev));
    const double* blob_d = tab.on_device(P, dev, L);
    const GergIsoHdr& h = *reinterpret_cast<const GergIsoHdr*>(tab.host.data());
    if (L == 4) return coop_launch_l<4, WANT_H, IP>(blob_d, h, Rgas, in, o, ip, st);
    if (L == 16) return coop_launch_l<16, WANT_H, IP>(blob_d, h, Rgas, in, o, ip, st);
    return coop_launch_l<8, WANT_H, IP>(blob_d, h, Rgas, in, o, ip, st);
}

}  // namespace

const double* GergIsoDeviceTables::on_device(const GergProgram& P, int dev, int L) const {
    std::lock_guard<std::mutex> g(mtx_);
    if (host.empty()) const_cast<std::vector<double>&>(host) = gerg_iso_build_blob(P, L);
    if (reinterpret_cast<const GergIsoHdr*>(host.data())->L != L) throw InvalidArgument("lane-cooperative tables were built for another number of lanes per point");
    auto it = dev_.find(dev);
    if (it != dev_.end()) return it->second;
    double* d = nullptr;
    TQ_CUDA(cudaMalloc(&d, host.size() * sizeof(double)));
    TQ_CUDA(cudaMemcpy(d, host.data(), host.size() * sizeof(double), cudaMemcpyHostToDevice));
    dev_[dev] = d;
    return d;
}
GergIsoDeviceTables::~GergIsoDeviceTables() {
    for (auto& kv : dev_) cudaFree(kv.second);
}

bool gerg_coop_launch_isochoric(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o,
                                cudaStream_t st) {
    if (in.npts == 0) return true;
    const IsoPropsOut none{};
    return o.hess ? coop_launch<true, false>(tab, P, Rgas, in, o, none, st) : coop_launch<false, false>(tab, P, Rgas, in, o, none, st);
}
bool gerg_coop_launch_isoprops(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsoPropsOut& ip,
                               cudaStream_t st) {
    if (in.npts == 0) return true;
    const IsochoricOut none{};
    return coop_launch<true, true>(tab, P, Rgas, in, none, ip, st);
}

bool gerg_coop_launch_fused(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o, const IsoPropsOut& ip,
                            cudaStream_t st) {
    if (in.npts == 0) return true;
    return coop_launch<true, true>(tab, P, Rgas, in, o, ip, st);
}

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
