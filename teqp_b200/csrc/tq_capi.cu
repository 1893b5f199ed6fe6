// C ABI of libteqpb200 (declared in include/teqpb200.h): the reference's extern "C" interface
// (/root/reference/interface/C/teqpc.cpp:41-225) re-implemented over the GPU kernels, plus the batched entry points and the
// host-side driver that shards a host-resident batch over the configured devices and streams it through each of them.
#include <array>
#include <atomic>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "../../include/teqpb200.h"
#include "tq_vle.cuh"
#include "tq_common.hpp"

namespace tq {

int sm_count() {
    static int cache[64] = {0};
    int dev = 0;
    TQ_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && cache[dev]) return cache[dev];
    int n = 0;
    TQ_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    if (dev < 64) cache[dev] = n;
    return n;
}

bool gerg_try_launch_fused(const Model&, const DeviceBatch&, const IsochoricOut&, const IsoPropsOut&, cudaStream_t);   // tq_gerg.cu
std::unique_ptr<Model> make_qcpr_model(const Json& spec);   // tq_qcpr.cu
std::unique_ptr<Model> make_rkpr_model(const Json& spec);   // tq_rkpr.cu
std::unique_ptr<Model> make_apr_model(const Json& spec);    // tq_apr.cu
// ideal-gas kinds (tq_ideal.cu)
std::unique_ptr<Model> make_ideal_helmholtz_model(const Json& jpures);
std::unique_ptr<Model> make_gerg_idealgas_model(const std::string& kind, const Json& spec);

std::unique_ptr<Model> make_model(const Json& j) {
    // kind -> factory; the in-scope subset of /root/reference/interface/CPP/teqp_impl_factory.cpp:54-102
    const std::string kind = j.at("kind").as_string();
    const Json& spec = j.at("model");
    if (kind == "vdW1" || kind == "vdW") return make_vdw_model(kind, spec);
    if (kind == "PR" || kind == "SRK" || kind == "cubic") return make_cubic_model(kind, spec);
    if (kind == "GERG2008resid" || kind == "GERG2004resid") return make_gerg_model(kind, spec);
    if (kind == "PCSAFT") return make_pcsaft_model(spec);
    if (kind == "SAFT-VR-Mie") return make_saftvrmie_model(spec);
    if (kind == "genericSAFT") return make_genericsaft_model(spec);
    if (kind == "multifluid") return make_multifluid_model(spec);
    if (kind == "QCPRAasen") return make_qcpr_model(spec);
    if (kind == "RKPRCismondi2005") return make_rkpr_model(spec);
    if (kind == "advancedPRaEres") return make_apr_model(spec);
    if (kind == "IdealHelmholtz") return make_ideal_helmholtz_model(spec);
    if (kind == "GERG2008idealgas" || kind == "GERG2004idealgas") return make_gerg_idealgas_model(kind, spec);
    throw InvalidArgument("Don't understand \"kind\" of: " + kind);
}

}  // namespace tq

using namespace tq;

namespace {

std::mutex g_lib_mtx;
std::unordered_map<long long, std::shared_ptr<Model>> g_library;
std::atomic<long long> g_next_index{0};
std::atomic<long long> g_launches{0};
std::mutex g_dev_mtx;
std::vector<int> g_devices;  // empty: the calling thread's current device
thread_local cudaStream_t tl_stream = nullptr;

struct CodeError : std::exception {  // the reference's teqpcException (teqpc.cpp:33-39)
    int code; std::string msg;
    CodeError(int c, std::string m) : code(c), msg(std::move(m)) {}
    const char* what() const noexcept override { return msg.c_str(); }
};

void write_error(const std::string& msg, char* buf, int len) {
    // same truncation rules as exception_handler in teqpc.cpp:45-73
    if (!buf || len <= 0) return;
    if (msg.size() < (size_t)len) { std::strcpy(buf, msg.c_str()); return; }
    const std::string toolong = "Error message too long for buffer";
    if (toolong.size() < (size_t)len) std::strcpy(buf, toolong.c_str());
    else if (len > 1) std::strcpy(buf, "?");
}

// reference entry points collapse every std::exception to 9999; the new batched ones keep the numeric teqp code
template<class F> int guarded(bool legacy, char* errmsg, int len, F&& f) {
    try { f(); return 0; }
    catch (const CodeError& e) { write_error(e.msg, errmsg, len); return e.code; }
    catch (const Error& e) { write_error(e.msg, errmsg, len); return legacy ? 9999 : e.code; }
    catch (const std::exception& e) { write_error(e.what(), errmsg, len); return 9999; }
    catch (...) { write_error("unknown error", errmsg, len); return 9999; }
}

std::shared_ptr<Model> lookup(long long uuid) {
    std::lock_guard<std::mutex> g(g_lib_mtx);
    auto it = g_library.find(uuid);
    if (it == g_library.end()) throw std::out_of_range("unordered_map::at");  // what .at() does in the reference (teqpc.cpp:111)
    return it->second;
}

void check_ncomp(const Model& m, int Ncomp) {
    if (Ncomp != m.ncomp()) throw InvalidArgument("mole fractions must be of size " + std::to_string(m.ncomp()) + " but are of size " + std::to_string(Ncomp));
}

// ---------------------------------------------------------------------------------------------------------------------
// host-resident batches: shard over devices, stream through each device in chunks
// ---------------------------------------------------------------------------------------------------------------------
struct HostIn { const double* p; int rows; bool broadcast; };
struct HostOut { double* p; int rows; };  // p == nullptr: device-only scratch rows
using ChunkFn = std::function<void(long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st)>;

constexpr int kSlots = 3;
constexpr long long kMaxChunk = 1LL << 22;
constexpr long long kMinChunk = 1LL << 16;
constexpr size_t kMaxSlotBytes = 384ull << 20;  // scratch per (device, slot): wide outputs (GERG21 tables, ~800 rows) shrink the chunk instead

struct DeviceArena {  // grow-only device scratch, one per (device, slot); released by teqpb200_release_scratch()
    double* p = nullptr; size_t cap = 0; cudaStream_t st = nullptr;
};
struct DeviceState {  // one per device: concurrent host-pointer calls serialise per device, not process-wide
    std::mutex mtx;
    std::array<DeviceArena, kSlots> arenas;
};
std::mutex g_state_mtx;
std::unordered_map<int, std::unique_ptr<DeviceState>> g_states;
std::atomic<long long> g_fault_chunk{-1};   // test hook (teqpb200_inject_fault): the chunk with this running index fails
std::atomic<long long> g_chunk_counter{0};
// device-side time of the last host-pointer call, per device: first H2D enqueued -> last D2H finished (CUDA events on the arena streams)
std::mutex g_time_mtx;
std::unordered_map<int, double> g_last_ms;
DeviceState& device_state(int dev) {
    std::lock_guard<std::mutex> g(g_state_mtx);
    auto& p = g_states[dev];
    if (!p) p = std::make_unique<DeviceState>();
    return *p;
}

// rows are contiguous n-double spans at p + r*npts + offset: one plain async copy per row (cudaMemcpy2DAsync rejects host pitches
// above cudaDevAttrMaxPitch = 2^31-1 bytes, i.e. batches above 2.68e8 points)
void copy_rows_h2d(double* dst, long long dst_stride, const double* src, long long src_stride, long long n, int rows, cudaStream_t st) {
    for (int r = 0; r < rows; ++r)
        TQ_CUDA(cudaMemcpyAsync(dst + (size_t)r * dst_stride, src + (size_t)r * src_stride, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, st));
}
void copy_rows_d2h(double* dst, long long dst_stride, const double* src, long long src_stride, long long n, int rows, cudaStream_t st) {
    for (int r = 0; r < rows; ++r)
        TQ_CUDA(cudaMemcpyAsync(dst + (size_t)r * dst_stride, src + (size_t)r * src_stride, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, st));
}

void run_on_device(int dev, long long k0, long long n_dev, long long npts, const std::vector<HostIn>& ins,
                   const std::vector<HostOut>& outs, const ChunkFn& fn) {
    if (n_dev <= 0) return;
    TQ_CUDA(cudaSetDevice(dev));
    DeviceState& S = device_state(dev);
    std::lock_guard<std::mutex> lock(S.mtx);
    auto& arenas = S.arenas;
    size_t rows_chunk = 0, fixed = 0;   // rows that scale with the chunk; doubles that do not (broadcast inputs)
    for (auto& a : ins) { if (a.broadcast) fixed += a.rows; else rows_chunk += a.rows; }
    for (auto& a : outs) rows_chunk += a.rows;
    long long chunk = std::min(kMaxChunk, std::max(kMinChunk, (n_dev + 7) / 8));
    const long long by_bytes = (long long)(kMaxSlotBytes / (sizeof(double) * std::max<size_t>(1, rows_chunk)));
    chunk = std::min(chunk, std::max<long long>(4096, by_bytes & ~127LL));
    chunk = std::min(chunk, n_dev);
    const size_t need = (rows_chunk * (size_t)chunk + fixed) * sizeof(double);
    for (int s = 0; s < kSlots; ++s) {
        DeviceArena& A = arenas[s];
        if (!A.st) TQ_CUDA(cudaStreamCreateWithFlags(&A.st, cudaStreamNonBlocking));
        if (A.cap < need) {
            if (A.p) { TQ_CUDA(cudaStreamSynchronize(A.st)); TQ_CUDA(cudaFree(A.p)); A.p = nullptr; A.cap = 0; }
            TQ_CUDA(cudaMalloc(&A.p, need));
            A.cap = need;
        }
    }
    cudaEvent_t ev0 = nullptr, ev1[kSlots] = {nullptr, nullptr, nullptr};
    TQ_CUDA(cudaEventCreate(&ev0));
    for (int s = 0; s < kSlots; ++s) TQ_CUDA(cudaEventCreate(&ev1[s]));
    auto drop_events = [&] { cudaEventDestroy(ev0); for (int s = 0; s < kSlots; ++s) cudaEventDestroy(ev1[s]); };
    try {
        int slot = 0;
        TQ_CUDA(cudaEventRecord(ev0, arenas[0].st));
        for (long long c0 = 0; c0 < n_dev; c0 += chunk, slot = (slot + 1) % kSlots) {
            const long long n = std::min(chunk, n_dev - c0);
            DeviceArena& A = arenas[slot];
            cudaStream_t st = A.st;
            double* cur = A.p;
            if (g_chunk_counter++ == g_fault_chunk.load()) { g_fault_chunk = -1; throw DeviceError("injected fault (teqpb200_inject_fault)"); }
            std::vector<const double*> din; std::vector<double*> dout;
            for (auto& a : ins) {
                din.push_back(cur);
                if (a.broadcast) {
                    TQ_CUDA(cudaMemcpyAsync(cur, a.p, sizeof(double) * a.rows, cudaMemcpyHostToDevice, st));
                    cur += a.rows;
                } else {
                    copy_rows_h2d(cur, n, a.p + k0 + c0, npts, n, a.rows, st);
                    cur += (size_t)a.rows * chunk;
                }
            }
            for (auto& a : outs) { dout.push_back(cur); cur += (size_t)a.rows * chunk; }
            fn(n, din, dout, st);
            for (size_t q = 0; q < outs.size(); ++q) {
                if (!outs[q].p) continue;
                copy_rows_d2h(outs[q].p + k0 + c0, npts, dout[q], n, n, outs[q].rows, st);
            }
        }
    } catch (...) {
        // copies into the caller's buffers may still be queued: drain them before the error leaves the library
        for (int s = 0; s < kSlots; ++s) if (arenas[s].st) cudaStreamSynchronize(arenas[s].st);
        cudaGetLastError();
        drop_events();
        throw;
    }
    for (int s = 0; s < kSlots; ++s) TQ_CUDA(cudaEventRecord(ev1[s], arenas[s].st));
    for (int s = 0; s < kSlots; ++s) TQ_CUDA(cudaStreamSynchronize(arenas[s].st));
    float ms = 0.f, m = 0.f;
    for (int s = 0; s < kSlots; ++s) if (cudaEventElapsedTime(&m, ev0, ev1[s]) == cudaSuccess) ms = std::max(ms, m);
    { std::lock_guard<std::mutex> g(g_time_mtx); g_last_ms[dev] = ms; }
    drop_events();
}

void run_host(long long npts, const std::vector<HostIn>& ins, const std::vector<HostOut>& outs, const ChunkFn& fn) {
    if (npts <= 0) return;
    std::vector<int> devs;
    { std::lock_guard<std::mutex> gd(g_dev_mtx); devs = g_devices; }
    int cur = 0;
    TQ_CUDA(cudaGetDevice(&cur));
    if (devs.empty() || npts < (long long)devs.size() * 1024) devs = {devs.empty() ? cur : devs[0]};
    const int nd = (int)devs.size();
    if (nd == 1) {
        try { run_on_device(devs[0], 0, npts, npts, ins, outs, fn); }
        catch (...) { cudaSetDevice(cur); throw; }
        TQ_CUDA(cudaSetDevice(cur));
        return;
    }
    // contiguous even split, remainder to the last device; one host thread per device
    const long long per = npts / nd;
    std::vector<std::thread> th;
    std::vector<std::string> errs(nd);
    std::vector<int> codes(nd, 0);
    for (int d = 0; d < nd; ++d) {
        const long long k0 = per * d, n = (d == nd - 1) ? npts - k0 : per;
        th.emplace_back([&, d, k0, n] {
            try { run_on_device(devs[d], k0, n, npts, ins, outs, fn); }
            catch (const Error& e) { codes[d] = e.code; errs[d] = e.msg; }
            catch (const std::exception& e) { codes[d] = 9999; errs[d] = e.what(); }
        });
    }
    for (auto& t : th) t.join();   // every device has drained its streams (run_on_device syncs on both exits) before an error is reported
    TQ_CUDA(cudaSetDevice(cur));
    for (int d = 0; d < nd; ++d) if (codes[d]) throw Error(codes[d], "device " + std::to_string(devs[d]) + ": " + errs[d]);
}

// Device-pointer calls take their scratch rows from the device's default stream-ordered memory pool (cudaMallocAsync).  With the pool's default release
// threshold of zero every synchronisation hands the freed blocks back to the driver, and the next call pays for mapping several hundred MB again
// (measured: 4 - 40 ms per call of pure_VLE_T_batch on 2 10^6 isotherms, against 1.5 ms of kernels).  The library therefore lets the pool keep up to
// 8 GiB cached between calls (once per device); teqpb200_release_scratch() trims it.
void scratch_pool_setup() {
    static std::mutex mtx;
    static bool done[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
    std::lock_guard<std::mutex> g(mtx);
    if (done[dev]) return;
    cudaMemPool_t pool = nullptr;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess && pool) {
        unsigned long long thr = 8ull << 30;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    cudaGetLastError();
    done[dev] = true;
}

// one entry point for both pointer kinds
void run(int flags, long long npts, const std::vector<HostIn>& ins, const std::vector<HostOut>& outs, const ChunkFn& fn) {
    if (npts < 0) throw InvalidArgument("npts must be non-negative");
    if (npts == 0) return;
    if (flags & TEQPB200_DEVICE_PTRS) {
        scratch_pool_setup();
        std::vector<const double*> din; std::vector<double*> dout;
        std::vector<double*> scratch;
        for (auto& a : ins) din.push_back(a.p);
        for (auto& a : outs) {
            if (a.p) { dout.push_back(a.p); continue; }
            double* s = nullptr;
            TQ_CUDA(cudaMallocAsync(&s, sizeof(double) * (size_t)a.rows * npts, tl_stream));
            scratch.push_back(s); dout.push_back(s);
        }
        fn(npts, din, dout, tl_stream);
        for (double* s : scratch) TQ_CUDA(cudaFreeAsync(s, tl_stream));
    } else {
        run_host(npts, ins, outs, fn);
    }
}

DeviceBatch make_batch(long long n, const double* T, const double* rho, const double* x, bool broadcast) {
    DeviceBatch b;
    b.npts = n; b.T = T; b.rho = rho; b.x = x; b.x_stride = broadcast ? 1 : n; b.x_bcast = broadcast ? 1 : 0;
    return b;
}

void ar_table_call(const Model& m, int NT, int ND, long long npts, const double* T, const double* rho, const double* x, int Ncomp,
                   double* out, int flags) {
    check_ncomp(m, Ncomp);
    const bool bc = flags & TEQPB200_BROADCAST_X;
    run(flags, npts, {{T, 1, false}, {rho, 1, false}, {x, Ncomp, bc}}, {{out, (NT + 1) * (ND + 1)}},
        [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
            m.launch_ar_table(NT, ND, make_batch(n, din[0], din[1], din[2], bc), dout[0], st);
            ++g_launches;
        });
}


// per-point status (SURVEY 8b "per-point int* status optional"): classifies each point of a finished batched call from its inputs and outputs
template<class OutT>
__global__ void __launch_bounds__(256) point_status_kernel(const DeviceBatch in, const int nc, const double* __restrict__ out, const int nrows,
                                                            OutT* __restrict__ status) {
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        int s = 0;
        const double T = __ldg(in.T + k), rho = __ldg(in.rho + k);
        if (!(T > 0) || !isfinite(T) || !(rho >= 0) || !isfinite(rho)) s |= TEQPB200_STATUS_BAD_INPUT;
        double sum = 0.0;
        for (int i = 0; i < nc; ++i) {
            const double x = __ldg(in.x + (long long)i * in.x_stride + (in.x_bcast ? 0 : k));
            if (!(x >= 0) || !isfinite(x)) s |= TEQPB200_STATUS_BAD_INPUT;
            sum += x;
        }
        if (!(fabs(sum - 1.0) <= 1e-6)) s |= TEQPB200_STATUS_COMPOSITION_SUM;
        for (int r = 0; r < nrows; ++r)
            if (!isfinite(__ldg(out + (long long)r * npts + k))) s |= TEQPB200_STATUS_NONFINITE_OUTPUT;
        status[k] = (OutT)s;
    }
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------------------------------------------------------------
// Part 1: reference ABI
// ---------------------------------------------------------------------------------------------------------------------
int build_model(const char* j, long long int* uuid, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        Json json = Json::parse(j);
        const long long uid = g_next_index++;
        std::shared_ptr<Model> m;
        try { m = tq::make_model(json); }
        catch (const std::exception& e) { throw CodeError(30, "Unable to load with error:" + std::string(e.what())); }
        { std::lock_guard<std::mutex> g(g_lib_mtx); g_library.emplace(uid, std::move(m)); }
        *uuid = uid;
    });
}

int free_model(const long long int uuid, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] { std::lock_guard<std::mutex> g(g_lib_mtx); g_library.erase(uuid); });
}

int get_Arxy(const long long int uuid, const int NT, const int ND, const double T, const double rho, const double* molefrac,
             const int Ncomp, double* val, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        // runtime dispatch of the reference accepts only these (derivs.hpp:627-682)
        if (NT < 0 || ND < 0 || NT + ND > 3) throw std::invalid_argument("Not valid: NT=" + std::to_string(NT) + ", ND=" + std::to_string(ND));
        std::vector<double> table((NT + 1) * (ND + 1));
        ar_table_call(*m, NT, ND, 1, &T, &rho, molefrac, Ncomp, table.data(), 0);
        *val = table[NT * (ND + 1) + ND];
    });
}

int get_ATrhoXi(const long long int uuid, const double T, const int NT, const double rhomolar, const int ND, const double* molefrac,
                const int Ncomp, const int i, const int NXi, double* val, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        // combinations the reference compiles (derivs.hpp:261-284): NT+ND <= 2, NXi <= 3, not all zero
        if (NT < 0 || ND < 0 || NXi < 0 || NT + ND > 2 || NXi > 3 || (NT + ND + NXi) == 0) throw std::invalid_argument("Can't match these derivative counts");
        if (i < 0 || i >= Ncomp) throw std::invalid_argument("component index out of range");
        std::vector<double> table((size_t)Ncomp * 36);
        run(0, 1, {{&T, 1, false}, {&rhomolar, 1, false}, {molefrac, Ncomp, false}}, {{table.data(), Ncomp * 36}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_ATrhoXi_table(make_batch(n, din[0], din[1], din[2], false), dout[0], st);
                ++g_launches;
            });
        *val = table[((i * 3 + NT) * 3 + ND) * 4 + NXi];
    });
}

#define TQ_NOT_YET(name)                                                                                              \
    return guarded(true, errmsg, errmsg_length, [&] {                                                                 \
        (void)lookup(uuid);                                                                                           \
        throw NotImplementedError(name " is not implemented in libteqpb200 yet");                                    \
    })

int get_ATrhoXiXj(const long long int uuid, const double T, const int NT, const double rhomolar, const int ND, const double* molefrac,
                  const int Ncomp, const int i, const int NXi, const int j, const int NXj, double* val, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (i == j) throw std::invalid_argument("i cannot equal j");  // derivs.hpp:425-427
        // combinations the reference compiles (derivs.hpp:294-309): NT+ND <= 1, (NXi, NXj) in {(1,0),(2,0),(0,1),(0,2),(1,1)}
        const bool ok = NT >= 0 && ND >= 0 && NT + ND <= 1 && NXi >= 0 && NXj >= 0 && NXi + NXj >= 1 && NXi + NXj <= 2;
        if (!ok) throw std::invalid_argument("Can't match these derivative counts");
        if (i < 0 || i >= Ncomp || j < 0 || j >= Ncomp) throw std::invalid_argument("component index out of range");
        double table[36];
        run(0, 1, {{&T, 1, false}, {&rhomolar, 1, false}, {molefrac, Ncomp, false}}, {{table, 36}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_ATrhoXiXj_table(make_batch(n, din[0], din[1], din[2], false), i, j, dout[0], st);
                ++g_launches;
            });
        *val = table[((NT * 2 + ND) * 3 + NXi) * 3 + NXj];
    });
}
int get_ATrhoXiXjXk(const long long int uuid, const double T, const int NT, const double rhomolar, const int ND, const double* molefrac,
                    const int Ncomp, const int i, const int NXi, const int j, const int NXj, const int k, const int NXk, double* val,
                    char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (i == j || j == k || i == k) throw std::invalid_argument("i, j, and k must all be unique");  // derivs.hpp:449-451
        // derivs.hpp:318-330: NT+ND <= 1, each composition order 0 or 1, at least two of them 1
        const bool ok = NT >= 0 && ND >= 0 && NT + ND <= 1 && NXi >= 0 && NXi <= 1 && NXj >= 0 && NXj <= 1 && NXk >= 0 && NXk <= 1 && NXi + NXj + NXk >= 2;
        if (!ok) throw std::invalid_argument("Can't match these derivative counts");
        if (i < 0 || i >= Ncomp || j < 0 || j >= Ncomp || k < 0 || k >= Ncomp) throw std::invalid_argument("component index out of range");
        double table[32];
        run(0, 1, {{&T, 1, false}, {&rhomolar, 1, false}, {molefrac, Ncomp, false}}, {{table, 32}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_ATrhoXiXjXk_table(make_batch(n, din[0], din[1], din[2], false), i, j, k, dout[0], st);
                ++g_launches;
            });
        *val = table[(((NT * 2 + ND) * 2 + NXi) * 2 + NXj) * 2 + NXk];
    });
}
// get_AtaudeltaXi / XiXj / XiXjXk (teqpc.h:14-18; derivs.hpp:341-414): as the three calls above with (tau, delta) given instead of (T, rho).
// Only models with alphar_taudelta (the JSON-driven multifluid kind) serve them; for the others the reference throws as well (derivs.hpp:137-140).
// The kernels are the same composition-derivative sweeps, driven with T := 1/tau and rho := delta (GergTauDeltaDev, tq_gerg_jet.cuh).
static void ataudelta_table(const Model& m, int which, int i, int j, int k, long long npts, const double* tau, const double* delta, const double* x,
                            int Ncomp, int rows, double* out, int flags) {
    check_ncomp(m, Ncomp);
    if (flags & TEQPB200_DEVICE_PTRS) throw NotImplementedError("get_Ataudelta tables take host pointers");
    const bool bc = flags & TEQPB200_BROADCAST_X;
    std::vector<double> Tinv((size_t)npts);
    for (long long q = 0; q < npts; ++q) Tinv[q] = 1.0 / tau[q];
    run(flags, npts, {{Tinv.data(), 1, false}, {delta, 1, false}, {x, Ncomp, bc}}, {{out, rows}},
        [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
            m.launch_Ataudelta_table(which, make_batch(n, din[0], din[1], din[2], bc), i, j, k, dout[0], st);
            ++g_launches;
        });
}
int get_AtaudeltaXi(const long long int uuid, const double tau, const int NT, const double delta, const int ND, const double* molefrac, const int Ncomp,
                    const int i, const int NXi, double* val, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (!(NT >= 0 && ND >= 0 && NT + ND <= 2 && NXi >= 0 && NXi <= 3)) throw std::invalid_argument("Can't match these derivative counts");
        if (i < 0 || i >= Ncomp) throw std::invalid_argument("component index out of range");
        std::vector<double> table((size_t)Ncomp * 36);
        ataudelta_table(*m, 1, 0, 0, 0, 1, &tau, &delta, molefrac, Ncomp, Ncomp * 36, table.data(), 0);
        *val = table[((i * 3 + NT) * 3 + ND) * 4 + NXi];
    });
}
int get_AtaudeltaXiXj(const long long int uuid, const double tau, const int NT, const double delta, const int ND, const double* molefrac, const int Ncomp,
                      const int i, const int NXi, const int j, const int NXj, double* val, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (i == j) throw std::invalid_argument("i cannot equal j");
        const bool ok = NT >= 0 && ND >= 0 && NT + ND <= 1 && NXi >= 0 && NXj >= 0 && NXi + NXj >= 1 && NXi + NXj <= 2;
        if (!ok) throw std::invalid_argument("Can't match these derivative counts");
        if (i < 0 || i >= Ncomp || j < 0 || j >= Ncomp) throw std::invalid_argument("component index out of range");
        double table[36];
        ataudelta_table(*m, 2, i, j, 0, 1, &tau, &delta, molefrac, Ncomp, 36, table, 0);
        *val = table[((NT * 2 + ND) * 3 + NXi) * 3 + NXj];
    });
}
int get_AtaudeltaXiXjXk(const long long int uuid, const double tau, const int NT, const double delta, const int ND, const double* molefrac, const int Ncomp,
                        const int i, const int NXi, const int j, const int NXj, const int k, const int NXk, double* val, char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (i == j || j == k || i == k) throw std::invalid_argument("i, j, and k must all be unique");
        const bool ok = NT >= 0 && ND >= 0 && NT + ND <= 1 && NXi >= 0 && NXi <= 1 && NXj >= 0 && NXj <= 1 && NXk >= 0 && NXk <= 1 && NXi + NXj + NXk >= 2;
        if (!ok) throw std::invalid_argument("Can't match these derivative counts");
        if (i < 0 || i >= Ncomp || j < 0 || j >= Ncomp || k < 0 || k >= Ncomp) throw std::invalid_argument("component index out of range");
        double table[32];
        ataudelta_table(*m, 3, i, j, k, 1, &tau, &delta, molefrac, Ncomp, 32, table, 0);
        *val = table[(((NT * 2 + ND) * 2 + NXi) * 2 + NXj) * 2 + NXk];
    });
}
// VirialDerivatives::get_dmBnvirdTm_runtime (derivs.hpp:808-841): d^m B_n / dT^m = 1/(n-2)! d^(n-1+m) alphar / dT^m drho^(n-1) at rho = 0.
// One unit-seeded jet sweep at rho = 0 gives D_k = (1/T)^k d^k/d(1/T)^k of d^(n-1) alphar / drho^(n-1); with u = 1/T:
//   dB/dT = -D_1 / T,   d2B/dT2 = (D_2 + 2 D_1) / T^2
static void virial_call(const Model& m, int Nvir, int NTd, long long npts, const double* T, const double* x, int Ncomp, double* out, int flags) {
    check_ncomp(m, Ncomp);
    if (Nvir < 2 || Nvir > 7) throw std::invalid_argument("Nderiv is invalid in get_dmBnvirdTm_runtime");
    if (NTd < 0 || NTd > 3) throw std::invalid_argument("NTderiv is invalid in get_dmBnvirdTm_runtime");
    if (NTd == 3 || (NTd > 0 && Nvir > 5)) throw NotImplementedError("get_dmBnvirdTm: third temperature derivatives (and T-derivatives of B_6, B_7) need a jet shape libteqpb200 does not compile");
    const int ND = Nvir - 1;
    const bool bc = flags & TEQPB200_BROADCAST_X;
    double fact = 1.0;
    for (int k = 2; k <= Nvir - 2; ++k) fact *= k;
    const int rows = (NTd + 1) * (ND + 1);
    // rho = 0 everywhere: a device scratch row (device-pointer calls) or a host vector (host calls) of zeros
    std::vector<double> zeros_h;
    double* zeros_d = nullptr;
    const double* rho = nullptr;
    if (flags & TEQPB200_DEVICE_PTRS) {
        scratch_pool_setup();
        TQ_CUDA(cudaMallocAsync(&zeros_d, sizeof(double) * (size_t)npts, tl_stream));
        TQ_CUDA(cudaMemsetAsync(zeros_d, 0, sizeof(double) * (size_t)npts, tl_stream));
        rho = zeros_d;
    } else { zeros_h.assign((size_t)npts, 0.0); rho = zeros_h.data(); }
    std::vector<double> tab_h;
    double* tab = nullptr;
    if (flags & TEQPB200_DEVICE_PTRS) TQ_CUDA(cudaMallocAsync(&tab, sizeof(double) * (size_t)rows * npts, tl_stream));
    else { tab_h.resize((size_t)rows * npts); tab = tab_h.data(); }
    run(flags, npts, {{T, 1, false}, {rho, 1, false}, {x, Ncomp, bc}}, {{tab, rows}},
        [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
            DeviceBatch b = make_batch(n, din[0], din[1], din[2], bc);
            b.rho_unit_seed = 1;
            m.launch_virial_table(NTd, ND, b, dout[0], st);
            ++g_launches;
        });
    if (flags & TEQPB200_DEVICE_PTRS) {
        // combine on the host for simplicity of this (not hot) path: copy the needed rows down, write the result back
        std::vector<double> h((size_t)rows * npts), T_h((size_t)npts), o((size_t)npts);
        TQ_CUDA(cudaMemcpyAsync(h.data(), tab, sizeof(double) * h.size(), cudaMemcpyDeviceToHost, tl_stream));
        TQ_CUDA(cudaMemcpyAsync(T_h.data(), T, sizeof(double) * (size_t)npts, cudaMemcpyDeviceToHost, tl_stream));
        TQ_CUDA(cudaStreamSynchronize(tl_stream));
        for (long long k = 0; k < npts; ++k) {
            const double D0 = h[(size_t)(0 * (ND + 1) + ND) * npts + k];
            const double D1 = NTd >= 1 ? h[(size_t)(1 * (ND + 1) + ND) * npts + k] : 0.0, D2 = NTd >= 2 ? h[(size_t)(2 * (ND + 1) + ND) * npts + k] : 0.0;
            const double Tk = T_h[k];
            o[k] = (NTd == 0 ? D0 : NTd == 1 ? -D1 / Tk : (D2 + 2.0 * D1) / (Tk * Tk)) / fact;
        }
        TQ_CUDA(cudaMemcpyAsync(out, o.data(), sizeof(double) * (size_t)npts, cudaMemcpyHostToDevice, tl_stream));
        TQ_CUDA(cudaStreamSynchronize(tl_stream));
        TQ_CUDA(cudaFreeAsync(tab, tl_stream));
        TQ_CUDA(cudaFreeAsync(zeros_d, tl_stream));
    } else {
        for (long long k = 0; k < npts; ++k) {
            const double D0 = tab[(size_t)(0 * (ND + 1) + ND) * npts + k];
            const double D1 = NTd >= 1 ? tab[(size_t)(1 * (ND + 1) + ND) * npts + k] : 0.0, D2 = NTd >= 2 ? tab[(size_t)(2 * (ND + 1) + ND) * npts + k] : 0.0;
            out[k] = (NTd == 0 ? D0 : NTd == 1 ? -D1 / T[k] : (D2 + 2.0 * D1) / (T[k] * T[k])) / fact;
        }
    }
}

int get_dmBnvirdTm(const long long int uuid, const int Nvir, const int NT, const double T, const double* molefrac, const int Ncomp, double* val,
                   char* errmsg, int errmsg_length) {
    return guarded(true, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        // the reference's runtime dispatch serves B_2..B_4 with 0..3 temperature derivatives (derivs.hpp:808-841)
        if (Nvir < 2 || Nvir > 4) throw std::invalid_argument("Nderiv is invalid in get_dmBnvirdTm_runtime");
        virial_call(*m, Nvir, NT, 1, &T, molefrac, Ncomp, val, 0);
    });
}

// ---------------------------------------------------------------------------------------------------------------------
// Part 2: batched
// ---------------------------------------------------------------------------------------------------------------------
int get_Ar0n_batch(const long long int uuid, const int nmax, const long long int npts, const double* T, const double* rho,
                   const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        if (nmax < 0 || nmax > kMaxDerivD) throw InvalidArgument("nmax must be 0..6");
        ar_table_call(*m, 0, nmax, npts, T, rho, molefrac, Ncomp, out, flags);
    });
}

int get_Arn0_batch(const long long int uuid, const int nmax, const long long int npts, const double* T, const double* rho,
                   const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        if (nmax < 0 || nmax > kMaxDerivT) throw InvalidArgument("nmax must be 0..4");
        ar_table_call(*m, nmax, 0, npts, T, rho, molefrac, Ncomp, out, flags);
    });
}

int get_Arxy_batch(const long long int uuid, const int npairs, const int* NT, const int* ND, const long long int npts,
                   const double* T, const double* rho, const double* molefrac, const int Ncomp, double* out, const int flags,
                   char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (npairs <= 0) throw InvalidArgument("npairs must be positive");
        // The requested pairs are partitioned over the sweeps every family compiles: the density row (NT = 0, ND <= 6), the temperature
        // column (ND = 0, NT <= 4) and the mixed block (NT <= 2, ND <= 4); each needed sweep runs once, sized by the largest order asked
        // of it.  Pairs outside all three (NT >= 3 with ND >= 1, or NT >= 1 with ND >= 5) are rejected -- the reference's own runtime
        // dispatch serves NT + ND <= 3 only (derivs.hpp:627-682), all of which are covered.
        int rowND = -1, colNT = -1, blkNT = -1, blkND = -1;
        std::vector<int> where(npairs);   // 0: row, 1: column, 2: block
        for (int p = 0; p < npairs; ++p) {
            const int nt = NT[p], nd = ND[p];
            if (nt < 0 || nd < 0 || nt > kMaxDerivT || nd > kMaxDerivD) throw InvalidArgument("derivative order out of range");
            if (nt == 0) { where[p] = 0; rowND = std::max(rowND, nd); }
            else if (nd == 0) { where[p] = 1; colNT = std::max(colNT, nt); }
            else if (nt <= 2 && nd <= 4) { where[p] = 2; blkNT = std::max(blkNT, nt); blkND = std::max(blkND, nd); }
            else throw InvalidArgument("mixed derivative (" + std::to_string(nt) + "," + std::to_string(nd) + ") is outside the compiled sweeps: NT = 0 with ND <= 6, ND = 0 with NT <= 4, or NT <= 2 with ND <= 4");
        }
        // when a mixed block is needed anyway it grows to absorb the row / column pairs it can hold (one (2,2) sweep is cheaper than a (1,1),
        // a (0,2) and a (2,0) sweep); what remains outside (ND > 4 on the row, NT > 2 on the column) keeps its own sweep
        if (blkNT >= 0) {
            rowND = -1; colNT = -1;
            for (int p = 0; p < npairs; ++p) {
                if (where[p] == 0 && ND[p] <= 4) { where[p] = 2; blkND = std::max(blkND, ND[p]); }
                else if (where[p] == 1 && NT[p] <= 2) { where[p] = 2; blkNT = std::max(blkNT, NT[p]); }
            }
            for (int p = 0; p < npairs; ++p) {
                if (where[p] == 0) rowND = std::max(rowND, ND[p]);
                if (where[p] == 1) colNT = std::max(colNT, NT[p]);
            }
            // a leftover row / column sweep also serves the lower orders of its own axis: no change needed, they were moved to the block
        }
        const bool bc = flags & TEQPB200_BROADCAST_X;
        const int rows_row = rowND >= 0 ? rowND + 1 : 0, rows_col = colNT >= 0 ? colNT + 1 : 0, rows_blk = blkNT >= 0 ? (blkNT + 1) * (blkND + 1) : 0;
        run(flags, npts, {{T, 1, false}, {rho, 1, false}, {molefrac, Ncomp, bc}}, {{out, npairs}, {nullptr, rows_row + rows_col + rows_blk}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                const DeviceBatch b = make_batch(n, din[0], din[1], din[2], bc);
                double* srow = dout[1];
                double* scol = srow + (size_t)rows_row * n;
                double* sblk = scol + (size_t)rows_col * n;
                if (rows_row) { m->launch_ar_table(0, rowND, b, srow, st); ++g_launches; }
                if (rows_col) { m->launch_ar_table(colNT, 0, b, scol, st); ++g_launches; }
                if (rows_blk) { m->launch_ar_table(blkNT, blkND, b, sblk, st); ++g_launches; }
                for (int p = 0; p < npairs; ++p) {
                    const double* src = where[p] == 0 ? srow + (size_t)ND[p] * n : where[p] == 1 ? scol + (size_t)NT[p] * n
                                                                                                 : sblk + (size_t)(NT[p] * (blkND + 1) + ND[p]) * n;
                    TQ_CUDA(cudaMemcpyAsync(dout[0] + (size_t)p * n, src, sizeof(double) * n, cudaMemcpyDeviceToDevice, st));
                }
            });
    });
}

int get_ATrhoXi_table_batch(const long long int uuid, const long long int npts, const double* T, const double* rho,
                            const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        const bool bc = flags & TEQPB200_BROADCAST_X;
        run(flags, npts, {{T, 1, false}, {rho, 1, false}, {molefrac, Ncomp, bc}}, {{out, Ncomp * 36}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_ATrhoXi_table(make_batch(n, din[0], din[1], din[2], bc), dout[0], st);
                ++g_launches;
            });
    });
}

int get_ATrhoXiXj_table_batch(const long long int uuid, const int i, const int j, const long long int npts, const double* T, const double* rho,
                              const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (i == j) throw InvalidArgument("i cannot equal j");
        if (i < 0 || i >= Ncomp || j < 0 || j >= Ncomp) throw InvalidArgument("component index out of range");
        const bool bc = flags & TEQPB200_BROADCAST_X;
        run(flags, npts, {{T, 1, false}, {rho, 1, false}, {molefrac, Ncomp, bc}}, {{out, 36}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_ATrhoXiXj_table(make_batch(n, din[0], din[1], din[2], bc), i, j, dout[0], st);
                ++g_launches;
            });
    });
}
int get_ATrhoXiXjXk_table_batch(const long long int uuid, const int i, const int j, const int k, const long long int npts, const double* T,
                                const double* rho, const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg,
                                int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (i == j || j == k || i == k) throw InvalidArgument("i, j, and k must all be unique");
        if (i < 0 || i >= Ncomp || j < 0 || j >= Ncomp || k < 0 || k >= Ncomp) throw InvalidArgument("component index out of range");
        const bool bc = flags & TEQPB200_BROADCAST_X;
        run(flags, npts, {{T, 1, false}, {rho, 1, false}, {molefrac, Ncomp, bc}}, {{out, 32}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_ATrhoXiXjXk_table(make_batch(n, din[0], din[1], din[2], bc), i, j, k, dout[0], st);
                ++g_launches;
            });
    });
}

static void isochoric_call(const Model& m, long long npts, const double* T, const double* rhovec, int Ncomp, double* f, double* g,
                           double* H, double* lnphi, double* phi, int flags) {
    check_ncomp(m, Ncomp);
    if (flags & TEQPB200_BROADCAST_X) throw InvalidArgument("TEQPB200_BROADCAST_X does not apply to rhovec inputs");
    std::vector<HostOut> outs;
    std::vector<int> slot(5, -1);
    auto add = [&](double* p, int rows, int which) { if (p) { slot[which] = (int)outs.size(); outs.push_back({p, rows}); } };
    add(f, 1, 0); add(g, Ncomp, 1); add(H, Ncomp * Ncomp, 2); add(lnphi, Ncomp, 3); add(phi, Ncomp, 4);
    if (outs.empty()) return;
    run(flags, npts, {{T, 1, false}, {rhovec, Ncomp, false}}, outs,
        [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
            IsochoricOut o;
            if (slot[0] >= 0) o.Psir = dout[slot[0]];
            if (slot[1] >= 0) o.grad = dout[slot[1]];
            if (slot[2] >= 0) o.hess = dout[slot[2]];
            if (slot[3] >= 0) o.lnphi = dout[slot[3]];
            if (slot[4] >= 0) o.phi = dout[slot[4]];
            m.launch_isochoric(make_batch(n, din[0], nullptr, din[1], false), o, st);
            ++g_launches;
        });
}

int isochoric_properties_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec, const int Ncomp,
                               double* pr, double* splus, double* dpdT, double* d2PsirdTdrhoi, double* dpdrhovec, double* pmv,
                               double* chempotVLE, double* dchempotdT, double* dlnphi_dT, double* dlnphi_drho, double* dlnphi_dv,
                               double* dlnphi_dx, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (flags & TEQPB200_BROADCAST_X) throw InvalidArgument("TEQPB200_BROADCAST_X does not apply to rhovec inputs");
        double* ptr[12] = {pr, splus, dpdT, d2PsirdTdrhoi, dpdrhovec, pmv, chempotVLE, dchempotdT, dlnphi_dT, dlnphi_drho, dlnphi_dv, dlnphi_dx};
        const int rows[12] = {1, 1, 1, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp * Ncomp};
        std::vector<HostOut> outs;
        int slot[12];
        for (int q = 0; q < 12; ++q) { slot[q] = -1; if (ptr[q]) { slot[q] = (int)outs.size(); outs.push_back({ptr[q], rows[q]}); } }
        if (outs.empty()) return;
        run(flags, npts, {{T, 1, false}, {rhovec, Ncomp, false}}, outs,
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                IsoPropsOut o;
                double** f[12] = {&o.pr, &o.splus, &o.dpdT, &o.d2PsirdTdrhoi, &o.dpdrhovec, &o.pmv, &o.chempotVLE, &o.dchempotdT, &o.dlnphi_dT,
                                  &o.dlnphi_drho, &o.dlnphi_dv, &o.dlnphi_dx};
                for (int q = 0; q < 12; ++q) if (slot[q] >= 0) *f[q] = dout[slot[q]];
                m->launch_isoprops(make_batch(n, din[0], nullptr, din[1], false), o, st);
                ++g_launches;
            });
    });
}

int build_Psir_fgradHessian_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec,
                                  const int Ncomp, double* f, double* g, double* H, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { isochoric_call(*lookup(uuid), npts, T, rhovec, Ncomp, f, g, H, nullptr, nullptr, flags); });
}
int get_ln_fugacity_coefficients_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec,
                                       const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { isochoric_call(*lookup(uuid), npts, T, rhovec, Ncomp, nullptr, nullptr, nullptr, out, nullptr, flags); });
}
int get_fugacity_coefficients_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec,
                                    const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { isochoric_call(*lookup(uuid), npts, T, rhovec, Ncomp, nullptr, nullptr, nullptr, nullptr, out, flags); });
}

int get_AtaudeltaXi_table_batch(const long long int uuid, const long long int npts, const double* tau, const double* delta, const double* molefrac,
                                const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { if (npts > 0) ataudelta_table(*lookup(uuid), 1, 0, 0, 0, npts, tau, delta, molefrac, Ncomp, Ncomp * 36, out, flags); });
}
int get_dmBnvirdTm_batch(const long long int uuid, const int Nvir, const int NT, const long long int npts, const double* T, const double* molefrac,
                         const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { if (npts > 0) virial_call(*lookup(uuid), Nvir, NT, npts, T, molefrac, Ncomp, out, flags); });
}
int get_Psir_sigma_derivs_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec, const double* v,
                                const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (flags & TEQPB200_BROADCAST_X) throw InvalidArgument("TEQPB200_BROADCAST_X does not apply to rhovec inputs");
        run(flags, npts, {{T, 1, false}, {rhovec, Ncomp, false}, {v, Ncomp, false}}, {{out, 5}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                m->launch_sigma_derivs(make_batch(n, din[0], nullptr, din[1], false), din[2], dout[0], st);
                ++g_launches;
            });
    });
}

int build_Psir_fgradHessian_and_properties_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec, const int Ncomp,
                                                 double* f, double* g, double* H, double* pr, double* splus, double* dpdT, double* d2PsirdTdrhoi,
                                                 double* dpdrhovec, double* pmv, double* chempotVLE, double* dchempotdT, double* dlnphi_dT, double* dlnphi_drho,
                                                 double* dlnphi_dv, double* dlnphi_dx, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (flags & TEQPB200_BROADCAST_X) throw InvalidArgument("TEQPB200_BROADCAST_X does not apply to rhovec inputs");
        double* ptr[15] = {f, g, H, pr, splus, dpdT, d2PsirdTdrhoi, dpdrhovec, pmv, chempotVLE, dchempotdT, dlnphi_dT, dlnphi_drho, dlnphi_dv, dlnphi_dx};
        const int rows[15] = {1, Ncomp, Ncomp * Ncomp, 1, 1, 1, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp, Ncomp * Ncomp};
        std::vector<HostOut> outs;
        int slot[15];
        for (int q = 0; q < 15; ++q) { slot[q] = -1; if (ptr[q]) { slot[q] = (int)outs.size(); outs.push_back({ptr[q], rows[q]}); } }
        if (outs.empty()) return;
        run(flags, npts, {{T, 1, false}, {rhovec, Ncomp, false}}, outs,
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                IsochoricOut io;
                IsoPropsOut o;
                double** fld[15] = {&io.Psir, &io.grad, &io.hess, &o.pr, &o.splus, &o.dpdT, &o.d2PsirdTdrhoi, &o.dpdrhovec, &o.pmv, &o.chempotVLE, &o.dchempotdT,
                                    &o.dlnphi_dT, &o.dlnphi_drho, &o.dlnphi_dv, &o.dlnphi_dx};
                bool any_iso = false, any_ip = false;
                for (int q = 0; q < 15; ++q) if (slot[q] >= 0) { *fld[q] = dout[slot[q]]; (q < 3 ? any_iso : any_ip) = true; }
                const DeviceBatch b = make_batch(n, din[0], nullptr, din[1], false);
                if (any_iso && any_ip && gerg_try_launch_fused(*m, b, io, o, st)) { ++g_launches; return; }
                if (any_iso) { m->launch_isochoric(b, io, st); ++g_launches; }
                if (any_ip) { m->launch_isoprops(b, o, st); ++g_launches; }
            });
    });
}

int pure_VLE_T_batch(const long long int uuid, const long long int npts, const double* T, const double* rhoL0, const double* rhoV0, const int maxiter,
                     const double* molefrac, const int Ncomp, double* rhoL, double* rhoV, double* psat, double* dpsatdT, const int flags,
                     char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (maxiter < 0) throw InvalidArgument("maxiter must be non-negative");
        if (!molefrac) throw InvalidArgument("molefrac must be given ([1.0] for a one-component model)");
        if (!rhoL || !rhoV) throw InvalidArgument("rhoL and rhoV must not be null");
        const bool want_extra = psat || dpsatdT;
        double R = 0.0;
        if (want_extra) {   // AbstractModel::get_R(molefrac) on the host (VLE_pure.hpp:163): the composition is one Ncomp-vector
            std::vector<double> xh((size_t)Ncomp, 1.0 / Ncomp), xalt((size_t)Ncomp, 0.0);
            xalt[0] = 1.0;
            const bool r_const = m->R(xh.data(), Ncomp) == m->R(xalt.data(), Ncomp);   // every kind but the mole-fraction-weighted multifluid R
            if (r_const) {
                // no need to read the composition back from the device (a stream synchronisation per call)
            } else if (flags & TEQPB200_DEVICE_PTRS) {
                TQ_CUDA(cudaMemcpyAsync(xh.data(), molefrac, sizeof(double) * Ncomp, cudaMemcpyDeviceToHost, tl_stream));
                TQ_CUDA(cudaStreamSynchronize(tl_stream));
            } else {
                for (int i = 0; i < Ncomp; ++i) xh[i] = molefrac[i];
            }
            R = m->R(xh.data(), Ncomp);
        }
        // scratch rows (device only): T2[2], rho2[2], done[1], A[3][2], B[4][2]
        run(flags, npts, {{T, 1, false}, {rhoL0, 1, false}, {rhoV0, 1, false}, {molefrac, Ncomp, true}},
            {{rhoL, 1}, {rhoV, 1}, {psat, 1}, {dpsatdT, 1}, {nullptr, 2}, {nullptr, 2}, {nullptr, 1}, {nullptr, 6}, {nullptr, want_extra ? 8 : 1}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                double *T2 = dout[4], *rho2 = dout[5], *done = dout[6], *A = dout[7], *B = dout[8];
                const int grid = grid_for(n, 256, sm_count() * 8);
                pure_vle_init_kernel<<<grid, 256, 0, st>>>(n, din[0], din[1], din[2], T2, rho2, done);
                TQ_CUDA(cudaGetLastError());
                ++g_launches;
                const DeviceBatch b2 = make_batch(2 * n, T2, rho2, din[3], true);
                for (int it = 0; it < maxiter; ++it) {
                    m->launch_ar_table(0, 2, b2, A, st);
                    pure_vle_update_kernel<<<grid, 256, 0, st>>>(n, A, rho2, done);
                    TQ_CUDA(cudaGetLastError());
                    g_launches += 2;
                }
                if (want_extra) { m->launch_ar_table(1, 1, b2, B, st); ++g_launches; }
                pure_vle_finish_kernel<<<grid, 256, 0, st>>>(n, din[0], rho2, want_extra ? B : nullptr, R, dout[0], dout[1], psat ? dout[2] : nullptr,
                                                             dpsatdT ? dout[3] : nullptr);
                TQ_CUDA(cudaGetLastError());
                ++g_launches;
            });
    });
}

int mix_VLE_Tx_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovecL0, const double* rhovecV0,
                     const double* xspec, const double atol, const double reltol, const double axtol, const double relxtol, const int maxiter,
                     double* rhovecL, double* rhovecV, double* return_code, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        if (m->ncomp() != 2) throw InvalidArgument("mix_VLE_Tx is for binary mixtures (VLE.hpp:105-138); this model has " + std::to_string(m->ncomp()) + " components");
        if (maxiter < 0) throw InvalidArgument("maxiter must be non-negative");
        if (!rhovecL || !rhovecV) throw InvalidArgument("rhovecL and rhovecV must not be null");
        if (flags & TEQPB200_BROADCAST_X) throw InvalidArgument("TEQPB200_BROADCAST_X does not apply: xspec is [2][npts]");
        // model.get_R(xspec) (VLE.hpp:90): one value for the batch, so it must not depend on the composition
        const double xa[2] = {0.25, 0.75}, xb[2] = {0.75, 0.25};
        const double R = m->R(xa, 2);
        if (R != m->R(xb, 2)) throw NotImplementedError("mix_VLE_Tx_batch needs a composition-independent gas constant (this model weights R by mole fractions)");
        // scratch rows (device only): T2[2], rv2[4], f[2], g[4], H[8]
        run(flags, npts, {{T, 1, false}, {rhovecL0, 2, false}, {rhovecV0, 2, false}, {xspec, 2, false}},
            {{rhovecL, 2}, {rhovecV, 2}, {return_code, 1}, {nullptr, 2}, {nullptr, 4}, {nullptr, 2}, {nullptr, 4}, {nullptr, 8}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                double *code = dout[2], *T2 = dout[3], *rv2 = dout[4], *f = dout[5], *g = dout[6], *H = dout[7];
                const int grid = grid_for(n, 256, sm_count() * 8);
                mix_vle_init_kernel<<<grid, 256, 0, st>>>(n, din[0], din[1], din[2], T2, rv2, code);
                TQ_CUDA(cudaGetLastError());
                ++g_launches;
                const DeviceBatch b2 = make_batch(2 * n, T2, nullptr, rv2, false);
                IsochoricOut io; io.Psir = f; io.grad = g; io.hess = H;
                for (int it = 0; it < maxiter; ++it) {
                    m->launch_isochoric(b2, io, st);
                    mix_vle_update_kernel<<<grid, 256, 0, st>>>(n, R, din[0], din[3], f, g, H, atol, reltol, axtol, relxtol, it == maxiter - 1 ? 1 : 0, rv2, code);
                    TQ_CUDA(cudaGetLastError());
                    g_launches += 2;
                }
                mix_vle_finish_kernel<<<grid, 256, 0, st>>>(n, rv2, dout[0], dout[1]);
                TQ_CUDA(cudaGetLastError());
                ++g_launches;
            });
    });
}

int mix_VLE_Tp_batch(const long long int uuid, const long long int npts, const double* T, const double* p, const double* rhovecL0, const double* rhovecV0,
                     const int maxiter, double* rhovecL, double* rhovecV, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        if (m->ncomp() != 2) throw InvalidArgument("mix_VLE_Tp is for binary mixtures (VLE.hpp:196-272); this model has " + std::to_string(m->ncomp()) + " components");
        if (maxiter < 0) throw InvalidArgument("maxiter must be non-negative");
        if (!rhovecL || !rhovecV) throw InvalidArgument("rhovecL and rhovecV must not be null");
        if (flags & TEQPB200_BROADCAST_X) throw InvalidArgument("TEQPB200_BROADCAST_X does not apply");
        const double xa[2] = {0.25, 0.75}, xb[2] = {0.75, 0.25};
        const double R = m->R(xa, 2);
        if (R != m->R(xb, 2)) throw NotImplementedError("mix_VLE_Tp_batch needs a composition-independent gas constant (this model weights R by mole fractions)");
        run(flags, npts, {{T, 1, false}, {p, 1, false}, {rhovecL0, 2, false}, {rhovecV0, 2, false}},
            {{rhovecL, 2}, {rhovecV, 2}, {nullptr, 1}, {nullptr, 2}, {nullptr, 4}, {nullptr, 2}, {nullptr, 4}, {nullptr, 8}},
            [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                double *code = dout[2], *T2 = dout[3], *rv2 = dout[4], *f = dout[5], *g = dout[6], *H = dout[7];
                const int grid = grid_for(n, 256, sm_count() * 8);
                mix_vle_init_kernel<<<grid, 256, 0, st>>>(n, din[0], din[2], din[3], T2, rv2, code);
                TQ_CUDA(cudaGetLastError());
                ++g_launches;
                const DeviceBatch b2 = make_batch(2 * n, T2, nullptr, rv2, false);
                IsochoricOut io; io.Psir = f; io.grad = g; io.hess = H;
                for (int it = 0; it < maxiter; ++it) {
                    m->launch_isochoric(b2, io, st);
                    mix_vle_tp_update_kernel<<<grid, 256, 0, st>>>(n, R, din[0], din[1], f, g, H, rv2);
                    TQ_CUDA(cudaGetLastError());
                    g_launches += 2;
                }
                mix_vle_finish_kernel<<<grid, 256, 0, st>>>(n, rv2, dout[0], dout[1]);
                TQ_CUDA(cudaGetLastError());
                ++g_launches;
            });
    });
}

int teqpb200_point_status_batch(const long long int uuid, const long long int npts, const double* T, const double* rho, const double* molefrac,
                                const int Ncomp, const int nrows, const double* out, int* status, const int flags, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (nrows < 0 || (nrows > 0 && !out)) throw InvalidArgument("nrows must be non-negative and out non-null when nrows > 0");
        if (!status) throw InvalidArgument("status must not be null");
        if (npts < 0) throw InvalidArgument("npts must be non-negative");
        if (npts == 0) return;
        const bool bc = flags & TEQPB200_BROADCAST_X;
        if (flags & TEQPB200_DEVICE_PTRS) {
            point_status_kernel<int><<<grid_for(npts, 256, sm_count() * 8), 256, 0, tl_stream>>>(make_batch(npts, T, rho, molefrac, bc), Ncomp, out, nrows, status);
            TQ_CUDA(cudaGetLastError());
            ++g_launches;
            return;
        }
        // host buffers: the staging pipeline moves rows of doubles, so the status row is staged as doubles and narrowed on the host
        std::vector<double> tmp((size_t)npts);
        std::vector<HostIn> ins = {{T, 1, false}, {rho, 1, false}, {molefrac, Ncomp, bc}};
        if (nrows > 0) ins.push_back({out, nrows, false});
        run_host(npts, ins, {{tmp.data(), 1}},
                 [&](long long n, const std::vector<const double*>& din, const std::vector<double*>& dout, cudaStream_t st) {
                     point_status_kernel<double><<<grid_for(n, 256, sm_count() * 8), 256, 0, st>>>(make_batch(n, din[0], din[1], din[2], bc), Ncomp,
                                                                                                   nrows > 0 ? din[3] : nullptr, nrows, dout[0]);
                     TQ_CUDA(cudaGetLastError());
                     ++g_launches;
                 });
        for (long long k = 0; k < npts; ++k) status[k] = (int)tmp[(size_t)k];
    });
}

// ---------------------------------------------------------------------------------------------------------------------
// utilities
// ---------------------------------------------------------------------------------------------------------------------
int teqpb200_get_reducing(const long long int uuid, const int which, const double* molefrac, const int Ncomp, double* val, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        double Tr = 0, rhor = 0;
        // the reference offers these for multi-fluid models only (deriv_adapter.hpp: has_meth_get_reducing_*)
        if (!m->reducing(molefrac, Ncomp, &Tr, &rhor)) throw NotImplementedError(std::string("reducing functions are not defined for model family ") + m->family());
        *val = which == 0 ? Tr : rhor;
    });
}
int teqpb200_get_R(const long long int uuid, const double* molefrac, const int Ncomp, double* R, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { auto m = lookup(uuid); check_ncomp(*m, Ncomp); *R = m->R(molefrac, Ncomp); });
}
int teqpb200_get_ncomp(const long long int uuid, int* Ncomp, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { *Ncomp = lookup(uuid)->ncomp(); });
}
int teqpb200_generated_kernel(const long long int uuid, const int NT, const int ND, const int which, void* buf, unsigned long long* nbytes,
                              char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        std::vector<char> bytes;
        if (which == 2) bytes = m->jit_cubin(NT, ND);
        else {
            std::vector<double> consts;
            const std::string src = m->jit_source(NT, ND, &consts);
            if (which == 0) bytes.assign(src.c_str(), src.c_str() + src.size() + 1);
            else bytes.assign(reinterpret_cast<const char*>(consts.data()), reinterpret_cast<const char*>(consts.data() + consts.size()));
        }
        if (bytes.empty() || (which == 0 && bytes.size() == 1)) throw NotImplementedError(std::string("no generated kernels for model family ") + m->family());
        if (buf) std::memcpy(buf, bytes.data(), (size_t)std::min<unsigned long long>(*nbytes, bytes.size()));
        *nbytes = bytes.size();
    });
}
int teqpb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
int teqpb200_set_devices(const int* device_ids, const int n, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        const int have = teqpb200_device_count();
        std::vector<int> d;
        for (int i = 0; i < n; ++i) {
            if (device_ids[i] < 0 || device_ids[i] >= have) throw InvalidArgument("no such CUDA device: " + std::to_string(device_ids[i]));
            d.push_back(device_ids[i]);
        }
        std::lock_guard<std::mutex> g(g_dev_mtx);
        g_devices = d;
    });
}
int teqpb200_set_stream(void* cuda_stream) { tl_stream = static_cast<cudaStream_t>(cuda_stream); return 0; }
void* teqpb200_host_alloc(unsigned long long nbytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, nbytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void teqpb200_host_free(void* p) { if (p) cudaFreeHost(p); }
long long int teqpb200_kernel_launches(void) { return g_launches.load(); }
int teqpb200_release_scratch(void) {
    // frees the per-device staging arenas of the host-pointer path (they only grow otherwise); safe between calls
    std::lock_guard<std::mutex> g(g_state_mtx);
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto& [dev, S] : g_states) {
        std::lock_guard<std::mutex> l(S->mtx);
        cudaSetDevice(dev);
        for (auto& A : S->arenas) {
            if (A.st) cudaStreamSynchronize(A.st);
            if (A.p) cudaFree(A.p);
            A.p = nullptr; A.cap = 0;
        }
    }
    cudaSetDevice(cur);
    {   // and the stream-ordered scratch the device-pointer calls left cached in the default memory pool (scratch_pool_setup)
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, cur) == cudaSuccess && pool) { cudaDeviceSynchronize(); cudaMemPoolTrimTo(pool, 0); }
    }
    cudaGetLastError();
    return 0;
}
void teqpb200_inject_fault(const long long int chunk_index) { g_chunk_counter = 0; g_fault_chunk = chunk_index; }
double teqpb200_last_call_device_ms(const int device) {
    std::lock_guard<std::mutex> g(g_time_mtx);
    auto it = g_last_ms.find(device);
    return it == g_last_ms.end() ? -1.0 : it->second;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// synthetic state points
// ---------------------------------------------------------------------------------------------------------------------
namespace {
__device__ __forceinline__ double u01(unsigned long long seed, unsigned long long k, unsigned long long j) {
    unsigned long long z = seed ^ (k * 0x9E3779B97F4A7C15ULL + j);
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z = z ^ (z >> 31);
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}
constexpr int kSynthMaxC = 32;
struct SynthParams { double z0[kSynthMaxC]; };
__global__ void synth_kernel(const __grid_constant__ SynthParams sp, long long npts, long long k0, unsigned long long seed, double Tlo, double Thi,
                             double lnlo, double lnhi, int nc, double* T, double* rho, double* x) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride) {
        const unsigned long long gk = (unsigned long long)(k0 + k);
        T[k] = Tlo + (Thi - Tlo) * u01(seed, gk, 0);
        rho[k] = exp(lnlo + (lnhi - lnlo) * u01(seed, gk, 1));
        double s = 0.0;
        for (int i = 0; i < nc; ++i) {
            const double v = sp.z0[i] * exp(0.2 * (2.0 * u01(seed, gk, 2 + i) - 1.0));
            x[(long long)i * npts + k] = v;
            s += v;
        }
        for (int i = 0; i < nc; ++i) x[(long long)i * npts + k] /= s;
    }
}
}  // namespace

extern "C" int teqpb200_synth_points(const long long int npts, const long long int k0, const unsigned long long seed, const double Tlo,
                                     const double Thi, const double rholo, const double rhohi, const double* z0_host, const int Ncomp,
                                     double* T, double* rho, double* x, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        if (Ncomp < 1 || Ncomp > kSynthMaxC) throw InvalidArgument("Ncomp out of range for the synthetic generator");
        if (npts <= 0) return;
        SynthParams sp{};
        for (int i = 0; i < Ncomp; ++i) sp.z0[i] = z0_host[i];
        const int block = 256;
        const int grid = grid_for(npts, block, sm_count() * 8);
        synth_kernel<<<grid, block, 0, tl_stream>>>(sp, npts, k0, seed, Tlo, Thi, std::log(rholo), std::log(rhohi), Ncomp, T, rho, x);
        TQ_CUDA(cudaGetLastError());
        ++g_launches;
    });
}

// ---------------------------------------------------------------------------------------------------------------------
// FP64 FMA-pipe peak of the current device, measured: 8 independent DFMA chains per thread, enough resident warps to saturate the
// pipe.  This is the denominator of the "% of FP64 peak" figure (MEASURED_PEAKS.json has no FP64 entry).
// ---------------------------------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) dfma_peak_kernel(double* sink, int iters, double a, double b) {
    double r0 = threadIdx.x, r1 = r0 + 1, r2 = r0 + 2, r3 = r0 + 3, r4 = r0 + 4, r5 = r0 + 5, r6 = r0 + 6, r7 = r0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            r0 = fma(r0, a, b); r1 = fma(r1, a, b); r2 = fma(r2, a, b); r3 = fma(r3, a, b);
            r4 = fma(r4, a, b); r5 = fma(r5, a, b); r6 = fma(r6, a, b); r7 = fma(r7, a, b);
        }
    }
    const double s = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
    if (s == 123.456) sink[0] = s;  // never true; keeps the chains alive
}
}  // namespace

extern "C" int teqpb200_fp64_peak(double* tflops, char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        double* sink = nullptr;
        TQ_CUDA(cudaMalloc(&sink, sizeof(double)));
        const int iters = 4096, block = 256, grid = sm_count() * 8;
        cudaEvent_t e0, e1;
        TQ_CUDA(cudaEventCreate(&e0)); TQ_CUDA(cudaEventCreate(&e1));
        double best = 0.0;
        for (int rep = 0; rep < 5; ++rep) {
            TQ_CUDA(cudaEventRecord(e0, 0));
            dfma_peak_kernel<<<grid, block>>>(sink, iters, 0.999999, 1e-9);
            TQ_CUDA(cudaEventRecord(e1, 0));
            TQ_CUDA(cudaEventSynchronize(e1));
            float ms = 0;
            TQ_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            const double flops = 2.0 * 64.0 * iters * (double)block * grid;
            if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
        }
        TQ_CUDA(cudaEventDestroy(e0)); TQ_CUDA(cudaEventDestroy(e1));
        TQ_CUDA(cudaFree(sink));
        *tflops = best;
    });
}

// ---------------------------------------------------------------------------------------------------------------------
// No-allocation variant of get_Arxy_batch (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA graph capture, tight serving loops).  get_Arxy_batch evaluates the covering sweeps into a device
// scratch table it obtains from cudaMallocAsync and gathers the requested rows from it; here the scratch table is a file-scope device
// array of kNoAllocMaxDoubles doubles (32 MiB, held by every GPU that loads the library until the process exits).  The array's name
// is a host-side stand-in in host code: its device address is resolved once per device with cudaGetSymbolAddress and passed on.
// Every call shares the array (the sweeps overwrite every scratch row they use before it is read, so nothing has to be cleared);
// calls that overlap on the GPU -- different streams, different host threads -- race on it.
// ---------------------------------------------------------------------------------------------------------------------
constexpr long long kNoAllocMaxDoubles = 4ll << 20;
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxDoubles];

namespace {
constexpr int kNoAllocMaxPairs = 32;
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mtx;
double* g_noalloc_ptr[kNoAllocMaxDevices] = {};

struct NoAllocGather { int npairs; int src_row[kNoAllocMaxPairs]; };

// out[p][k] = scratch[src_row[p]][k]; scratch is the DEVICE address of g_noalloc_scratch (from cudaGetSymbolAddress)
__global__ void noalloc_gather_kernel(const double* __restrict__ scratch, const NoAllocGather g, const long long npts, double* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < npts; k += stride)
        for (int p = 0; p < g.npairs; ++p) out[(long long)p * npts + k] = scratch[(long long)g.src_row[p] * npts + k];
}

// device address of the scratch array on the CURRENT device (resolved once per device, with that device current)
double* noalloc_scratch_ptr() {
    int dev = 0;
    TQ_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) throw InvalidArgument("device ordinal out of range for the no-allocation entry point");
    std::lock_guard<std::mutex> g(g_noalloc_mtx);
    if (!g_noalloc_ptr[dev]) {
        double* p = nullptr;
        TQ_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) throw std::runtime_error("cudaGetSymbolAddress returned a null address for the no-allocation scratch array");
        g_noalloc_ptr[dev] = p;
    }
    return g_noalloc_ptr[dev];
}
}  // namespace

extern "C" {

long long int teqpb200_noalloc_max_doubles(void) { return kNoAllocMaxDoubles; }

int teqpb200_noalloc_setup(char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] { (void)noalloc_scratch_ptr(); });
}

int get_Arxy_batch_noalloc(const long long int uuid, const int npairs, const int* NT, const int* ND, const long long int npts,
                           const double* T, const double* rho, const double* molefrac, const int Ncomp, double* out, const int flags,
                           char* errmsg, int errmsg_length) {
    return guarded(false, errmsg, errmsg_length, [&] {
        auto m = lookup(uuid);
        check_ncomp(*m, Ncomp);
        if (!(flags & TEQPB200_DEVICE_PTRS)) throw InvalidArgument("get_Arxy_batch_noalloc takes device pointers only (TEQPB200_DEVICE_PTRS)");
        if (npairs <= 0 || npairs > kNoAllocMaxPairs) throw InvalidArgument("npairs must be in 1.." + std::to_string(kNoAllocMaxPairs));
        if (npts < 0) throw InvalidArgument("npts must be non-negative");
        if (npts == 0) return;
        // same partition of the pairs over the compiled sweeps as get_Arxy_batch (row NT = 0, column ND = 0, block NT <= 2, ND <= 4)
        int rowND = -1, colNT = -1, blkNT = -1, blkND = -1;
        int where[kNoAllocMaxPairs];
        for (int p = 0; p < npairs; ++p) {
            const int nt = NT[p], nd = ND[p];
            if (nt < 0 || nd < 0 || nt > kMaxDerivT || nd > kMaxDerivD) throw InvalidArgument("derivative order out of range");
            if (nt == 0) where[p] = 0;
            else if (nd == 0) where[p] = 1;
            else if (nt <= 2 && nd <= 4) { where[p] = 2; blkNT = std::max(blkNT, nt); blkND = std::max(blkND, nd); }
            else throw InvalidArgument("mixed derivative (" + std::to_string(nt) + "," + std::to_string(nd) + ") is outside the compiled sweeps: NT = 0 with ND <= 6, ND = 0 with NT <= 4, or NT <= 2 with ND <= 4");
        }
        for (int p = 0; p < npairs; ++p) {
            if (blkNT >= 0 && where[p] == 0 && ND[p] <= 4) { where[p] = 2; blkND = std::max(blkND, ND[p]); }
            else if (blkNT >= 0 && where[p] == 1 && NT[p] <= 2) { where[p] = 2; blkNT = std::max(blkNT, NT[p]); }
        }
        for (int p = 0; p < npairs; ++p) {
            if (where[p] == 0) rowND = std::max(rowND, ND[p]);
            if (where[p] == 1) colNT = std::max(colNT, NT[p]);
        }
        const int rows_row = rowND >= 0 ? rowND + 1 : 0, rows_col = colNT >= 0 ? colNT + 1 : 0, rows_blk = blkNT >= 0 ? (blkNT + 1) * (blkND + 1) : 0;
        const long long rows = rows_row + rows_col + rows_blk;
        if (rows * npts > kNoAllocMaxDoubles)
            throw InvalidArgument("get_Arxy_batch_noalloc: " + std::to_string(rows) + " scratch rows x " + std::to_string(npts) + " points exceed the static scratch array of " +
                                  std::to_string(kNoAllocMaxDoubles) + " doubles; split the batch or call get_Arxy_batch");
        // per-model generated kernels are compiled and loaded (which allocates) at the first use of a shape: refuse rather than allocate
        if ((rows_row && !m->ar_table_ready(0, rowND)) || (rows_col && !m->ar_table_ready(colNT, 0)) || (rows_blk && !m->ar_table_ready(blkNT, blkND)))
            throw InvalidArgument("get_Arxy_batch_noalloc: this model's kernel for the requested shape has not been generated yet; call get_Arxy_batch once with the same pairs first");
        double* const scratch = noalloc_scratch_ptr();
        const bool bc = flags & TEQPB200_BROADCAST_X;
        const DeviceBatch b = make_batch(npts, T, rho, molefrac, bc);
        double* srow = scratch;
        double* scol = srow + (size_t)rows_row * npts;
        double* sblk = scol + (size_t)rows_col * npts;
        cudaStream_t st = tl_stream;
        if (rows_row) { m->launch_ar_table(0, rowND, b, srow, st); ++g_launches; }
        if (rows_col) { m->launch_ar_table(colNT, 0, b, scol, st); ++g_launches; }
        if (rows_blk) { m->launch_ar_table(blkNT, blkND, b, sblk, st); ++g_launches; }
        NoAllocGather g{};
        g.npairs = npairs;
        for (int p = 0; p < npairs; ++p)
            g.src_row[p] = where[p] == 0 ? ND[p] : where[p] == 1 ? rows_row + NT[p] : rows_row + rows_col + NT[p] * (blkND + 1) + ND[p];
        noalloc_gather_kernel<<<grid_for(npts, 256, 8 * sm_count()), 256, 0, st>>>(scratch, g, npts, out);
        TQ_CUDA(cudaGetLastError());
        ++g_launches;
    });
}

}  // extern "C"
