// NVRTC runtime of libteqpb200: CUDA C source -> sm_100a cubin -> loaded kernel.  libnvrtc is opened with dlopen at first use, so the
// library itself loads (and the CPU test tier can inspect generated sources) on machines without it; a model whose kernel cannot be
// compiled fails loudly -- there is no interpreter or CPU fallback behind this path.
// The cubin is loaded through the runtime's context-independent library API (cudaLibraryLoadData / cudaLibraryGetKernel), so one
// compiled kernel serves every device the host driver shards over.
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "tq_common.hpp"
#include "tq_gerg_jit.hpp"

namespace tq {

namespace {

// the handful of NVRTC entry points used, resolved at run time (nvrtc.h:  nvrtcCreateProgram, nvrtcCompileProgram, ...)
struct Nvrtc {
    void* h = nullptr;
    int (*CreateProgram)(void**, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
    int (*CompileProgram)(void*, int, const char* const*) = nullptr;
    int (*GetProgramLogSize)(void*, size_t*) = nullptr;
    int (*GetProgramLog)(void*, char*) = nullptr;
    int (*GetCUBINSize)(void*, size_t*) = nullptr;
    int (*GetCUBIN)(void*, char*) = nullptr;
    int (*DestroyProgram)(void**) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string tried;
};

Nvrtc& nvrtc() {
    static Nvrtc N;
    static std::once_flag once;
    std::call_once(once, [] {
        std::vector<std::string> names;
        if (const char* e = std::getenv("TEQPB200_NVRTC")) names.push_back(e);
        for (const char* n : {"libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so", "libnvrtc.so"}) names.push_back(n);
        for (const auto& n : names) {
            N.h = dlopen(n.c_str(), RTLD_NOW | RTLD_LOCAL);
            if (N.h) break;
            N.tried += n + " ";
        }
        if (!N.h) return;
#define TQ_SYM(f) N.f = reinterpret_cast<decltype(N.f)>(dlsym(N.h, "nvrtc" #f))
        TQ_SYM(CreateProgram); TQ_SYM(CompileProgram); TQ_SYM(GetProgramLogSize); TQ_SYM(GetProgramLog); TQ_SYM(GetCUBINSize); TQ_SYM(GetCUBIN);
        TQ_SYM(DestroyProgram); TQ_SYM(GetErrorString);
#undef TQ_SYM
    });
    if (!N.h || !N.CreateProgram || !N.CompileProgram || !N.GetCUBIN)
        throw DeviceError("libteqpb200 needs NVRTC to build per-model kernels and could not load it (tried: " + N.tried + "; set TEQPB200_NVRTC)");
    return N;
}

}  // namespace

struct JitKernel {
    cudaLibrary_t lib = nullptr;
    cudaKernel_t kern = nullptr;
    int regs = 0;
};

JitKernel* jit_compile(const std::string& code, const char* entry, std::string* log_out, std::vector<char>* cubin_out, bool load) {
    Nvrtc& N = nvrtc();
    void* prog = nullptr;
    int rc = N.CreateProgram(&prog, code.c_str(), "tqj_model.cu", 0, nullptr, nullptr);
    if (rc) throw DeviceError(std::string("nvrtcCreateProgram: ") + N.GetErrorString(rc));
    const char* opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "-lineinfo", "--fmad=true", "-default-device"};
    rc = N.CompileProgram(prog, 5, opts);
    std::string log;
    size_t ls = 0;
    if (N.GetProgramLogSize(prog, &ls) == 0 && ls > 1) { log.resize(ls); N.GetProgramLog(prog, log.data()); }
    if (log_out) *log_out = log;
    if (rc) {
        N.DestroyProgram(&prog);
        throw DeviceError(std::string("NVRTC failed to compile a generated kernel: ") + N.GetErrorString(rc) + "\n" + log.substr(0, 2000));
    }
    size_t cs = 0;
    N.GetCUBINSize(prog, &cs);
    std::vector<char> cubin(cs);
    N.GetCUBIN(prog, cubin.data());
    N.DestroyProgram(&prog);
    if (cubin_out) *cubin_out = cubin;
    if (!load) return nullptr;
    auto k = new JitKernel();
    try {
        TQ_CUDA(cudaLibraryLoadData(&k->lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0));
        TQ_CUDA(cudaLibraryGetKernel(&k->kern, k->lib, entry));
        cudaFuncAttributes fa{};
        if (cudaFuncGetAttributes(&fa, (const void*)k->kern) == cudaSuccess) k->regs = fa.numRegs; else cudaGetLastError();
    } catch (...) {
        if (k->lib) cudaLibraryUnload(k->lib);
        delete k;
        throw;
    }
    return k;
}

void jit_release(JitKernel* k) {
    if (!k) return;
    if (k->lib) cudaLibraryUnload(k->lib);
    delete k;
}

int jit_kernel_regs(const JitKernel* k) { return k ? k->regs : 0; }

void jit_launch_table(const JitKernel* k, const std::vector<double>& consts, size_t smem_bytes, const DeviceBatch& in, double* out, cudaStream_t st) {
    if (in.npts == 0) return;
    long long npts = in.npts, xstride = in.x_stride;
    int xbcast = in.x_bcast;
    const double *T = in.T, *rho = in.rho, *x = in.x;
    // the first parameter is the constant pool BY VALUE (a __grid_constant__ struct of consts.size() doubles)
    void* args[] = {(void*)consts.data(), &npts, &T, &rho, &x, &xstride, &xbcast, &out};
    if (smem_bytes > 48 * 1024) TQ_CUDA(cudaFuncSetAttribute((const void*)k->kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    const int grid = grid_for(npts, 128, 1 << 30);
    TQ_CUDA(cudaLaunchKernel((const void*)k->kern, dim3(grid), dim3(128), args, smem_bytes, st));
}

}  // namespace tq
