// Local-memory (stack) sizing for kernels with large per-thread frames.
//
// Measured on B200 (driver 580, CUDA 12.9; tools/gpu_debug_fault.sh, profiles/r02w_stack_limit.txt): the jet-sweep kernels of the GERG / multifluid family
// (3.3 KB frame, 32 KB __grid_constant__ parameter block) fault with "illegal memory access" when they are the FIRST large-frame kernel a process
// launches -- cudaLimitStackSize is still 1024 bytes and the driver does not grow the per-thread local memory for them -- and run correctly once the limit
// covers the frame (either because an earlier kernel of another family made the driver grow it, which is what hid the fault in every full test run, or
// because the limit was set) -- with some headroom: frame + 24 bytes still faulted, frame + 288 / + 536 ran.  The cause of the extra need is not understood
// (the host replay of the same device functions is clean under AddressSanitizer; kernels of the other families with larger frames and smaller parameter blocks
// are sized by the driver without help).  ensure_stack() queries the frame of each kernel a launcher may start (cudaFuncGetAttributes: localSizeBytes) and raises
// the limit of the current device when needed, once per (device, launcher).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <initializer_list>
#include <map>
#include <mutex>

#include "tq_common.hpp"

namespace tq {

inline void ensure_stack(std::initializer_list<const void*> kernels) {
    static std::mutex mtx;
    static std::map<std::pair<int, const void*>, bool> done;
    int dev = 0;
    TQ_CUDA(cudaGetDevice(&dev));
    const void* key = *kernels.begin();
    {
        std::lock_guard<std::mutex> g(mtx);
        if (done.count({dev, key})) return;
    }
    size_t need = 0;
    for (const void* k : kernels) {
        cudaFuncAttributes a{};
        TQ_CUDA(cudaFuncGetAttributes(&a, k));
        if (a.localSizeBytes > need) need = a.localSizeBytes;
    }
    size_t cur = 0;
    TQ_CUDA(cudaDeviceGetLimit(&cur, cudaLimitStackSize));
    if (std::getenv("TEQPB200_DEBUG_STACK")) std::fprintf(stderr, "[teqpb200] ensure_stack: frame %zu B, cudaLimitStackSize %zu B\n", need, cur);
    // headroom: a limit that merely covers the reported frame is NOT enough (measured: the 3048-byte sigma kernel still faults at 3072 and runs at 3584),
    // so the limit is the frame plus 1 KB, rounded up to 1 KB
    const size_t want = (need + 1024 + 1023) & ~size_t(1023);
    if (cur < want) TQ_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, want));
    std::lock_guard<std::mutex> g(mtx);
    done[{dev, key}] = true;
}
#define TQ_K(...) reinterpret_cast<const void*>(&__VA_ARGS__)

}  // namespace tq
