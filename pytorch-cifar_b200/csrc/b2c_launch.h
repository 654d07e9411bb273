// b2c_launch.h — kernel launches with programmatic dependent launch (PDL).
//
// A training step is ~150 dependent launches; between two of them the GPU otherwise idles for the launch latency
// plus the next kernel's prologue (barrier init, TMEM allocation, descriptor prefetch). Every kernel of this library
// starts with `pdl_launch_dependents()` — its successor may be scheduled as soon as all of this grid's CTAs have
// started — and calls `pdl_wait()` before its first access to global memory, which blocks until the predecessor grid
// has completed and its writes are visible. The launch attribute below is what makes the pair effective; without it
// (B2C_PDL=0, or a predecessor that is not one of ours) both instructions are no-ops and launches serialise as usual.
// Works under stream capture (programmatic edges in the CUDA graph).
#pragma once
#include <cuda_runtime.h>

#include <cstdlib>
#include <utility>

namespace b2c {

inline bool pdl_enabled() {
    static const bool on = [] {
        const char* e = std::getenv("B2C_PDL");
        return !(e != nullptr && e[0] == '0');
    }();
    return on;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace b2c

#define B2C_LAUNCH(kern, grid, block, smem, stream, ...) \
    (void)::b2c::launch_pdl(kern, dim3(grid), dim3(block), smem, stream, __VA_ARGS__)
