// pdl.cuh — programmatic dependent launch for the small kernels of a decode step.
// A decode token through llama.cpp is ~700 dependent launches of kernels that run for 1-2 us; launched plainly (also inside a
// CUDA graph) every one of them pays the full launch latency after its predecessor has drained.  With the programmatic
// stream-serialization attribute the next kernel is set up while the previous one runs: every kernel here first lets its
// dependents launch (griddepcontrol.launch_dependents) and then waits for ALL earlier work of the stream to complete and
// become visible (griddepcontrol.wait) before it touches memory — ordering and visibility are exactly those of a plain
// launch, only the launch latency overlaps.  Rule: a kernel launched through launch_pdl() must call pdl_grid_sync() as its
// first statement.  B200Q_NO_PDL=1 launches plainly (measurement).
#pragma once
#include <cuda_runtime.h>

#include <cstdlib>
#include <utility>

namespace b200q {

__device__ __forceinline__ void pdl_grid_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&... args) {
    static const bool off = getenv("B200Q_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim          = grid;
    cfg.blockDim         = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[1];
    attr[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs                                          = attr;
    cfg.numAttrs                                       = off ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace b200q
