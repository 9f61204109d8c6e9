// dense_mm.cuh — unquantized weights (f16 / bf16 / f32) outside the tensor-core kernel's shapes: see dense_mm.cu
#pragma once
#include "../../include/b200q.h"
#include "mmq.cuh"

#include <cuda_runtime.h>

namespace b200q {
bool dense_mm_supported(int type, int64_t k, int64_t m, int64_t n, int64_t nbatch);
int  dense_mm_launch(const MmqArgs & a, cudaStream_t st);  // strides in bytes; allocates nothing, needs no workspace
}  // namespace b200q
