// dequant.cu — bit-exact dequantize and GET_ROWS for GGUF block formats.
//
// Replaces the reference's dequantize_block_* / k_get_rows kernels (ggml-cuda/convert.cu:82-290,
// ggml-cuda/getrows.cu:5-67) for every type in b200q.h — including the K-quants and MXFP4 that the
// reference CUDA GET_ROWS declines (ggml-cuda.cu:3332-3348) but ggml-cpu supports (ops.cpp:5121-5162).
// Results are bit-identical to dequantize_row_* (ggml-quants.c) because decode_elem() performs the same
// fp32 operations with the same roundings (no FMA contraction; see blocks.cuh).
//
// HBM-bound byte work: each thread produces 4 consecutive f32 outputs (one 16-byte store, fully coalesced
// across the warp); the quantized bytes of a row are re-read by neighbouring threads out of L1.
#include "blocks.cuh"
#include "dequant.cuh"

#include "pdl.cuh"
#include <cuda_runtime.h>

namespace b200q {

namespace {

template <int T> __device__ __forceinline__ float decode_any(const uint8_t * row, int64_t i) {
    return decode_elem<T>(row, i);
}
template <> __device__ __forceinline__ float decode_any<T_F32>(const uint8_t * row, int64_t i) {
    return ((const float *) row)[i];
}

// dst[(i12, i11, i10)][i00] = decode(src0 row ids[i10, i11, i12] of plane (i11, i12))[i00]
template <int T> __global__ void __launch_bounds__(256) get_rows_kernel(const GetRowsParams p) {
    pdl_grid_sync();
    const int64_t quads_per_row = (p.ne00 + 3) >> 2;
    const int64_t total         = quads_per_row * p.n_rows;
    for (int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t) gridDim.x * blockDim.x) {
        const int64_t rowi = idx / quads_per_row;
        const int64_t i00  = (idx - rowi * quads_per_row) << 2;
        int64_t       i10 = rowi, i11 = 0, i12 = 0;
        const uint8_t * src;
        float *         out;
        if (p.ids) {
            i12                 = rowi / (p.ne10 * p.ne11);
            const int64_t rem   = rowi - i12 * p.ne10 * p.ne11;
            i11                 = rem / p.ne10;
            i10                 = rem - i11 * p.ne10;
            const int32_t r     = *(const int32_t *) ((const char *) p.ids + i10 * p.nb10 + i11 * p.nb11 + i12 * p.nb12);
            src                 = p.src + (int64_t) r * p.nb01 + i11 * p.nb02 + i12 * p.nb03;
            out                 = (float *) ((char *) p.dst + i10 * p.nb1 + i11 * p.nb2 + i12 * p.nb3);
        } else {
            src = p.src + rowi * p.nb01;
            out = (float *) ((char *) p.dst + rowi * p.nb1);
        }
        if (i00 + 3 < p.ne00 && p.dst_vec4) {
            float4 v;
            v.x = decode_any<T>(src, i00 + 0);
            v.y = decode_any<T>(src, i00 + 1);
            v.z = decode_any<T>(src, i00 + 2);
            v.w = decode_any<T>(src, i00 + 3);
            *(float4 *) (out + i00) = v;
        } else {
            for (int64_t i = i00; i < p.ne00 && i < i00 + 4; ++i) {
                out[i] = decode_any<T>(src, i);
            }
        }
    }
}

}  // namespace

get_rows_kernel_t get_rows_kernel_for(int type) {
    switch (type) {
        case T_F32:   return get_rows_kernel<T_F32>;
        case T_Q4_0:  return get_rows_kernel<T_Q4_0>;
        case T_Q4_1:  return get_rows_kernel<T_Q4_1>;
        case T_Q5_0:  return get_rows_kernel<T_Q5_0>;
        case T_Q5_1:  return get_rows_kernel<T_Q5_1>;
        case T_Q8_0:  return get_rows_kernel<T_Q8_0>;
        case T_MXFP4: return get_rows_kernel<T_MXFP4>;
        case T_Q4_K:  return get_rows_kernel<T_Q4_K>;
        case T_Q5_K:  return get_rows_kernel<T_Q5_K>;
        case T_Q6_K:  return get_rows_kernel<T_Q6_K>;
        case T_F16:   return get_rows_kernel<T_F16>;
        case T_BF16:  return get_rows_kernel<T_BF16>;
        case T_Q2_K:  return get_rows_kernel<T_Q2_K>;
        case T_Q3_K:  return get_rows_kernel<T_Q3_K>;
        case T_IQ4_NL: return get_rows_kernel<T_IQ4_NL>;
        case T_IQ4_XS: return get_rows_kernel<T_IQ4_XS>;
        default: return nullptr;
    }
}

}  // namespace b200q
