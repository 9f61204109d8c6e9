// dequant.cuh — launch parameters of the dequantize / GET_ROWS kernel (dequant.cu)
#pragma once
#include <stdint.h>

namespace b200q {

struct GetRowsParams {
    const uint8_t * src;
    const int32_t * ids;   // NULL: plain dequantize of n_rows consecutive rows (stride nb01)
    float *         dst;
    int64_t ne00;          // row length (elements)
    int64_t n_rows;        // ne10*ne11*ne12 (or number of rows when ids == NULL)
    int64_t ne10, ne11;
    int64_t nb01, nb02, nb03;
    int64_t nb10, nb11, nb12;
    int64_t nb1, nb2, nb3;
    int     dst_vec4;      // dst rows are 16-byte aligned
};

typedef void (*get_rows_kernel_t)(const GetRowsParams);
get_rows_kernel_t get_rows_kernel_for(int type);

}  // namespace b200q
