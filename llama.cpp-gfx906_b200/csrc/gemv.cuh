// gemv.cuh — launch parameters of the streaming GEMV kernel (gemv.cu)
#pragma once
#include <stdint.h>

namespace b200q {

enum { GEMV_MODE_BATCH = 0, GEMV_MODE_MOE = 1 };

struct GemvParams {
    const uint8_t * w;     // quantized weights
    const float *   act;   // f32 activations
    float *         dst;   // f32 result
    const int32_t * ids;   // MoE expert ids (device) or NULL
    int k, m, nch;         // row length, rows per channel, channels
    int mode;
    int ne12, r2, r3;      // batch mode: channel -> (i12, i13), broadcast ratios
    int n_used, n_expert, b_bcast;  // MoE mode
    long long w_nb1, w_nb2, w_nb3;  // byte strides
    long long a_nb1, a_nb2, a_nb3;
    long long d_nb1, d_nb2, d_nb3;
    long long ids_nb1;
    int row_bytes;         // bytes of one quantized row
    int contiguous;        // w_nb1 == row_bytes: one bulk copy per stage
    int smem_pitch;        // row slot size inside a stage when !contiguous
    int wpr;               // warps cooperating on one row pair (1, 2, 4, 8)
    int passes;            // row passes per stage
    int stage_bytes, nstages, stage_off;
    int act_vec4;          // activation columns are 16-byte aligned
    int early_prefetch;    // weights immutable: stream them before griddepcontrol.wait
};

typedef void (*gemv_kernel_t)(const GemvParams);

gemv_kernel_t gemv_kernel_for(int type, int ncols);
int           gemv_threads();
int           gemv_fixed_smem();
int           gemv_pass_rows(int wpr);

}  // namespace b200q
