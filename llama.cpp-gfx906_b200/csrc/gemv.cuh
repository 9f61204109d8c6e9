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
    int smem_pitch;        // distance between rows inside a stage
    // consumer tiling (see gemv.cu): 8 consumer warps = `groups` groups of (8/groups) warps; a group owns every
    // groups-th stage; inside a group `wpr` warps share one row pair; a warp step covers (32/lpr) * 2 rows
    int lpr;               // lanes per row (power of two <= 32)
    int wpr;               // warps per row pair (1, 2, 4, 8); > 1 only when lpr == 32
    int groups;            // 1, 2, 4, 8 (1 when nch > 1)
    int passes;            // row passes per stage
    int stage_rows;        // rows per stage = passes * pass_rows
    int stage_bytes, nstages, stage_off;
    int act_col_bytes;     // bytes between quantized activation columns in smem (k rounded up to 128)
    int ds_off;            // smem offset of the (d, s) pairs
    int act_vec4;          // activation columns are 16-byte aligned in global memory
    int early_stages;      // stages queued before the consumers' activation loads are in flight
    int early_prefetch;    // weights immutable: stream them before griddepcontrol.wait
    unsigned long long * ts;  // debug: per-CTA %globaltimer stamps [grid][8] (NULL in production)
};

typedef void (*gemv_kernel_t)(const GemvParams);

gemv_kernel_t gemv_kernel_for(int type, int ncols);
int           gemv_threads();
int           gemv_consumer_warps();
int           gemv_act_off();        // smem offset where the quantized activations start
int           gemv_unit_elems(int type);

}  // namespace b200q
