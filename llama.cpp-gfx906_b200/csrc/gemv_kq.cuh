// gemv_kq.cuh — launch parameters of the K-quant decode GEMV (gemv_kq.cu): n = 1, q4_K / q6_K, up to 4 weight
// matrices that share one activation vector (q/k/v, gate/up) in ONE launch.
#pragma once
#include <stdint.h>

namespace b200q {

constexpr int KQ_MAX_SEG = 8;

struct KqSegment {
    const uint8_t * w;    // [m][row_bytes] quantized rows, contiguous, 16-byte aligned
    float *         dst;  // [m] f32
    int             m;
    int             stage0;  // first stage index of this segment (stages never straddle segments)
};

struct KqParams {
    KqSegment     seg[KQ_MAX_SEG];
    int           nseg;
    const float * act;  // [k] f32 (16-byte aligned)
    int           k, nsb;        // row length; units per row: 256-element super-blocks (K-quants) | 32-element blocks
    int           row_bytes;     // nsb * block bytes (K-quants: a multiple of 16; 32-element block formats: any)
    int           lpr;           // lanes per row: 8, 16 or 32
    int           rows_per_stage;  // 32 / lpr: one stage = one warp pass
    int           total_stages;
    int           nstages;       // ring depth
    int           stage_bytes;   // rows_per_stage * row_bytes + alignment lead / over-read, rounded up to 128
    // shared-memory offsets
    int           off_q, off_dx, off_bs16, off_bs32, off_stage;
    int           early_prefetch;  // weights immutable: stream them before griddepcontrol.wait
    // row split: every result is also stored into the same place of the dst buffer of each peer GPU (peer_delta = the
    // byte distance from this GPU's buffer to the peer's mapping of its own) — the all-gather is the kernel's epilogue
    int             n_peers;
    long long       peer_delta[7];
    // MUL_MAT_ID decode (one token): segment g multiplies expert ids[g]'s matrix; every segment may have its own
    // activation vector (b->ne[1] == n_used) with its own quantized image in shared memory
    const int32_t * ids;           // nullptr: plain segments
    long long       expert_stride; // bytes between expert matrices (seg[g].w holds the base of expert 0)
    long long       act_seg_stride;  // floats between the segments' activation vectors (0: shared)
    const float *   bias;          // MoE only, may be nullptr: dst += bias[ids[g] * bias_stride + row] (the ADD_ID behind the matmul)
    long long       bias_stride;   // floats between the experts' bias rows
    int             act_smem_stride; // bytes between the segments' quantized activation images (0: shared)
};

typedef void (*kq_kernel_t)(const KqParams);
kq_kernel_t kq_kernel_for(int type, bool moe = false);
int         kq_max_segments(bool moe);
void        kq_set_hang_buffer(unsigned long long * host_mapped);  // development aid
int         kq_threads();
int         kq_consumer_warps();

}  // namespace b200q
