// gemv_list.cuh — launch parameters of the persistent decode kernel (gemv_list.cu): ONE launch walks a whole list of
// n = 1 matmuls (a decode token's layers), the weight ring streams across matrix boundaries, true data dependencies
// are device-side barrier points.
#pragma once
#include <stdint.h>

namespace b200q {

// geometry rule (depends on (type, k) only, so results do not depend on how a list is composed or sharded):
// a stage is `rps` whole rows (rps a power of two, <= 32) with rps * row_bytes <= LIST_STAGE_PAYLOAD, or — for rows
// longer than that — one k-chunk of a single row; the chunks of a row are consecutive stages of the same warp.
// K-quants with k <= LIST_REG_ACT_NSB * 256: a lane owns a super-block COLUMN (lane % lpr), keeps its quantized
// activations in registers for the whole generation and walks the rows of the stage two at a time;
// lpr = the power of two >= columns, 32 / lpr rows run side by side.  Everything else: a lane owns a row of the stage and
// walks its blocks (gemv_kq.cu's scheme), activations read from shared memory.
constexpr int LIST_WARPS         = 8;                         // compute warps (own the ring slots, walk the units)
constexpr int LIST_QWARPS        = 4;                         // + one warpgroup that quantizes the activations one generation ahead
constexpr int LIST_STAGE_PAYLOAD = 9216;                      // 4 rows of q4_K at k = 4096
constexpr int LIST_SLOT_BYTES    = LIST_STAGE_PAYLOAD + 128;  // payload + <= 15 bytes of alignment lead + 16 bytes of over-read
constexpr int LIST_MIN_SLOTS     = 2 * LIST_WARPS;            // every warp owns at least two ring slots
constexpr int LIST_MAX_SLOTS     = 32;
constexpr int LIST_REG_ACT_NSB   = 32;                        // columns whose activations fit one register set per lane

// One matmul of the launch.  Its units (a unit = rps rows, all k-chunks) are cut into G contiguous ranges, one per CTA:
// CTA c owns base + (c' < rem) units from c' * base + min(c', rem), c' = (c + rot) % G; rot walks on by rem from op to op,
// so the surplus units rotate over the CTAs and the whole launch is balanced to one unit.  Inside a CTA the units of
// its range go round-robin to the 8 warps, continuing where the previous op stopped.
struct ListOp {  // 48 bytes; the table lives in shared memory during the launch
    const uint8_t * w;    // [m][row_bytes] dense quantized rows, 16-byte aligned base
    float *         dst;  // [m] f32
    int32_t         m;
    int32_t         base, rem, rot;
    uint32_t        rb;     // row bytes
    uint16_t        nsb;    // units of lane work per row: 256-element super-blocks (K-quants) | 32-element blocks
    uint16_t        csb;    // lane-work units per k-chunk (== nsb when the row is one chunk)
    uint16_t        gen;    // activation generation this op reads
    uint8_t         type;   // ggml type
    uint8_t         geo;    // bits 0..3: log2(rps), bits 4..7: chunks per row - 1
    uint8_t         bb;     // block bytes
    uint8_t         pad[3];
};
static_assert(sizeof(ListOp) == 48, "ListOp is 48 bytes");

struct ListGen {         // 32 bytes: one activation vector, quantized once per CTA into shared memory
    const float * act;   // [k] f32, 16-byte aligned
    uint16_t      nsb;   // k / (256 | 32)
    uint8_t       qmode; // 0: per-256 (K-quant partners), 1: per-32, 2: per-32 interleaved (mxfp4)
    uint8_t       wait;  // 1: barrier point before this generation (every earlier op complete on every CTA and GPU)
    uint16_t      bp;    // 1-based index of that barrier point within the launch
    uint16_t      xform; // 0: the vector is act; 1: silu(act[i]) * act2[i] (B200Q_LIST_SWIGLU)
    const float * act2;
    uint64_t      pad;
};
static_assert(sizeof(ListGen) == 32, "ListGen is 32 bytes");

// words of the 256-byte sync block (device memory, zero-initialised once)
// PROGRESS / EXIT: CTAs of this launch that passed the current barrier point / left (re-armed by the last one);
// ARRIVE: row split — CTAs of ALL GPUs counted in here, monotonic over launches; BASE: its value when this launch started
enum : int { LIST_SYNC_PROGRESS = 0, LIST_SYNC_EXIT = 1, LIST_SYNC_BASE = 2, LIST_SYNC_ARRIVE = 4 };

struct ListParams {
    const ListOp *  ops;    // [n_ops]
    const ListGen * gens;   // [n_gens]
    int             n_ops, n_gens;
    int             n_bp;       // barrier points of this launch (the last one is the final cross-GPU one when world > 1)
    int             final_bp;   // 1: all warps pass barrier point n_bp after the last op (row split)
    int             nslots;
    // shared-memory offsets
    int             off_ops, off_gens, off_act, act_stride, off_ring;
    unsigned *      sync;       // this GPU's sync block
    // row split
    int             world, rank, n_peers;
    unsigned *      peer_sync[7];
    long long       peer_delta[7];
    // development aid (b200q_debug_list_stamps, builds with -DB200Q_LIST_TRACE): [8 warps][4096] x (tag, clock64) of CTA 0
    unsigned long long * ts;
};

typedef void (*list_kernel_t)(const ListParams);
list_kernel_t list_kernel(bool with_q5k);  // with_q5k: some op of the launch is q5_K
int           list_threads();
bool          list_type_ok(int type);

}  // namespace b200q
