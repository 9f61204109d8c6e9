// mmq.cu — prompt-processing matmul (n > 8 activation columns) for GGUF block formats on the 5th-generation
// tensor cores of sm_100a (tcgen05.mma, accumulators in TMEM, operands fed by 1-D TMA bulk copies).
//
// Replaces the reference's quantize_mmq_q8_1 + mul_mat_q pair (ggml-cuda/quantize.cu:50-146, mmq.cuh:3134-3370,
// tile loaders mmq.cuh:277-336 q4_0, :637-697 q8_0, :699-762 mxfp4, :1726-1833 q4_K, :2027-2111 q6_K, stream-K fixup
// :3373-3522, host side mmq.cu:71-143).  It is NOT a port: the reference multiplies int8 tiles with mma.sync and pays an
// fp32 scale FMA per accumulator per 32-element block in registers (mmq.cuh:923-926).  On tcgen05 the accumulator lives
// in TMEM and a kind::i8 MMA consumes exactly one 32-block per instruction, so per-block scales would force a TMEM
// read-out + 2 CUDA-core ops per accumulator element per block (4x over the FP32-pipe budget at the i8 MMA rate, DESIGN.md
// "H2").  Instead the per-block scales are folded into the A operand where they are exact to 2^-9: weight tiles are
// dequantized on the fly by CUDA-core warps to bf16 (d * q computed in fp32, one rounding) into the canonical
// 128-byte-swizzled K-major UMMA layout, activations are rounded f32 -> bf16 once (no q8_1 quantization at all, which
// is MORE accurate than the reference's int8 activations), and kind::f16 MMAs accumulate the whole K extent in fp32 in
// TMEM with no per-block epilogue.
//
// CTA = 26 warps, one CTA per SM, persistent over work units of 128 weight rows x 512 activation columns (two 128x256
// fp32 accumulators fill the 512 TMEM columns and share ONE dequantized A chunk; 128x256 units, double-buffered across
// tiles, when that fills the SMs better; the k range halved across two CTAs + order-independent atomics for m <= 4096):
//   warps 0-3   epilogue: tcgen05.ld of the finished accumulators (TMEM lane = weight row, column = token) and
//               coalesced stores to dst (ggml layout: rows contiguous); MoE: scatter through the expert's row list
//   warps 4-7   W producers: coalesced 16-byte cp.async of the raw quantized bytes of 128/256 k-elements per row,
//               16-byte-aligned windows around arbitrarily aligned 17/18/34/210-byte blocks
//   warp  8     B producer: one bulk copy per 64-wide k-chunk of the bf16 activation tiles (both column tiles, adjacent
//               in the workspace), which mmq_pack_act has already written in the exact swizzled shared-memory image
//   warp  9     MMA issuer: the WHOLE warp runs the loop (warp-uniform control flow keeps descriptors in uniform
//               registers), one elected lane issues 8 x tcgen05.mma (M=128, N<=256, K=16) per k-chunk and the commits
//   warps 10-25 dequantizers, two groups on alternate k-chunks: thread = (row, 32-element block); raw bytes -> fp32
//               (PRMT into the mantissa of 2^23, no I2F) -> bf16 -> 4 x STS.128 under the 128B swizzle,
//               fence.proxy.async, arrive
// MUL_MAT_ID for many tokens runs through the same kernel (moe_route / moe_pack_act below).
#include "mmq.cuh"

#include "../../include/b200q.h"
#include "blocks.cuh"
#include "sm100.cuh"

#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdlib>
#include <mutex>
#include <map>
#include <set>
#include <tuple>

namespace b200q {

namespace {

using namespace sm100;

// development aids (B200Q_MMQ_DBG knock-out flags, clock stamps) are compiled in only with -DB200Q_MMQ_DEVTOOLS=1: in the
// one-warp issue loops every predicated-off instruction still costs issue slots
#ifndef B200Q_MMQ_DEVTOOLS
#define B200Q_MMQ_DEVTOOLS 0
#endif
constexpr bool DEV = B200Q_MMQ_DEVTOOLS != 0;
#define DEVF (DEV ? p.dbg_flags : 0)

constexpr int TILE_M  = 128;  // weight rows per tile (UMMA M)
constexpr int TILE_N  = 256;  // activation columns per tile (UMMA N, <= 256)
constexpr int CHUNK_K = 64;   // bf16 elements per 128-byte swizzled row
constexpr int A_STAGES = 3;      // dequantized weight chunks (bf16)
constexpr int MAX_B_STAGES = 4;  // packed activation chunks (bf16); 3 when the W ring needs the room
constexpr int MAX_W_STAGES = 4;
constexpr int A_STAGE_BYTES = TILE_M * 128;  // 16 KB
constexpr int B_STAGE_BYTES = TILE_N * 128;  // 32 KB
constexpr int N_EPI_WARPS = 4, N_WP_WARPS = 4, N_DQ_WARPS = 16, N_DQ_GROUPS = 2;  // dequant groups take alternate k-chunks
constexpr int WARP_WPROD = 4, WARP_BPROD = 8, WARP_MMA = 9, WARP_DQ0 = 10;
constexpr int NTHREADS = (WARP_DQ0 + N_DQ_WARPS) * 32;  // 832
constexpr int TMEM_COLS = 512;

// barrier slots (8 bytes each) inside the barrier block
enum {
    BAR_W_FULL  = 0,                           // [MAX_W_STAGES] count 128 (cp.async.mbarrier.arrive.noinc per producer thread)
    BAR_W_EMPTY = BAR_W_FULL + MAX_W_STAGES,   // [MAX_W_STAGES] count N_DQ_WARPS
    BAR_A_FULL  = BAR_W_EMPTY + MAX_W_STAGES,  // [A_STAGES] count N_DQ_WARPS / N_DQ_GROUPS
    BAR_A_EMPTY = BAR_A_FULL + A_STAGES,       // [A_STAGES] count 1 (tcgen05.commit after the chunk's last MMA)
    BAR_B_FULL  = BAR_A_EMPTY + A_STAGES,      // [MAX_B_STAGES] count 1 + tx bytes
    BAR_B_EMPTY = BAR_B_FULL + MAX_B_STAGES,   // [MAX_B_STAGES] count 1 (tcgen05.commit)
    BAR_ACC_FULL = BAR_B_EMPTY + MAX_B_STAGES,  // [2] count 1 (tcgen05.commit)
    BAR_ACC_EMPTY = BAR_ACC_FULL + 2,         // [2] count N_EPI_WARPS
    BAR_COUNT = BAR_ACC_EMPTY + 2,
};

struct MmqParams {
    const uint8_t * w;
    const uint8_t * actp;  // packed bf16 activation tiles (mmq_pack_act)
    float *         dst;
    int   k, m, n;
    int   ne12, r2, r3, nbatch;
    long long w_nb1, w_nb2, w_nb3;
    long long d_nb1, d_nb2, d_nb3;
    int   m_tiles, n_tiles, k_chunks, w_stages_per_tile;
    int   w_stage_k;     // k-elements of raw weights per W stage (128 for 32-element block types, 256 for K-quants)
    int   w_stages;      // W ring depth (2..MAX_W_STAGES, what fits in shared memory)
    int   b_stages;      // B ring depth in stages of n_sub x 32 KB (the column tiles of one k-chunk travel as ONE stage)
    int   b_stage_bytes; // n_sub * B_STAGE_BYTES
    int   n_super;       // column super-tiles per row tile: two 256-column accumulators share ONE dequantized A chunk
    int   n_sub;         // column tiles per super-tile (2, or 1 when there are too few tiles to fill the GPU otherwise)
    int   cluster;       // CTAs per cluster (1, 2 or 4): they take adjacent row tiles and share every B chunk by multicast
    int   m_groups;      // ceil(m_tiles / cluster)
    // MoE (MUL_MAT_ID) mode: the column tiles are per-expert lists of (token, slot) rows built on the device (moe_route)
    const int * moe;         // header: [0] total column tiles, then the tables below (nullptr: plain MUL_MAT)
    const int * moe_tile_expert, * moe_tile_col0, * moe_count, * moe_offset, * moe_rows;
    int   n_used;
    long long w_nb_expert;
    int   ksplit;        // 1 or 2: k-halves of one tile go to different CTAs and are summed with fp32 atomics into a zeroed dst
                         // (two addends on zero: the result does not depend on the arrival order)
    int   ws_per_split;  // W stages per k-split
    int   piece_bytes;   // raw bytes of w_stage_k elements of one row
    int   blk_bytes;     // bytes per 32 elements (32-block types) — used for the ragged last stage
    int   w_pitch;       // shared-memory pitch of one row piece
    int   w_stage_bytes; // TILE_M * w_pitch
    uint32_t off_b, off_w, off_bar;  // shared-memory offsets from the 1024-aligned base
    int dbg_flags;                   // development aid (timing experiments only): 1 = skip W loads, 2 = skip dequant math, 4 = skip B loads
    unsigned long long * dbg;        // development aid: CTA 0 stores clock64 stamps [chunk][8] for its first 256 chunks
};
__device__ __forceinline__ void dbg_phase(const MmqParams & p, int slot) {  // globaltimer (ns) stamps of CTA 0's phases
    if (DEV && p.dbg && blockIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.dbg[255 * 8 + slot] = t;
    }
}
__device__ __forceinline__ void dbg_stamp(const MmqParams & p, int g, int slot) {
    if (DEV && p.dbg && blockIdx.x == 0 && g < 256) p.dbg[g * 8 + slot] = (unsigned long long) clock64();
}

struct PackParams {
    const float * act;
    uint8_t *     out;
    float *       zero_dst;  // != nullptr: also clear dst [nbatch][n][m] (k-split results are accumulated with atomics)
    int           m;
    long long     d_nb1, d_nb2, d_nb3;
    int   k, n, ne12, nbatch, n_tiles, k_chunks;
    int   n_sub, n_super;  // blocks are ordered (batch, super-tile, k-chunk, column tile of the super-tile)
    long long a_nb1, a_nb2, a_nb3;
    int   vec4;
    int   zero_only;  // the packed image is already in the workspace (same activations as the previous call): only clear dst
};

// f32 activations [n][k] (byte strides) -> bf16 tiles in the shared-memory image the MMA reads:
//   block (batch, n_tile, k_chunk) = TILE_N rows x 128 bytes; 16-byte chunk c of row j sits at chunk (c ^ (j & 7)).
// One thread = one 16-byte chunk (8 elements).  Rows >= n are written as zeros.
__global__ void __launch_bounds__(256) mmq_pack_act(const PackParams p) {
    if (p.zero_dst) {
        const long long cells = (long long) p.nbatch * p.n * p.m;
        for (long long idx = (long long) blockIdx.x * blockDim.x + threadIdx.x; idx < cells; idx += (long long) gridDim.x * blockDim.x) {
            const int       i   = (int) (idx % p.m);
            const long long r   = idx / p.m;
            const int       col = (int) (r % p.n), b = (int) (r / p.n);
            const int       i13 = b / p.ne12, i12 = b - i13 * p.ne12;
            *(float *) ((char *) p.zero_dst + (long long) i * 4 + col * p.d_nb1 + i12 * p.d_nb2 + i13 * p.d_nb3) = 0.0f;
        }
    }
    if (p.zero_only) return;
    // one CTA per 32 KB block (batch, super-tile, k-chunk, column tile): the block's coordinates are decoded once (uniform),
    // 8 consecutive threads read the 256 contiguous bytes of one row's k-chunk and write its eight swizzled 16-byte chunks
    // (one thread per chunk with five 64-bit divisions each made the pack 12 us for 512 x 4096 — 15 % of a prompt step)
    const long long n_blocks = (long long) p.nbatch * p.n_super * p.n_sub * p.k_chunks;
    for (long long blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        const int       sn   = (int) (blk % p.n_sub);
        const long long r1   = blk / p.n_sub;
        const int       kc   = (int) (r1 % p.k_chunks);
        const long long rest = r1 / p.k_chunks;
        const int       nt   = (int) (rest % p.n_super) * p.n_sub + sn;
        const int       b    = (int) (rest / p.n_super);
        const int       i13  = b / p.ne12, i12 = b - i13 * p.ne12;
        const char *    base = (const char *) p.act + (long long) i12 * p.a_nb2 + (long long) i13 * p.a_nb3 + (long long) kc * CHUNK_K * 4;
        uint8_t *       out  = p.out + blk * (long long) B_STAGE_BYTES;
#pragma unroll 2
        for (int idx = threadIdx.x; idx < TILE_N * 8; idx += 256) {
            const int c = idx & 7, j = idx >> 3;
            const int col = nt * TILE_N + j;  // >= n for the padding tile of an odd tile count
            uint4     o   = make_uint4(0, 0, 0, 0);
            if (col < p.n) {
                const float * src = (const float *) (base + (long long) col * p.a_nb1) + c * 8;
                float v[8];
                if (p.vec4) {
                    const float4 x = __ldg((const float4 *) src), y = __ldg((const float4 *) src + 1);
                    v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w; v[4] = y.x; v[5] = y.y; v[6] = y.z; v[7] = y.w;
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = __ldg(src + i);
                }
                __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
                __nv_bfloat162 h2 = __floats2bfloat162_rn(v[4], v[5]), h3 = __floats2bfloat162_rn(v[6], v[7]);
                o.x = *(uint32_t *) &h0; o.y = *(uint32_t *) &h1; o.z = *(uint32_t *) &h2; o.w = *(uint32_t *) &h3;
            }
            *(uint4 *) (out + (long long) j * 128 + ((c ^ (j & 7)) << 4)) = o;
        }
    }
}

// ---- dequantizers: 32 consecutive k-elements of one row piece -> fp32 ------------------------------------------
// `piece` is the SHARED-WINDOW address (32-bit) of the first byte of the row's W-stage piece (a block / super-block
// boundary, ANY byte alignment); b32 = index of the 32-element block inside the piece (0..7).  All loads are explicit
// ld.shared (a generic pointer here costs an LD.E through the L1TEX path per access: measured 3x slower).
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
// 32-bit little-endian word at any byte address: the two aligned words that cover it + funnel shift
__device__ __forceinline__ uint32_t ldu32(uint32_t a) {
    const uint32_t w = a & ~3u;
    return __funnelshift_r(lds_u32(w), lds_u32(w + 4), (int) (a & 3u) * 8);
}
// N consecutive words from any byte address: N+1 aligned loads
template <int N> __device__ __forceinline__ void ldu32_n(uint32_t a, uint32_t (&out)[N]) {
    const uint32_t w  = a & ~3u;
    const int      sh = (int) (a & 3u) * 8;
    uint32_t       prev = lds_u32(w);
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const uint32_t next = lds_u32(w + 4 * (i + 1));
        out[i]              = __funnelshift_r(prev, next, sh);
        prev                = next;
    }
}
__device__ __forceinline__ float ldh(uint32_t a) {  // fp16 at a 1-byte aligned address
    return half_bits_to_float((uint16_t) (ldu32(a) & 0xffffu));
}

// int -> float without the I2F instruction (quarter-rate XU pipe): byte j of `word` dropped into the mantissa of 2^23
// gives the float 8388608 + byte exactly; one FADD / FFMA on the FP32 pipe removes the bias.
constexpr float MAGIC = 8388608.0f;
__device__ __forceinline__ float magic_u8(uint32_t word, int j) {
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u + j));
}

template <int T> struct Dq {
    // generic (slow) path: the bit-exact scalar decoder of blocks.cuh on a generic pointer
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const uint8_t * gp = (const uint8_t *) __cvta_shared_to_generic((size_t) piece);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = decode_elem<T>(gp, (int64_t) b32 * 32 + i);
    }
};
template <> struct Dq<T_Q8_0> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const uint32_t b = piece + b32 * 34;
        uint32_t       w[9];  // [d | q0 q1] ... [q30 q31 | next]
        ldu32_n<9>(b, w);
        const float d  = half_bits_to_float((uint16_t) (w[0] & 0xffffu));
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t q = __funnelshift_r(w[i], w[i + 1], 16) ^ 0x80808080u;  // q + 128 per byte
#pragma unroll
            for (int j = 0; j < 4; ++j) v[4 * i + j] = d * (magic_u8(q, j) - (MAGIC + 128.0f));
        }
    }
};
template <> struct Dq<T_Q4_0> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const uint32_t b = piece + b32 * 18;
        uint32_t       w[5];  // [d | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        ldu32_n<5>(b, w);
        const float d  = half_bits_to_float((uint16_t) (w[0] & 0xffffu));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q  = __funnelshift_r(w[i], w[i + 1], 16);
            const uint32_t lo = q & 0x0F0F0F0Fu, hi = (q >> 4) & 0x0F0F0F0Fu;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = d * (magic_u8(lo, j) - (MAGIC + 8.0f));
                v[16 + 4 * i + j] = d * (magic_u8(hi, j) - (MAGIC + 8.0f));
            }
        }
    }
};
template <> struct Dq<T_MXFP4> {
    static __device__ __forceinline__ float kv(uint32_t nib) {  // doubled E2M1 code point (ggml-common.h:1094-1096)
        const uint32_t mag = __byte_perm(0x03020100u, 0x0C080604u, nib & 7u);
        const float    f   = magic_u8(mag, 0) - MAGIC;
        return (nib & 8u) ? -f : f;
    }
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const uint32_t b = piece + b32 * 17;
        uint32_t       w[5];  // [e | qs0 qs1 qs2] ... [qs15 | next]
        ldu32_n<5>(b, w);
        const float d = e8m0_half_to_float((uint8_t) (w[0] & 0xff));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = __funnelshift_r(w[i], w[i + 1], 8);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = d * kv((q >> (8 * j)) & 0xF);
                v[16 + 4 * i + j] = d * kv((q >> (8 * j + 4)) & 0xF);
            }
        }
    }
};
template <> struct Dq<T_Q4_K> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t h[4];  // [d dmin] [scales x3]
        ldu32_n<4>(piece, h);
        const float d = half_bits_to_float((uint16_t) (h[0] & 0xffffu)), dmin = half_bits_to_float((uint16_t) (h[0] >> 16));
        // get_scale_min_k4 (ggml-quants.c:703-710) on the three scale words
        int sc, mn;
        if (b32 < 4) {
            sc = (int) ((h[1] >> (8 * b32)) & 63);
            mn = (int) ((h[2] >> (8 * b32)) & 63);
        } else {
            const int      jj = b32 - 4;
            const uint32_t hi = (h[3] >> (8 * jj)) & 0xff;
            sc                = (int) ((hi & 0xF) | (((h[1] >> (8 * jj + 6)) & 3) << 4));
            mn                = (int) ((hi >> 4) | (((h[2] >> (8 * jj + 6)) & 3) << 4));
        }
        const float dsc = d * (float) sc, dm = -dmin * (float) mn;
        const int   sh  = 4 * (b32 & 1);
        uint32_t    q[8];
        ldu32_n<8>(piece + 16 + 32 * (b32 >> 1), q);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t w = (q[i] >> sh) & 0x0F0F0F0Fu;
#pragma unroll
            for (int j = 0; j < 4; ++j) v[4 * i + j] = fmaf(magic_u8(w, j) - MAGIC, dsc, dm);
        }
    }
};
template <> struct Dq<T_Q5_K> {  // q4_K's layout with 32 bytes of fifth bits behind the scales: element l of sub-block b32 takes bit b32 of qh[l]
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t h[4];  // [d dmin] [scales x3]
        ldu32_n<4>(piece, h);
        const float d = half_bits_to_float((uint16_t) (h[0] & 0xffffu)), dmin = half_bits_to_float((uint16_t) (h[0] >> 16));
        int sc, mn;
        if (b32 < 4) {
            sc = (int) ((h[1] >> (8 * b32)) & 63);
            mn = (int) ((h[2] >> (8 * b32)) & 63);
        } else {
            const int      jj = b32 - 4;
            const uint32_t hi = (h[3] >> (8 * jj)) & 0xff;
            sc                = (int) ((hi & 0xF) | (((h[1] >> (8 * jj + 6)) & 3) << 4));
            mn                = (int) ((hi >> 4) | (((h[2] >> (8 * jj + 6)) & 3) << 4));
        }
        const float dsc = d * (float) sc, dm = -dmin * (float) mn;
        const int   sh  = 4 * (b32 & 1);
        uint32_t    q[8], fh[8];
        ldu32_n<8>(piece + 48 + 32 * (b32 >> 1), q);
        ldu32_n<8>(piece + 16, fh);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t w = ((q[i] >> sh) & 0x0F0F0F0Fu) | (((fh[i] >> b32) & 0x01010101u) << 4);
#pragma unroll
            for (int j = 0; j < 4; ++j) v[4 * i + j] = fmaf(magic_u8(w, j) - MAGIC, dsc, dm);
        }
    }
};
// the 32-element formats with a minimum and / or a fifth bit (dequantize_row_q4_1 / q5_0 / q5_1, ggml-quants.c): bit l of the
// 32-bit qh belongs to element l; four of them are spread to bit 4 of four bytes by one multiplication (bit k lands on 8 k)
__device__ __forceinline__ uint32_t spread4(uint32_t qh, int first) { return ((((qh >> first) & 0xFu) * 0x00204081u) & 0x01010101u) << 4; }
template <> struct Dq<T_Q4_1> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[5];  // [d m] [qs x4]
        ldu32_n<5>(piece + b32 * 20, w);
        const float d = half_bits_to_float((uint16_t) (w[0] & 0xffffu)), m = half_bits_to_float((uint16_t) (w[0] >> 16));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t lo = w[1 + i] & 0x0F0F0F0Fu, hi = (w[1 + i] >> 4) & 0x0F0F0F0Fu;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = fmaf(magic_u8(lo, j) - MAGIC, d, m);
                v[16 + 4 * i + j] = fmaf(magic_u8(hi, j) - MAGIC, d, m);
            }
        }
    }
};
template <> struct Dq<T_Q5_0> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[6];  // [d | qh0 qh1] [qh2 qh3 | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        ldu32_n<6>(piece + b32 * 22, w);
        const float    d  = half_bits_to_float((uint16_t) (w[0] & 0xffffu));
        const uint32_t qh = __funnelshift_r(w[0], w[1], 16);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q  = __funnelshift_r(w[1 + i], w[2 + i], 16);
            const uint32_t lo = (q & 0x0F0F0F0Fu) | spread4(qh, 4 * i), hi = ((q >> 4) & 0x0F0F0F0Fu) | spread4(qh, 16 + 4 * i);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = d * (magic_u8(lo, j) - (MAGIC + 16.0f));
                v[16 + 4 * i + j] = d * (magic_u8(hi, j) - (MAGIC + 16.0f));
            }
        }
    }
};
template <> struct Dq<T_Q5_1> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[6];  // [d m] [qh] [qs x4]
        ldu32_n<6>(piece + b32 * 24, w);
        const float    d  = half_bits_to_float((uint16_t) (w[0] & 0xffffu)), m = half_bits_to_float((uint16_t) (w[0] >> 16));
        const uint32_t qh = w[1];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t lo = (w[2 + i] & 0x0F0F0F0Fu) | spread4(qh, 4 * i), hi = ((w[2 + i] >> 4) & 0x0F0F0F0Fu) | spread4(qh, 16 + 4 * i);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = fmaf(magic_u8(lo, j) - MAGIC, d, m);
                v[16 + 4 * i + j] = fmaf(magic_u8(hi, j) - MAGIC, d, m);
            }
        }
    }
};
// kvalues_iq4nl (ggml-common.h:1088-1090) + 128 as a 16-byte table in four words; four nibbles (one per byte of `nibs`) are looked
// up with three PRMTs: entries 0..7 and 8..15 by the low three bits, then bit 3 picks per byte
__device__ __forceinline__ uint32_t iq4_lut4(uint32_t nibs) {
    const uint32_t sel = (nibs & 0xFu) | ((nibs >> 4) & 0xF0u) | ((nibs >> 8) & 0xF00u) | ((nibs >> 12) & 0xF000u);
    const uint32_t lo  = __byte_perm(0x3F2D1801u, 0x766A5D4Fu, sel & 0x7777u);
    const uint32_t hi  = __byte_perm(0xA6998D81u, 0xF1D9C5B5u, sel & 0x7777u);
    return __byte_perm(lo, hi, 0x3210u | ((sel & 0x8888u) >> 1));
}
template <> struct Dq<T_IQ4_NL> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[5];  // [d | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        ldu32_n<5>(piece + b32 * 18, w);
        const float d = half_bits_to_float((uint16_t) (w[0] & 0xffffu));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q  = __funnelshift_r(w[i], w[i + 1], 16);
            const uint32_t lo = iq4_lut4(q & 0x0F0F0F0Fu), hi = iq4_lut4((q >> 4) & 0x0F0F0F0Fu);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = d * (magic_u8(lo, j) - (MAGIC + 128.0f));
                v[16 + 4 * i + j] = d * (magic_u8(hi, j) - (MAGIC + 128.0f));
            }
        }
    }
};
template <> struct Dq<T_IQ4_XS> {  // d:0 scales_h:2 scales_l:4 qs:8 (136 bytes per 256 values); sub-block b32 has a 6-bit scale
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t h[2], q[4];
        ldu32_n<2>(piece, h);  // [d | scales_h] [scales_l x4]
        ldu32_n<4>(piece + 8 + 16 * b32, q);
        const int   ls = (int) (((h[1] >> (4 * b32)) & 0xFu) | ((((h[0] >> 16) >> (2 * b32)) & 3u) << 4));
        const float dl = half_bits_to_float((uint16_t) (h[0] & 0xffffu)) * (float) (ls - 32);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t lo = iq4_lut4(q[i] & 0x0F0F0F0Fu), hi = iq4_lut4((q[i] >> 4) & 0x0F0F0F0Fu);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v[4 * i + j]      = dl * (magic_u8(lo, j) - (MAGIC + 128.0f));
                v[16 + 4 * i + j] = dl * (magic_u8(hi, j) - (MAGIC + 128.0f));
            }
        }
    }
};
// q2_K (scales:0 qs:16 d:80 dmin:82) and q3_K (hmask:0 qs:32 scales:96 d:108): 32 values = 2-bit field j of 32 consecutive qs
// bytes of 128-value half n, two 16-value halves with a scale each (dequantize_row_q2_K / _q3_K, ggml-quants.c:784-814, 1128-1176)
template <> struct Dq<T_Q2_K> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const int n = b32 >> 2, j = b32 & 3;
        uint32_t  q[8];
        ldu32_n<8>(piece + 16 + 32 * n, q);
        const uint32_t dd = ldu32(piece + 80), sc = ldu32(piece + 8 * n + 2 * j);  // [d dmin]; scale bytes of the two halves
        const float    d = half_bits_to_float((uint16_t) (dd & 0xffffu)), dmin = half_bits_to_float((uint16_t) (dd >> 16));
        const float    dl0 = d * (float) (sc & 0xFu), ml0 = -dmin * (float) ((sc >> 4) & 0xFu);
        const float    dl1 = d * (float) ((sc >> 8) & 0xFu), ml1 = -dmin * (float) ((sc >> 12) & 0xFu);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t w  = (q[i] >> (2 * j)) & 0x03030303u;
            const float    dl = i < 4 ? dl0 : dl1, ml = i < 4 ? ml0 : ml1;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) v[4 * i + jj] = fmaf(magic_u8(w, jj) - MAGIC, dl, ml);
        }
    }
};
template <> struct Dq<T_Q3_K> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const int n = b32 >> 2, j = b32 & 3;
        uint32_t  q[8], hm[8], scw[3];
        ldu32_n<8>(piece + 32 + 32 * n, q);
        ldu32_n<8>(piece, hm);
        ldu32_n<3>(piece + 96, scw);
        const float d = ldh(piece + 108);
        auto byte = [&](int k) { return ((k < 4 ? scw[0] : k < 8 ? scw[1] : scw[2]) >> (8 * (k & 3))) & 0xffu; };
        auto scale = [&](int s) {  // the aux[] shuffle of dequantize_row_q3_K: low 4 bits in bytes 0..7, high 2 bits in bytes 8..11
            const uint32_t lo = s < 8 ? (byte(s) & 0xFu) : (byte(s - 8) >> 4);
            const uint32_t hi = (byte(8 + (s & 3)) >> (2 * (s >> 2))) & 3u;
            return (int) (lo | (hi << 4)) - 32;
        };
        const float dl0 = d * (float) scale(8 * n + 2 * j), dl1 = d * (float) scale(8 * n + 2 * j + 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t w  = ((q[i] >> (2 * j)) & 0x03030303u) | (((hm[i] >> (4 * n + j)) & 0x01010101u) << 2);
            const float    dl = i < 4 ? dl0 : dl1;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) v[4 * i + jj] = dl * (magic_u8(w, jj) - (MAGIC + 4.0f));
        }
    }
};
template <> struct Dq<T_Q6_K> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        const int n128 = b32 >> 2, quad = b32 & 3;
        const int shl  = 4 * (quad >> 1), shh = 2 * quad;
        uint32_t  ql[8], qh[8];
        ldu32_n<8>(piece + 64 * n128 + 32 * (quad & 1), ql);
        ldu32_n<8>(piece + 128 + 32 * n128, qh);
        const float    d  = ldh(piece + 208);
        const uint32_t sw = ldu32(piece + 192 + 8 * n128 + 2 * quad);
        const float    s0 = d * (float) (int8_t) (sw & 0xff);
        const float    s1 = d * (float) (int8_t) ((sw >> 8) & 0xff);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t q = ((ql[i] >> shl) & 0x0F0F0F0Fu) | (((qh[i] >> shh) & 0x03030303u) << 4);
            const float    s = i < 4 ? s0 : s1;
#pragma unroll
            for (int j = 0; j < 4; ++j) v[4 * i + j] = s * (magic_u8(q, j) - (MAGIC + 32.0f));
        }
    }
};

// unquantized weights (f16 / bf16 / f32 tensors of a model, the reference's mmvf.cu / mmf.cu / cuBLAS cases): 32 values per
// "block", rounded to bf16 like every other operand of this kernel
template <> struct Dq<T_F16> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[16];
        ldu32_n<16>(piece + b32 * 64, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float2 f = __half22float2(*(const __half2 *) &w[i]);
            v[2 * i] = f.x; v[2 * i + 1] = f.y;
        }
    }
};
template <> struct Dq<T_BF16> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[16];
        ldu32_n<16>(piece + b32 * 64, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
};
template <> struct Dq<T_F32W> {
    static __device__ __forceinline__ void run(uint32_t piece, int b32, float (&v)[32]) {
        uint32_t w[32];
        ldu32_n<32>(piece + b32 * 128, w);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(w[i]);
    }
};

extern __shared__ __align__(1024) uint8_t mmq_smem_raw[];

struct TileCoord {
    int b, mt, nt, ns;  // batch, row tile, first column tile, column tiles in this super-tile (1 or 2)
    int ws0, ws1;       // W stages [ws0, ws1) of the row pieces
    int c0, c1;         // k-chunks [c0, c1)
    int ncols[2];       // valid columns of the (up to two) column tiles
    long long w_off;    // byte offset of the unit's weight matrix (batch broadcast | expert)
    int row_base;       // MoE: index of the tile's first entry in moe_rows
};
// work unit `t` of a cluster; CTA `rank` of the cluster takes row tile group * cluster + rank, which may lie past the
// last row tile (a phantom tile: the CTA still feeds its share of B to its peers, its results are not stored)
__device__ __forceinline__ TileCoord tile_coord(const MmqParams & p, int t, int rank) {
    TileCoord c;
    if (p.moe) {  // unit = (column tile of one expert, row tile); row tiles fastest so that neighbours share the B tile
        c.mt           = t % p.m_tiles;
        const int ts   = t / p.m_tiles;
        const int e    = p.moe_tile_expert[ts], col0 = p.moe_tile_col0[ts];
        c.b = 0; c.nt = ts; c.ns = 1;
        c.ws0 = 0; c.ws1 = p.w_stages_per_tile; c.c0 = 0; c.c1 = p.k_chunks;
        c.ncols[0]     = min(TILE_N, p.moe_count[e] - col0);
        c.ncols[1]     = 0;
        c.w_off        = (long long) e * p.w_nb_expert;
        c.row_base     = p.moe_offset[e] + col0;
        return c;
    }
    const int ks   = t % p.ksplit;
    t /= p.ksplit;
    c.ws0          = ks * p.ws_per_split;
    c.ws1          = min(p.w_stages_per_tile, c.ws0 + p.ws_per_split);
    const int cps  = p.w_stage_k / CHUNK_K;
    c.c0           = c.ws0 * cps;
    c.c1           = min(p.k_chunks, c.ws1 * cps);
    c.nt           = (t % p.n_super) * p.n_sub;
    const int rest = t / p.n_super;
    c.mt           = (rest % p.m_groups) * p.cluster + rank;
    c.b            = rest / p.m_groups;
    c.ns           = min(p.n_sub, p.n_tiles - c.nt);
    c.ncols[0]     = min(TILE_N, p.n - c.nt * TILE_N);
    c.ncols[1]     = max(0, min(TILE_N, p.n - (c.nt + 1) * TILE_N));
    const int i13  = c.b / p.ne12, i12 = c.b - i13 * p.ne12;
    c.w_off        = (long long) (i12 / p.r2) * p.w_nb2 + (long long) (i13 / p.r3) * p.w_nb3;
    c.row_base     = 0;
    return c;
}
// bytes of the raw row piece of W stage `ws` (the last stage of a row may be ragged for 32-element block types)
__device__ __forceinline__ int piece_bytes_of(const MmqParams & p, int ws) {
    const int elems = min(p.w_stage_k, p.k - ws * p.w_stage_k);
    return elems == p.w_stage_k ? p.piece_bytes : (elems >> 5) * p.blk_bytes;
}

template <int T> __global__ void __launch_bounds__(NTHREADS, 1) mmq_kernel(const MmqParams p) {
    // 1024-byte aligned base (128B-swizzle atoms)
    const uint32_t raw_addr = smem_u32(mmq_smem_raw);
    const uint32_t base     = (raw_addr + 1023u) & ~1023u;
    uint8_t *      sm       = mmq_smem_raw + (base - raw_addr);
    const uint32_t bars     = base + p.off_bar;
    volatile uint32_t * tmem_slot = (volatile uint32_t *) (sm + p.off_bar + BAR_COUNT * 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int  num_tiles = p.moe ? p.moe[0] * p.m_tiles : p.nbatch * p.m_groups * p.n_super * p.ksplit;  // work units of a cluster
    const bool dual      = p.n_sub > 1;  // both accumulators belong to one super-tile (else: double-buffered tiles)
    const int  crank     = p.cluster > 1 ? (int) cluster_ctarank() : 0;
    const int  cid       = blockIdx.x / p.cluster, ncl = gridDim.x / p.cluster;
    const uint16_t cmask = (uint16_t) ((1u << p.cluster) - 1u);
    if (threadIdx.x == 0) dbg_phase(p, 0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.w_stages; ++s) {
            mbar_init(bars + 8 * (BAR_W_FULL + s), N_WP_WARPS * 32);
            mbar_init(bars + 8 * (BAR_W_EMPTY + s), N_DQ_WARPS);
        }
        for (int s = 0; s < A_STAGES; ++s) {
            mbar_init(bars + 8 * (BAR_A_FULL + s), N_DQ_WARPS / N_DQ_GROUPS);
            mbar_init(bars + 8 * (BAR_A_EMPTY + s), 1);
        }
        for (int s = 0; s < p.b_stages; ++s) {
            mbar_init(bars + 8 * (BAR_B_FULL + s), 1);
            mbar_init(bars + 8 * (BAR_B_EMPTY + s), p.cluster);  // every CTA of the cluster releases every copy
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(bars + 8 * (BAR_ACC_FULL + s), 1);
            mbar_init(bars + 8 * (BAR_ACC_EMPTY + s), N_EPI_WARPS);
        }
        mbar_fence_init();
    }
    if (warp == WARP_MMA) {
        tmem_alloc(smem_u32((const void *) tmem_slot), TMEM_COLS);
    }
    tc_fence_before();
    __syncthreads();
    if (p.cluster > 1) cluster_sync();  // peers' barriers are initialised before anything is multicast to them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (threadIdx.x == 0) dbg_phase(p, 1);

    if (warp >= WARP_WPROD && warp < WARP_WPROD + N_WP_WARPS) {
        if (!(DEVF & 64)) {
        // ------------------------------------------------------------------ raw weight rows -> W ring
        // coalesced 16-byte cp.async (LDGSTS): a half-warp covers the aligned window around one row piece (<= 16 chunks
        // of 16 bytes), so one warp instruction touches 2 rows x <= 3 cache lines instead of 32 rows
        const int wp  = warp - WARP_WPROD;
        const int l16 = lane & 15, hf = lane >> 4;
        const int cpr = p.w_pitch >> 4;  // 16-byte chunks copied per row (the whole pitch; over-read stays inside the tensor + pad)
        int wsi = 0;
        for (int t = cid; t < num_tiles; t += ncl) {
            const TileCoord tc  = tile_coord(p, t, crank);
            const uint8_t * wb  = p.w + tc.w_off;
            for (int ws = tc.ws0; ws < tc.ws1; ++ws, ++wsi) {
                const int slot = wsi % p.w_stages;
                mbar_wait(bars + 8 * (BAR_W_EMPTY + slot), ((wsi / p.w_stages) & 1) ^ 1);
                const uint32_t sbase = base + p.off_w + (uint32_t) slot * p.w_stage_bytes + 16u * l16;
                if (l16 < cpr && !((DEVF & 1) && wsi >= p.w_stages)) {
#pragma unroll 4
                    for (int it = 0; it < TILE_M / 8; ++it) {
                        const int       row = it * 8 + wp * 2 + hf;
                        const uint8_t * a   = wb + (long long) min(tc.mt * TILE_M + row, p.m - 1) * p.w_nb1 + (long long) ws * p.piece_bytes;
                        const uint8_t * src = (const uint8_t *) ((uintptr_t) a & ~(uintptr_t) 15) + 16 * l16;
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sbase + (uint32_t) row * p.w_pitch), "l"(src) : "memory");
                    }
                }
                asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bars + 8 * (BAR_W_FULL + slot)) : "memory");
            }
        }
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    } else if (warp == WARP_BPROD) {
        if (!(DEVF & 128))
        // ------------------------------------------------------------------ packed activation tiles -> B ring
        {
            // (the whole warp runs the loop, one elected lane issues; ring slot / phase are running counters and the source
            // pointer advances by one stage per chunk: runtime div / mod here cost more than the copy they issue)
            int      slot = 0, issued = 0;
            uint32_t ph   = 0;
            for (int t = cid; t < num_tiles; t += ncl) {
                const TileCoord tc = tile_coord(p, t, crank);
                // the column tiles of one k-chunk are adjacent in the packed workspace: one copy per chunk
                const int ncols_last = tc.ncols[tc.ns - 1];
                uint32_t  bytes      = (uint32_t) (tc.ns - 1) * B_STAGE_BYTES + (uint32_t) ((ncols_last + 15) & ~15) * 128u;
                const uint8_t * src  = p.moe ? p.actp + ((long long) tc.nt * p.k_chunks) * B_STAGE_BYTES
                                             : p.actp + (((long long) (tc.b * p.n_super + tc.nt / p.n_sub) * p.k_chunks + tc.c0) * p.n_sub) * B_STAGE_BYTES;
                uint32_t doff = 0;
                if (p.cluster > 1) {  // this CTA fetches 1/cluster of every chunk and multicasts it to its peers
                    bytes /= (uint32_t) p.cluster;  // a multiple of 2048 before: the slices stay 16-byte multiples
                    doff = bytes * (uint32_t) crank;
                    src += doff;
                }
                for (int c = tc.c0; c < tc.c1; ++c) {
                    mbar_wait(bars + 8 * (BAR_B_EMPTY + slot), ph ^ 1);
                    const uint32_t full = bars + 8 * (BAR_B_FULL + slot);
                    const uint32_t dst  = base + p.off_b + (uint32_t) slot * (uint32_t) p.b_stage_bytes + doff;
                    if (elect_one()) {
                        if (!(DEVF & 1024)) dbg_stamp(p, c, 7);
                        if ((DEVF & 4) && issued >= p.b_stages) {
                            mbar_arrive(full);
                        } else if (p.cluster == 1) {
                            mbar_expect_tx(full, bytes);
                            bulk_g2s(dst, src, bytes, full);
                        } else {
                            mbar_expect_tx(full, bytes * (uint32_t) p.cluster);
                            bulk_g2s_multicast(dst, src, bytes, full, cmask);
                        }
                    }
                    __syncwarp();
                    src += p.b_stage_bytes;
                    ++issued;
                    if (++slot == p.b_stages) {
                        slot = 0;
                        ph ^= 1;
                    }
                }
            }
        }
    } else if (warp == WARP_MMA) {
        // ------------------------------------------------------------------ tensor-core issue (whole warp loops, one lane issues)
        {
            int      g = 0, ti = 0, aslot = 0, bslot = 0;
            uint32_t a_ph = 0, b_ph = 0;
            // shared-memory descriptors advance with the ring slots (address field = bytes >> 4)
            const uint64_t adesc0 = umma_desc_k_sw128(base), bdesc0 = umma_desc_k_sw128(base + p.off_b);
            uint64_t       adesc = adesc0, bdesc = bdesc0;
            for (int t = cid; t < num_tiles; t += ncl, ++ti) {
                const TileCoord tc  = tile_coord(p, t, crank);
                const int       abi = dual ? 0 : (ti & 1);              // accumulator barrier pair
                const uint32_t  aph = dual ? (ti & 1) : ((ti >> 1) & 1);
                uint32_t idesc[2], d_tmem[2];
                for (int sn = 0; sn < 2; ++sn) {
                    idesc[sn]  = umma_idesc_bf16_f32(TILE_M, (max(tc.ncols[sn], 1) + 15) & ~15);
                    d_tmem[sn] = tmem_base + (uint32_t) (dual ? sn : abi) * TILE_N;
                }
                mbar_wait(bars + 8 * (BAR_ACC_EMPTY + abi), aph ^ 1);
                tc_fence_after();
                // The commits of chunk c are issued right in FRONT of the MMAs of chunk c + 1 (and after the last chunk of the
                // unit), not behind their own MMAs: the tcgen05 queue is in order and only a few entries deep, so a commit
                // behind eight 128-cycle MMAs blocks the issuing warp until most of them have run, and the warp's own work
                // for the next chunk (ring counters, two barrier waits, descriptors: ~500 cycles) then starts with a nearly
                // empty queue — the tensor pipe idled ~500 of every 1500 cycles (profiles/r02_mmq_dev_v2.txt).  Deferred,
                // the commits still complete when their MMAs do (they are queued before those finish), and the warp does
                // its bookkeeping while four MMAs are still queued.
                uint32_t pend_a = 0, pend_b = 0;
                for (int c = tc.c0; c < tc.c1; ++c, ++g) {
                    if (lane == 0) dbg_stamp(p, g, 0);
                    mbar_wait(bars + 8 * (BAR_A_FULL + aslot), a_ph);
                    if (lane == 0) dbg_stamp(p, g, 1);
                    if (!(DEVF & 128)) mbar_wait(bars + 8 * (BAR_B_FULL + bslot), b_ph);
                    if (lane == 0) dbg_stamp(p, g, 2);
                    // no tcgen05.fence here: operands arrive through shared memory (mbarrier-ordered); the fence is only
                    // needed around TMEM hand-offs (accumulator release above)
                    if (elect_one()) {
                        if (pend_a) {
                            if (p.cluster == 1) umma_commit(pend_b);
                            else umma_commit_multicast(pend_b, cmask);
                            umma_commit(pend_a);
                        }
                        if (!(DEVF & 32)) {
#pragma unroll
                            for (int sn = 0; sn < 2; ++sn) {
                                if (sn < tc.ns) {
#pragma unroll
                                    for (int kk = 0; kk < CHUNK_K / 16; ++kk) {
                                        // +32 bytes (16 bf16) along K inside the 128-byte swizzle atom = +2 in the address
                                        // field; the second column tile sits B_STAGE_BYTES further
                                        umma_f16(d_tmem[sn], adesc + 2 * kk, bdesc + 2 * kk + sn * (B_STAGE_BYTES >> 4), idesc[sn],
                                                 c != tc.c0 || kk != 0);
                                    }
                                }
                            }
                        }
                        if (DEVF & 1024) dbg_stamp(p, g, 7);  // timing aid: MMAs issued (replaces the B producer's stamp)
                        if (c == tc.c1 - 1) {  // end of the unit: nothing follows to hide behind
                            if (p.cluster == 1) umma_commit(bars + 8 * (BAR_B_EMPTY + bslot));
                            else umma_commit_multicast(bars + 8 * (BAR_B_EMPTY + bslot), cmask);
                            umma_commit(bars + 8 * (BAR_A_EMPTY + aslot));
                            umma_commit(bars + 8 * (BAR_ACC_FULL + abi));
                        }
                    }
                    __syncwarp();
                    pend_a = bars + 8 * (BAR_A_EMPTY + aslot);
                    pend_b = bars + 8 * (BAR_B_EMPTY + bslot);
                    if (c == tc.c1 - 1) pend_a = 0;
                    if (lane == 0) dbg_stamp(p, g, 3);
                    bdesc += (uint64_t) (p.b_stage_bytes >> 4);
                    if (++bslot == p.b_stages) {
                        bslot = 0;
                        b_ph ^= 1;
                        bdesc = bdesc0;
                    }
                    adesc += A_STAGE_BYTES >> 4;
                    if (++aslot == A_STAGES) {
                        aslot = 0;
                        a_ph ^= 1;
                        adesc = adesc0;
                    }
                }
                if (lane == 0) dbg_phase(p, 2);
            }
        }
    } else if (warp >= WARP_DQ0) {
        // ------------------------------------------------------------------ dequantizers: W ring -> A ring (bf16)
        const int dt   = (threadIdx.x - WARP_DQ0 * 32) & 255;  // 0..255 inside the group
        const int grp  = (threadIdx.x - WARP_DQ0 * 32) >> 8;   // group g handles k-chunks with (chunk & 1) == g
        const int row  = dt >> 1, half = dt & 1;
        const uint32_t a_row_off = (uint32_t) row * 128u;
        const uint32_t sw        = (uint32_t) (row & 7);
        int wsi = 0, g = 0;
        for (int t = cid; t < num_tiles; t += ncl) {
            const TileCoord tc  = tile_coord(p, t, crank);
            const uint8_t * wb  = p.w + tc.w_off;
            const int       grow = min(tc.mt * TILE_M + row, p.m - 1);
            const uint8_t * rowp = wb + (long long) grow * p.w_nb1;
            for (int ws = tc.ws0; ws < tc.ws1; ++ws, ++wsi) {
                const int wslot = wsi % p.w_stages;
                if (!(DEVF & 64)) mbar_wait(bars + 8 * (BAR_W_FULL + wslot), (wsi / p.w_stages) & 1);
                const uint32_t  lead  = (uint32_t) ((uintptr_t) (rowp + (long long) ws * p.piece_bytes) & 15);
                const uint32_t  piece = base + p.off_w + (uint32_t) wslot * p.w_stage_bytes + (uint32_t) row * p.w_pitch + lead;
                const int cps     = p.w_stage_k / CHUNK_K;
                const int nchunks = min(cps, tc.c1 - ws * cps);
                for (int c = 0; c < nchunks; ++c, ++g) {
                    if ((g & (N_DQ_GROUPS - 1)) != grp) continue;
                    const int slot = g % A_STAGES;
                    float v[32];
                    if (dt == 0 && !(DEVF & 512)) dbg_stamp(p, g, 4);
                    if (DEVF & 2) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = (float) (i + row);
                    } else {
                        Dq<T>::run(piece, 2 * c + half, v);
                    }
                    if (dt == 0 && !(DEVF & 512)) dbg_stamp(p, g, 5);
                    mbar_wait(bars + 8 * (BAR_A_EMPTY + slot), ((g / A_STAGES) & 1) ^ 1);
                    if (dt == 0 && !(DEVF & 512)) dbg_stamp(p, g, 6);
                    uint8_t * arow = sm + (uint32_t) slot * A_STAGE_BYTES + a_row_off;
                    if (!(DEVF & 8))
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * i + 0], v[8 * i + 1]);
                        __nv_bfloat162 h1 = __floats2bfloat162_rn(v[8 * i + 2], v[8 * i + 3]);
                        __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * i + 4], v[8 * i + 5]);
                        __nv_bfloat162 h3 = __floats2bfloat162_rn(v[8 * i + 6], v[8 * i + 7]);
                        uint4 o;
                        o.x = *(uint32_t *) &h0; o.y = *(uint32_t *) &h1; o.z = *(uint32_t *) &h2; o.w = *(uint32_t *) &h3;
                        *(uint4 *) (arow + ((((uint32_t) (4 * half + i)) ^ sw) << 4)) = o;
                    }
                    if (!(DEVF & 16)) fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bars + 8 * (BAR_A_FULL + slot));
                }
                __syncwarp();
                if (lane == 0 && !(DEVF & 64)) mbar_arrive(bars + 8 * (BAR_W_EMPTY + wslot));
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue: TMEM -> dst
        int ti = 0;
        for (int t = cid; t < num_tiles; t += ncl, ++ti) {
            const TileCoord tc  = tile_coord(p, t, crank);
            const int       i13 = tc.b / p.ne12, i12 = tc.b - i13 * p.ne12;
            const int       ab  = dual ? 0 : (ti & 1);
            const uint32_t  aph = dual ? (ti & 1) : ((ti >> 1) & 1);
            const int       row = tc.mt * TILE_M + warp * 32 + lane;
            char * dbase = (char *) p.dst + (long long) i12 * p.d_nb2 + (long long) i13 * p.d_nb3 + (long long) row * 4;
            mbar_wait(bars + 8 * (BAR_ACC_FULL + ab), aph);
            tc_fence_after();
            if (threadIdx.x == 0) dbg_phase(p, 3);
            for (int sn = 0; sn < tc.ns; ++sn) {
                const int n0    = (tc.nt + sn) * TILE_N;
                const int ncols = tc.ncols[sn];
                const int acol  = (dual ? sn : ab) * TILE_N;
                for (int c0 = 0; c0 < ncols; c0 += 32) {
                    float v[32];
                    tmem_ld32(tmem_base + ((uint32_t) (warp * 32) << 16) + (uint32_t) (acol + c0), v);
                    if (row < p.m && p.moe) {
                        // scatter: column j of this tile is the (token, slot) row moe_rows[row_base + j] of dst
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (c0 + j < ncols) {
                                const int r  = __ldg(p.moe_rows + tc.row_base + c0 + j);
                                const int tk = r / p.n_used, sl = r - tk * p.n_used;
                                *(float *) ((char *) p.dst + (long long) tk * p.d_nb2 + (long long) sl * p.d_nb1 + (long long) row * 4) = v[j];
                            }
                        }
                    } else if (row < p.m) {
                        char * dp = dbase + (long long) (n0 + c0) * p.d_nb1;  // one coalesced 128-byte warp store per column
                        if (p.ksplit > 1) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                if (c0 + j < ncols) atomicAdd((float *) dp, v[j]);  // RED.ADD.F32 (result unused)
                                dp += p.d_nb1;
                            }
                        } else if (c0 + 32 <= ncols) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                *(float *) dp = v[j];
                                dp += p.d_nb1;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                if (c0 + j < ncols) *(float *) dp = v[j];
                                dp += p.d_nb1;
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * (BAR_ACC_EMPTY + ab));
            if (threadIdx.x == 0) dbg_phase(p, 4);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (p.cluster > 1) cluster_sync();  // no CTA leaves while a peer may still multicast into it / arrive on its barriers
    if (threadIdx.x == 0) dbg_phase(p, 5);
    if (warp == WARP_MMA) {
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---- MoE routing: ids [T][n_used] -> per-expert row lists and the column-tile table ---------------------------------
// header layout (int32): [0] total tiles | [16, 16+E) count | [.., +E) offset | [.., +maxT) tile_expert |
// [.., +maxT) tile_col0 | [.., +R) rows  (R = T * n_used, maxT = ceil(R / TILE_N) + E)
struct MoeTables {
    int * total, * count, * offset, * tile_expert, * tile_col0, * rows;
    size_t bytes;  // header bytes, rounded up to 1024
};
inline MoeTables moe_tables(void * ws, int n_expert, int rows, int max_tiles) {
    MoeTables t;
    int *     b   = (int *) ws;
    t.total       = b;
    t.count       = b + 16;
    t.offset      = t.count + n_expert;
    t.tile_expert = t.offset + n_expert;
    t.tile_col0   = t.tile_expert + max_tiles;
    t.rows        = t.tile_col0 + max_tiles;
    t.bytes       = ((size_t) (16 + 2 * n_expert + 2 * max_tiles + rows) * 4 + 1023) / 1024 * 1024;
    return t;
}
struct MoeRouteParams {
    const int32_t * ids;
    long long       ids_nb1;
    int             n_tokens, n_used, n_expert;
    MoeTables       t;
};
// one CTA.  Rows of an expert are listed in increasing (token, slot) order (ranks by a prefix count, no atomics), so
// the layout — and with it every byte the MMQ kernel reads or writes — is reproducible.
__global__ void __launch_bounds__(1024) moe_route(const MoeRouteParams p) {
    extern __shared__ int rs[];  // [n_expert] counts, then per-thread-chunk partial counts are avoided: two passes
    const int R = p.n_tokens * p.n_used;
    for (int e = threadIdx.x; e < p.n_expert; e += blockDim.x) rs[e] = 0;
    __syncthreads();
    for (int r = threadIdx.x; r < R; r += blockDim.x) {
        const int t = r / p.n_used, sl = r - t * p.n_used;
        const int e = *(const int32_t *) ((const char *) p.ids + (long long) t * p.ids_nb1 + sl * 4);
        if (e >= 0 && e < p.n_expert) atomicAdd(&rs[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int off = 0, ts = 0;
        for (int e = 0; e < p.n_expert; ++e) {
            const int c   = rs[e];
            p.t.count[e]  = c;
            p.t.offset[e] = off;
            for (int c0 = 0; c0 < c; c0 += TILE_N, ++ts) {
                p.t.tile_expert[ts] = e;
                p.t.tile_col0[ts]   = c0;
            }
            off += c;
        }
        p.t.total[0] = ts;
    }
    __syncthreads();
    // stable fill: one warp per expert scans the rows in order (ballot + popc ranks)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int e = warp; e < p.n_expert; e += nwarps) {
        if (rs[e] == 0) continue;
        int pos = p.t.offset[e];
        for (int r0 = 0; r0 < R; r0 += 32) {
            const int r = r0 + lane;
            bool      mine = false;
            if (r < R) {
                const int t = r / p.n_used, sl = r - t * p.n_used;
                mine = *(const int32_t *) ((const char *) p.ids + (long long) t * p.ids_nb1 + sl * 4) == e;
            }
            const unsigned bal = __ballot_sync(0xffffffffu, mine);
            if (mine) p.t.rows[pos + __popc(bal & ((1u << lane) - 1u))] = r;
            pos += __popc(bal);
        }
    }
}

struct MoePackParams {
    const float * act;
    uint8_t *     out;
    MoeTables     t;
    int           k_chunks, n_used, b_bcast, max_tiles;
    long long     a_nb1, a_nb2;  // byte strides of b: slot, token
};
// gathered f32 rows -> the same swizzled bf16 tile image as mmq_pack_act; block (tile, k_chunk)
__global__ void __launch_bounds__(256) moe_pack_act(const MoePackParams p) {
    const int       total_tiles = p.t.total[0];
    const long long per_tile    = (long long) p.k_chunks * TILE_N * 8;
    const long long total       = (long long) total_tiles * per_tile;
    for (long long idx = (long long) blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long) gridDim.x * blockDim.x) {
        const int c  = (int) (idx & 7);
        const int j  = (int) ((idx >> 3) & (TILE_N - 1));
        const long long blk = idx / (TILE_N * 8);
        const int kc = (int) (blk % p.k_chunks);
        const int ts = (int) (blk / p.k_chunks);
        const int e  = p.t.tile_expert[ts], col = p.t.tile_col0[ts] + j, cnt = p.t.count[e];
        if (col >= ((cnt + 15) & ~15)) continue;  // the MMA reads whole groups of 16 columns only
        uint4 o = make_uint4(0, 0, 0, 0);
        if (col < cnt) {
            const int     r  = p.t.rows[p.t.offset[e] + col];
            const int     tk = r / p.n_used, sl = r - tk * p.n_used;
            const float * src = (const float *) ((const char *) p.act + (long long) tk * p.a_nb2 + (p.b_bcast ? 0ll : (long long) sl * p.a_nb1)) +
                                kc * CHUNK_K + c * 8;
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldg(src + i);
            __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v[4], v[5]), h3 = __floats2bfloat162_rn(v[6], v[7]);
            o.x = *(uint32_t *) &h0; o.y = *(uint32_t *) &h1; o.z = *(uint32_t *) &h2; o.w = *(uint32_t *) &h3;
        }
        *(uint4 *) (p.out + blk * (long long) B_STAGE_BYTES + (long long) j * 128 + ((c ^ (j & 7)) << 4)) = o;
    }
}

typedef void (*mmq_kernel_t)(const MmqParams);
unsigned long long * g_mmq_dbg = nullptr;

mmq_kernel_t mmq_kernel_for(int type) {
    switch (type) {
        case T_Q4_0:  return mmq_kernel<T_Q4_0>;
        case T_Q4_1:  return mmq_kernel<T_Q4_1>;
        case T_Q5_0:  return mmq_kernel<T_Q5_0>;
        case T_Q5_1:  return mmq_kernel<T_Q5_1>;
        case T_Q8_0:  return mmq_kernel<T_Q8_0>;
        case T_MXFP4: return mmq_kernel<T_MXFP4>;
        case T_Q4_K:  return mmq_kernel<T_Q4_K>;
        case T_Q5_K:  return mmq_kernel<T_Q5_K>;
        case T_Q6_K:  return mmq_kernel<T_Q6_K>;
        case T_F16:   return mmq_kernel<T_F16>;
        case T_BF16:  return mmq_kernel<T_BF16>;
        case T_F32W:  return mmq_kernel<T_F32W>;
        case T_Q2_K:  return mmq_kernel<T_Q2_K>;
        case T_Q3_K:  return mmq_kernel<T_Q3_K>;
        case T_IQ4_NL: return mmq_kernel<T_IQ4_NL>;
        case T_IQ4_XS: return mmq_kernel<T_IQ4_XS>;
        default: return nullptr;
    }
}

inline int round_up_i(int x, int a) { return (x + a - 1) / a * a; }

}  // namespace

void mmq_set_debug(unsigned long long * buf) { g_mmq_dbg = buf; }

bool mmq_supported(int type, int64_t k) {
    return mmq_kernel_for(type) != nullptr && k > 0 && k % CHUNK_K == 0 && k % blck_elems(type) == 0;
}

size_t mmq_workspace_bytes(int64_t k, int64_t n, int64_t nbatch) {
    const int64_t n_tiles = (n + TILE_N - 1) / TILE_N;
    return (size_t) (nbatch * ((n_tiles + 1) / 2 * 2) * (k / CHUNK_K)) * B_STAGE_BYTES;  // room for the padding tile of a pair
}

size_t mmq_moe_workspace_bytes(int64_t k, int64_t n_tokens, int64_t n_used, int64_t n_expert) {
    const int64_t rows = n_tokens * n_used, max_tiles = (rows + TILE_N - 1) / TILE_N + n_expert;
    return moe_tables(nullptr, (int) n_expert, (int) rows, (int) max_tiles).bytes + (size_t) (max_tiles * (k / CHUNK_K)) * B_STAGE_BYTES;
}

static int mmq_launch_impl(const MmqArgs & a, const MmqMoeArgs * moe, void * workspace, cudaStream_t st, int sms, std::string & err);

// what the last plain (non-MoE) launch of this thread packed, and where (see the reuse in mmq_launch)
struct LastPack {
    const void * ws = nullptr; const float * act = nullptr; cudaStream_t st = nullptr;
    long long k = 0, n = 0, ne12 = 0, ne13 = 0, nb1 = 0, nb2 = 0, nb3 = 0; int n_sub = 0, n_super = 0;
};
static thread_local LastPack g_last_pack;
void mmq_forget_pack() { g_last_pack = LastPack{}; }

int mmq_launch(const MmqArgs & a, void * workspace, cudaStream_t st, int sms, std::string & err) {
    return mmq_launch_impl(a, nullptr, workspace, st, sms, err);
}

// MUL_MAT_ID for many tokens: the (token, slot) rows are grouped per expert on the device and every expert's rows run
// through the tensor-core kernel as column tiles of that expert's matrix (replaces the reference's mmq ids path,
// ggml-cuda/mmq.cu:161-, mmid.cu)
int mmq_moe_launch(const MmqMoeArgs & m, void * workspace, cudaStream_t st, int sms, std::string & err) {
    if (m.n_expert > 1024) {
        err = "mmq: more than 1024 experts";
        return B200Q_ENOTSUP;
    }
    MmqArgs a{};
    a.type = m.type; a.k = m.k; a.m = m.m; a.n = TILE_N;
    a.ne02 = a.ne03 = a.ne12 = a.ne13 = 1;
    a.w_nb1 = m.w_nb1; a.w_nb2 = a.w_nb3 = 0;
    a.d_nb1 = m.d_nb1; a.d_nb2 = m.d_nb2; a.d_nb3 = 0;
    a.w = m.w; a.act = m.act; a.dst = m.dst;
    return mmq_launch_impl(a, &m, workspace, st, sms, err);
}

static int mmq_launch_impl(const MmqArgs & a, const MmqMoeArgs * moe, void * workspace, cudaStream_t st, int sms, std::string & err) {
    mmq_kernel_t fn = mmq_kernel_for(a.type);
    if (!fn || !mmq_supported(a.type, a.k)) {
        err = "mmq: unsupported type / k";
        return B200Q_ENOTSUP;
    }
    if (((uintptr_t) workspace & 15) != 0) {
        err = "mmq: workspace must be 16-byte aligned";
        return B200Q_EINVAL;
    }
    const int nbatch = (int) (a.ne12 * a.ne13);

    PackParams pp{};
    pp.act = a.act; pp.out = (uint8_t *) workspace;
    pp.k = (int) a.k; pp.n = (int) a.n; pp.ne12 = (int) a.ne12; pp.nbatch = nbatch;
    pp.n_tiles  = (int) ((a.n + TILE_N - 1) / TILE_N);
    pp.k_chunks = (int) (a.k / CHUNK_K);
    pp.a_nb1 = a.a_nb1; pp.a_nb2 = a.a_nb2; pp.a_nb3 = a.a_nb3;
    pp.vec4  = ((uintptr_t) a.act & 15) == 0 && (a.a_nb1 & 15) == 0 && (a.a_nb2 & 15) == 0 && (a.a_nb3 & 15) == 0;

    MmqParams p{};
    MoeTables mt{};
    int       moe_max_tiles = 0;
    if (moe) {
        const int rows = (int) (moe->n_tokens * moe->n_used);
        moe_max_tiles  = (rows + TILE_N - 1) / TILE_N + (int) moe->n_expert;
        mt             = moe_tables(workspace, (int) moe->n_expert, rows, moe_max_tiles);
        p.moe = mt.total; p.moe_tile_expert = mt.tile_expert; p.moe_tile_col0 = mt.tile_col0;
        p.moe_count = mt.count; p.moe_offset = mt.offset; p.moe_rows = mt.rows;
        p.n_used = (int) moe->n_used; p.w_nb_expert = moe->w_nb2;
    }
    p.w = a.w; p.actp = (const uint8_t *) workspace + mt.bytes; p.dst = a.dst;
    p.k = (int) a.k; p.m = (int) a.m; p.n = (int) a.n;
    p.ne12 = (int) a.ne12; p.r2 = (int) (a.ne12 / a.ne02); p.r3 = (int) (a.ne13 / a.ne03); p.nbatch = nbatch;
    p.w_nb1 = a.w_nb1; p.w_nb2 = a.w_nb2; p.w_nb3 = a.w_nb3;
    p.d_nb1 = a.d_nb1; p.d_nb2 = a.d_nb2; p.d_nb3 = a.d_nb3;
    p.m_tiles  = (int) ((a.m + TILE_M - 1) / TILE_M);
    p.n_tiles  = pp.n_tiles;
    p.k_chunks = pp.k_chunks;
    const int be = blck_elems(a.type), bb = blck_bytes(a.type);
    p.w_stage_k     = a.type == T_F32W ? 64 : be == 32 ? 128 : 256;  // raw bytes per row and stage stay <= 256
    p.w_stages_per_tile = (int) ((a.k + p.w_stage_k - 1) / p.w_stage_k);
    p.piece_bytes   = p.w_stage_k / be * bb;
    p.blk_bytes     = be == 32 ? bb : 0;
    p.w_pitch       = round_up_i(p.piece_bytes + 15 + 4, 16);  // lead <= 15 bytes + one word of funnel-shift over-read
    if (!((p.w_pitch >> 4) & 1)) p.w_pitch += 16;              // odd multiple of 16: rows spread over the banks
    p.w_stage_bytes = TILE_M * p.w_pitch;
    // unit shape: two accumulators share each dequantized chunk (n_sub = 2) when that still leaves a unit for most SMs;
    // otherwise one accumulator per unit, and the k range is halved (ksplit = 2) if even that underfills the GPU
    static const int force_sub = getenv("B200Q_MMQ_NSUB") ? atoi(getenv("B200Q_MMQ_NSUB")) : 0;
    static const int force_ks  = getenv("B200Q_MMQ_KSPLIT") ? atoi(getenv("B200Q_MMQ_KSPLIT")) : 0;
    static const int force_cl  = getenv("B200Q_MMQ_CLUSTER") ? atoi(getenv("B200Q_MMQ_CLUSTER")) : 0;
    {
        const long long fill       = (long long) sms * 6 / 10;
        const long long units_dual = (long long) nbatch * p.m_tiles * ((p.n_tiles + 1) / 2);
        const long long units_one  = (long long) nbatch * p.m_tiles * p.n_tiles;
        p.n_sub  = p.n_tiles > 1 && units_dual >= fill ? 2 : 1;
        p.ksplit = p.n_sub == 1 && units_one < fill && p.w_stages_per_tile >= 4 ? 2 : 1;
        if (force_sub) p.n_sub = p.n_tiles > 1 ? std::min(force_sub, 2) : 1;
        if (force_ks) p.ksplit = std::max(1, std::min(force_ks, std::min(8, p.w_stages_per_tile / 2)));
        if (moe) p.n_sub = p.ksplit = 1;
    }
    p.off_b = A_STAGES * A_STAGE_BYTES;
    auto fit_rings = [&](int n_sub) {  // B ring (stages of n_sub x 32 KB, at most 128 KB) + W ring in what is left
        p.b_stage_bytes = n_sub * B_STAGE_BYTES;
        for (int b32 = MAX_B_STAGES; b32 >= 2 * n_sub; b32 -= n_sub) {
            p.b_stages      = b32 / n_sub;
            p.off_w         = p.off_b + p.b_stages * p.b_stage_bytes;
            const int avail = 227 * 1024 - 1024 - (int) p.off_w - BAR_COUNT * 8 - 16;
            p.w_stages      = std::min(MAX_W_STAGES, avail / p.w_stage_bytes);
            if (p.w_stages >= 2) return true;
        }
        return false;
    };
    if (!fit_rings(p.n_sub) && p.n_sub == 2) {  // wide raw rows (q6_K, q5_K): one accumulator per unit
        p.n_sub = 1;
        p.w_stages = 0;
        fit_rings(1);
    }
    p.ws_per_split = (p.w_stages_per_tile + p.ksplit - 1) / p.ksplit;
    p.n_super = (p.n_tiles + p.n_sub - 1) / p.n_sub;
    p.cluster = force_cl && !moe ? force_cl : 1;  // multicast of the B chunks measured no gain (L2 already dedups the reads)
    while (p.cluster > 1 && sms / p.cluster < 1) p.cluster >>= 1;
    p.m_groups = (p.m_tiles + p.cluster - 1) / p.cluster;
    if (p.w_stages < 2) {
        err = "mmq: W ring does not fit in shared memory";
        return B200Q_ENOTSUP;
    }
    p.off_bar = p.off_w + p.w_stages * p.w_stage_bytes;
    p.dbg = g_mmq_dbg;
    {
        static const int flags = getenv("B200Q_MMQ_DBG") ? atoi(getenv("B200Q_MMQ_DBG")) : 0;
        p.dbg_flags = flags;
    }
    const int smem = 1024 + (int) p.off_bar + BAR_COUNT * 8 + 16;

    static std::mutex                            mu;
    static std::set<std::pair<int, const void *>> done;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lk(mu);
        if (!done.count({ dev, (const void *) fn })) {
            cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem > 227 * 1024 ? smem : 227 * 1024);
            if (e != cudaSuccess) {
                err = std::string("mmq: cudaFuncSetAttribute: ") + cudaGetErrorString(e);
                return B200Q_ECUDA;
            }
            done.insert({ dev, (const void *) fn });
        }
    }
    const int num_units = moe ? moe_max_tiles * p.m_tiles : nbatch * p.m_groups * p.n_super * p.ksplit;  // MoE: upper bound
    cudaLaunchConfig_t cfg{};
    cfg.blockDim         = dim3(NTHREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[1];
    attr[0].id               = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = p.cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs                = attr;
    cfg.numAttrs             = 1;
    // persistent grid = the clusters that are co-resident (a GPC whose SM count is not a multiple of the cluster size
    // leaves SMs out, so this can be less than sms / cluster)
    int max_clusters = sms / p.cluster;
    {
        static std::map<std::tuple<int, const void *, int, int>, int> cache;
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lk(mu);
        auto key = std::make_tuple(dev, (const void *) fn, p.cluster, smem);
        auto it  = cache.find(key);
        if (it == cache.end()) {
            int n = 0;
            cfg.gridDim = dim3(sms / p.cluster * p.cluster);
            if (cudaOccupancyMaxActiveClusters(&n, fn, &cfg) != cudaSuccess || n < 1) {
                cudaGetLastError();
                n = max_clusters;
            }
            it = cache.emplace(key, n).first;
        }
        max_clusters = std::min(max_clusters, it->second);
    }
    const int grid = std::min(num_units, max_clusters) * p.cluster;
    cfg.gridDim    = dim3(grid);
    static const bool debug = getenv("B200Q_DEBUG") != nullptr;
    if (debug) {
        fprintf(stderr, "[b200q] mmq type=%d m=%d n=%d k=%d cluster=%d n_sub=%d ksplit=%d units=%d grid=%d w_stages=%d b_stages=%d pitch=%d\n",
                a.type, p.m, p.n, p.k, p.cluster, p.n_sub, p.ksplit, num_units, grid, p.w_stages, p.b_stages, p.w_pitch);
    }
    if (moe) {
        g_last_pack = LastPack{};  // the workspace now holds the MoE tables and per-expert tiles
        MoeRouteParams rp{};
        rp.ids = moe->ids; rp.ids_nb1 = moe->ids_nb1;
        rp.n_tokens = (int) moe->n_tokens; rp.n_used = (int) moe->n_used; rp.n_expert = (int) moe->n_expert;
        rp.t = mt;
        moe_route<<<1, 1024, (size_t) moe->n_expert * sizeof(int), st>>>(rp);
        MoePackParams mp{};
        mp.act = moe->act; mp.out = (uint8_t *) workspace + mt.bytes; mp.t = mt;
        mp.k_chunks = p.k_chunks; mp.n_used = (int) moe->n_used; mp.b_bcast = moe->b_bcast ? 1 : 0; mp.max_tiles = moe_max_tiles;
        mp.a_nb1 = moe->a_nb1; mp.a_nb2 = moe->a_nb2;
        moe_pack_act<<<sms * 8, 256, 0, st>>>(mp);
    } else {
        if (p.ksplit > 1) {
            pp.zero_dst = a.dst; pp.m = (int) a.m;
            pp.d_nb1 = a.d_nb1; pp.d_nb2 = a.d_nb2; pp.d_nb3 = a.d_nb3;
        }
        pp.n_sub = p.n_sub; pp.n_super = p.n_super;
        // consecutive matmuls of one activation matrix (wq / wk / wv, ffn_gate / ffn_up): the caller may state that `act` is
        // unchanged since the previous call (a.reuse_pack); the image is reused when that call packed the same matrix in
        // the same layout into the same workspace on the same stream (per thread: a backend drives its stream from one)
        LastPack & last = g_last_pack;
        LastPack   cur;
        cur.ws = workspace; cur.act = a.act; cur.st = st; cur.k = a.k; cur.n = a.n; cur.ne12 = a.ne12; cur.ne13 = a.ne13;
        cur.nb1 = a.a_nb1; cur.nb2 = a.a_nb2; cur.nb3 = a.a_nb3; cur.n_sub = p.n_sub; cur.n_super = p.n_super;
        const bool same = a.reuse_pack && last.ws == cur.ws && last.act == cur.act && last.st == cur.st && last.k == cur.k && last.n == cur.n &&
                          last.ne12 == cur.ne12 && last.ne13 == cur.ne13 && last.nb1 == cur.nb1 && last.nb2 == cur.nb2 && last.nb3 == cur.nb3 &&
                          last.n_sub == cur.n_sub && last.n_super == cur.n_super;
        last = cur;
        pp.zero_only = same ? 1 : 0;
        if (!same || pp.zero_dst) {
            // zeroing: a thread per dst value; packing: a CTA per 32 KB block (both grid-stride)
            const long long zero_ctas = pp.zero_dst ? ((long long) nbatch * a.n * a.m + 255) / 256 : 0;
            const long long pack_ctas = same ? 0 : (long long) nbatch * pp.n_super * pp.n_sub * pp.k_chunks;
            const int       pgrid     = (int) std::max<long long>(1, std::min<long long>(std::max(zero_ctas, pack_ctas), (long long) sms * 8));
            mmq_pack_act<<<pgrid, 256, 0, st>>>(pp);
        }
    }
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, p);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        err = std::string("mmq: launch failed: ") + cudaGetErrorString(e);
        return B200Q_ECUDA;
    }
    return 0;
}

}  // namespace b200q
