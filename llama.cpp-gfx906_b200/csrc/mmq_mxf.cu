// mmq_mxf.cu — MXFP4 weights on the NATIVE block-scaled tensor-core path (north star: "MXFP4 uses native block-scaled MMA";
// SURVEY H3; replaces load_tiles_mxfp4 + vec_dot_q8_0_q8_1_mma of ggml-cuda/mmq.cuh:699-762, 796-931 and the generic
// bf16 route of csrc/mmq.cu for this type).
//
// tcgen05.mma kind::mxf8f6f4.block_scale multiplies e2m1 weights by e4m3 activations, both with UE8M0 scales per 32
// k-elements (csrc/mxf_probe.cu validated the operand layouts against a CPU emulation and measured 128 cycles per
// 128 x 256 x 32 MMA = twice the bf16 rate).  ggml's block_mxfp4 IS that weight format — the nibble codes are e2m1, the
// exponent byte is the UE8M0 scale (kvalues_mxfp4 are the doubled code points and ggml halves the scale: the product is
// the same) — so the weight side needs NO dequantization, only a nibble re-order (ggml keeps element j in the low and
// element j + 16 in the high nibble of byte j; the MMA wants 16 consecutive elements in 8 bytes).
// One e4m3 term per activation is not enough for the 5e-4 NMSE gate (5.5e-4 .. 6.9e-4 measured): activations
// go in as hi + lo (x ~= s_hi * hi + s_lo * lo, both e4m3 with their own block scales; NMSE 3e-7) and every k-chunk issues
// two MMAs into the same accumulator.  Effective ceiling = the bf16 rate, but the CUDA cores do ~25 instructions per 32
// weights instead of ~130, which is what bounded MoE prompt processing (64 tokens per expert: 8 % of the bf16 peak).
//
// CTA = 10 warps, persistent over units of 128 weight rows x <= 128 columns, k in chunks of 128:
//   warps 0-3  epilogue: TMEM -> dst (MoE: scattered through the expert's row list)
//   warps 4-7  weight producers: warp w takes the chunks g = w (mod 4) and owns A stage w: coalesced 16-byte cp.async bring the
//              128 row pieces (68 bytes each, any alignment: the 96-byte aligned window around them) into a raw stage, then a
//              lane re-orders the nibbles of its (row, 32-block) items into the swizzled K-major tile and writes the scale
//              bytes.  (Word loads straight from global memory — 32 lanes on 32 different rows — cost ~2400 LSU cycles per
//              chunk against 512 cycles of MMA; one 96-byte bulk copy per row was slower still: ~15 ns per small copy.)
//   warp 8     activation producer: one bulk copy per chunk brings [hi tile | lo tile | scales hi | scales lo] (33 KB),
//              laid out by mxf_pack_act as the exact shared-memory image
//   warp 9     issues, per chunk, 3 tcgen05.cp (scale factors -> TMEM) and 8 MMAs (4 x K = 32, hi and lo)
// TMEM: two 128-column accumulators (the epilogue of one unit overlaps the next unit's MMAs) + the scale-factor columns.
#include "mmq_mxf.cuh"

#include "blocks.cuh"
#include "sm100.cuh"

#include "../../include/b200q.h"

#include <cuda_fp8.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <map>
#include <mutex>

namespace b200q {

namespace {
using namespace sm100;

constexpr int TM = 128, TN = 128, KC = 128;      // unit: rows x columns, k per chunk
constexpr int A_STAGES = 4, B_STAGES = 3;
constexpr int RAW_PITCH = 112;                     // raw row piece: 16-byte aligned window of <= 96 bytes; 7 x 16: conflict-free LDS.128
constexpr int RAW_BYTES = TM * RAW_PITCH;          // one raw stage per producer warp
constexpr int A_STAGE_BYTES = TM * 128;            // 16 KB: one byte per k-element slot (4-bit values in 16-byte chunks of 8 + 8 pad)
constexpr int B_TILE_BYTES  = TN * 128;            // 16 KB
constexpr int SF_BYTES      = 512;                 // 32 x 16: scale bytes of 128 rows x 4 k-blocks
constexpr int B_STAGE_BYTES = 2 * B_TILE_BYTES + 2 * SF_BYTES;  // hi | lo | sf hi | sf lo = 33792 (a multiple of 1024)
constexpr int NTHREADS = 320;
constexpr int WARP_WPROD = 4, WARP_BPROD = 8, WARP_MMA = 9;
constexpr int TMEM_COLS = 512, T_ACC1 = 128, T_SF = 256;

enum { BAR_A_FULL = 0, BAR_A_EMPTY = 4, BAR_B_FULL = 8, BAR_B_EMPTY = 12, BAR_ACC_FULL = 16, BAR_ACC_EMPTY = 18, BAR_RAW = 20, BAR_COUNT = 24 };

struct MxfTables {  // MoE routing (device memory, head of the workspace)
    int * total, * count, * offset, * tile_expert, * tile_col0, * rows;
    size_t bytes;
};
inline MxfTables mxf_tables(void * ws, int n_expert, int rows, int max_tiles) {
    MxfTables t;
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

struct MxfParams {
    const uint8_t * w;      // block_mxfp4 rows
    const uint8_t * actp;   // packed activation stages
    float *         dst;
    int   k, m, n, k_chunks, m_tiles, n_tiles;
    long long w_nb1, w_nb_expert;
    long long d_nb1, d_nb2;
    // MoE (moe != nullptr): unit = (expert column tile, row tile)
    const int * moe;        // [0] = number of column tiles
    const int * moe_tile_expert, * moe_tile_col0, * moe_count, * moe_offset, * moe_rows;
    int   n_used;
    int   dbg;            // development: 1 producers skip fetch + re-order, 2 no activation copies, 4 no tcgen05.cp, 8 no MMAs, 16 no fetch
};

// ---- activations: f32 -> (hi, lo) e4m3 with UE8M0 block-32 scales, in the shared-memory image of a stage -------------------
// smallest power of two 2^E with amax / 2^E <= 448 (the largest e4m3 magnitude); blocks below 2^-100 are stored as zeros
__device__ __forceinline__ int e4m3_block_exp(float amax) {
    if (!(amax > 7.9e-31f)) return -127;
    int E = ((__float_as_int(amax) >> 23) & 0xff) - 127 - 8;             // amax / 2^E in [256, 512)
    if (amax * __int_as_float((127 - E) << 23) > 448.0f) ++E;
    return E;
}
__device__ __forceinline__ float pow2f(int e) { return __int_as_float((127 + e) << 23); }  // -126 <= e <= 127
__device__ __forceinline__ float e4m3_to_float(uint32_t v) {
    return __half2float(__half(__nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t) v, __NV_E4M3)));
}
// 8 consecutive lanes: the 32 activations of one (column, k-block) -> 2 x 32 bytes + 2 scale bytes.  `sub` = lane within the
// group (4 values each); all 8 lanes of a group must call this together (shuffles)
__device__ __forceinline__ void pack_block(const float * src, bool valid, uint8_t * stage, int j, int b, int sub) {
    const unsigned gmask = 0xFFu << (threadIdx.x & 24);  // the 8 lanes of this group
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
        if (((uintptr_t) src & 15) == 0) {
            x = __ldg((const float4 *) src + sub);
        } else {
            x = make_float4(__ldg(src + 4 * sub), __ldg(src + 4 * sub + 1), __ldg(src + 4 * sub + 2), __ldg(src + 4 * sub + 3));
        }
    }
    float amax = fmaxf(fmaxf(fabsf(x.x), fabsf(x.y)), fmaxf(fabsf(x.z), fabsf(x.w)));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(gmask, amax, o));
    const int E = e4m3_block_exp(amax);
    float     v[4] = { x.x, x.y, x.z, x.w };
    uint32_t  hiw = 0, low = 0;
    float     rmax = 0.0f;
    const float is = E > -127 ? pow2f(-E) : 0.0f;  // exact scaling; an all-(near-)zero block becomes zeros
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float    xs = v[i] * is;
        const uint32_t h  = (uint32_t) __nv_cvt_float_to_fp8(xs, __NV_SATFINITE, __NV_E4M3);
        hiw |= h << (8 * i);
        v[i] = xs - e4m3_to_float(h);  // exact residual
        rmax = fmaxf(rmax, fabsf(v[i]));
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) rmax = fmaxf(rmax, __shfl_xor_sync(gmask, rmax, o));
    int E2 = e4m3_block_exp(rmax);
    if (E + E2 < -127 || E == -127) E2 = -127;  // the combined scale must stay a UE8M0 value: drop the low term
    const float is2 = E2 > -127 ? pow2f(-E2) : 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) low |= (uint32_t) __nv_cvt_float_to_fp8(v[i] * is2, __NV_SATFINITE, __NV_E4M3) << (8 * i);
    const int c16 = 2 * b + (sub >> 2);
    const int off = j * 128 + ((c16 ^ (j & 7)) << 4) + (sub & 3) * 4;
    *(uint32_t *) (stage + off)                = hiw;
    *(uint32_t *) (stage + B_TILE_BYTES + off) = low;
    if (sub == 0) {
        stage[2 * B_TILE_BYTES + (j & 31) * 16 + (j >> 5) * 4 + b]            = (uint8_t) (E + 127);
        stage[2 * B_TILE_BYTES + SF_BYTES + (j & 31) * 16 + (j >> 5) * 4 + b] = (uint8_t) (E2 > -127 ? E + E2 + 127 : 0);
    }
}

struct PackParams {
    const float * act;
    uint8_t *     out;
    int           k, n, n_tiles, k_chunks;
    long long     a_nb1;
};
// dense: stage (n_tile, k_chunk); 8 threads = (column j of the tile, k-block b of the chunk)
__global__ void __launch_bounds__(256) mxf_pack_act(const PackParams p) {
    const long long total = (long long) p.n_tiles * p.k_chunks * TN * 4 * 8;  // a multiple of 256: whole warps stay together
    for (long long idx = (long long) blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long) gridDim.x * blockDim.x) {
        const int       sub = (int) (idx & 7);
        const int       b   = (int) ((idx >> 3) & 3);
        const int       j   = (int) ((idx >> 5) & (TN - 1));
        const long long blk = idx / (TN * 32);
        const int       kc  = (int) (blk % p.k_chunks);
        const int       nt  = (int) (blk / p.k_chunks);
        const int       col = nt * TN + j, k0 = kc * KC + b * 32;
        const bool      valid = col < p.n && k0 < p.k;
        const float *   src = (const float *) ((const char *) p.act + (long long) (valid ? col : 0) * p.a_nb1) + (valid ? k0 : 0);
        pack_block(src, valid, p.out + blk * (long long) B_STAGE_BYTES, j, b, sub);
    }
}

struct RouteParams {
    const int32_t * ids;
    long long       ids_nb1;
    int             n_tokens, n_used, n_expert;
    MxfTables       t;
};
// one CTA; rows of an expert in increasing (token, slot) order (ballot + popc ranks): reproducible layout.  The ids are staged
// in shared memory first (the rank scan re-reads them once per expert; from global memory that was 28 us for 2048 rows)
constexpr int ROUTE_SMEM_IDS = 8192;  // 32 KB: within the default dynamic shared-memory limit
__global__ void __launch_bounds__(1024) mxf_route(const RouteParams p) {
    extern __shared__ int rs[];  // [n_expert] counts | [min(R, ROUTE_SMEM_IDS)] expert of row r
    int *     sid = rs + p.n_expert;
    const int R = p.n_tokens * p.n_used;
    const bool staged = R <= ROUTE_SMEM_IDS;
    for (int e = threadIdx.x; e < p.n_expert; e += blockDim.x) rs[e] = 0;
    __syncthreads();
    for (int r = threadIdx.x; r < R; r += blockDim.x) {
        const int t = r / p.n_used, sl = r - t * p.n_used;
        const int e = *(const int32_t *) ((const char *) p.ids + (long long) t * p.ids_nb1 + sl * 4);
        if (staged) sid[r] = e;
        if (e >= 0 && e < p.n_expert) atomicAdd(&rs[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int off = 0, ts = 0;
        for (int e = 0; e < p.n_expert; ++e) {
            const int c   = rs[e];
            p.t.count[e]  = c;
            p.t.offset[e] = off;
            for (int c0 = 0; c0 < c; c0 += TN, ++ts) {
                p.t.tile_expert[ts] = e;
                p.t.tile_col0[ts]   = c0;
            }
            off += c;
        }
        p.t.total[0] = ts;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int e = warp; e < p.n_expert; e += nwarps) {
        if (rs[e] == 0) continue;
        int pos = p.t.offset[e];
        for (int r0 = 0; r0 < R; r0 += 32) {
            const int r = r0 + lane;
            bool      mine = false;
            if (r < R) {
                if (staged) {
                    mine = sid[r] == e;
                } else {
                    const int t = r / p.n_used, sl = r - t * p.n_used;
                    mine = *(const int32_t *) ((const char *) p.ids + (long long) t * p.ids_nb1 + sl * 4) == e;
                }
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
    MxfTables     t;
    int           k, k_chunks, n_used, b_bcast;
    long long     a_nb1, a_nb2;  // byte strides of b: slot, token
};
__global__ void __launch_bounds__(256) mxf_moe_pack_act(const MoePackParams p) {
    const int       tiles = p.t.total[0];
    const long long total = (long long) tiles * p.k_chunks * TN * 4 * 8;
    for (long long idx = (long long) blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long) gridDim.x * blockDim.x) {
        const int       sub = (int) (idx & 7);
        const int       b   = (int) ((idx >> 3) & 3);
        const int       j   = (int) ((idx >> 5) & (TN - 1));
        const long long blk = idx / (TN * 32);
        const int       kc  = (int) (blk % p.k_chunks);
        const int       ts  = (int) (blk / p.k_chunks);
        const int       e   = p.t.tile_expert[ts], col = p.t.tile_col0[ts] + j, cnt = p.t.count[e];
        if (col >= ((cnt + 15) & ~15)) continue;  // the MMA reads whole groups of 16 columns only (uniform over the 8 lanes of a group)
        const int  k0    = kc * KC + b * 32;
        const bool valid = col < cnt && k0 < p.k;
        const float * src = p.act;
        if (valid) {
            const int r  = p.t.rows[p.t.offset[e] + col];
            const int tk = r / p.n_used, sl = r - tk * p.n_used;
            src = (const float *) ((const char *) p.act + (long long) tk * p.a_nb2 + (p.b_bcast ? 0ll : (long long) sl * p.a_nb1)) + k0;
        }
        pack_block(src, valid, p.out + blk * (long long) B_STAGE_BYTES, j, b, sub);
    }
}

// ---- tcgen05 pieces the bf16 kernel does not need --------------------------------------------------------------------------
__device__ __forceinline__ uint64_t desc_k_noswizzle(uint32_t smem_addr) {  // 8-row x 16-byte core matrices stacked at 128 bytes
    return (uint64_t) ((smem_addr >> 4) & 0x3FFFu) | ((uint64_t) (16 >> 4) << 16) | ((uint64_t) (128 >> 4) << 32) | ((uint64_t) 1 << 46);
}
// cute::UMMA::InstrDescriptorBlockScaled: A e2m1 (5), B e4m3 (0), UE8M0 scales, both operands K-major
__device__ __forceinline__ uint32_t idesc_mxf(int N, int sf_id) {
    return ((uint32_t) sf_id << 4) | (5u << 7) | ((uint32_t) (N >> 3) << 17) | (1u << 23) | ((uint32_t) (TM >> 4) << 24) | ((uint32_t) sf_id << 29);
}
__device__ __forceinline__ void tmem_cp_sf(uint32_t taddr, uint64_t sdesc) {
    asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_mxf(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t sfa, uint32_t sfb, uint32_t acc) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc), "r"(sfa), "r"(sfb)
        : "memory");
}

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
__device__ __forceinline__ void sts_v2(uint32_t a, uint32_t x, uint32_t y) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void sts_u8(uint32_t a, uint32_t v) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

struct Unit {
    int mt, ncols, n0, row_base;  // row tile; valid columns; dense: first column | MoE: first entry in moe_rows
    long long w_off;              // byte offset of the expert's matrix
    long long stage0;             // index of the unit's first packed stage
};
__device__ __forceinline__ Unit unit_of(const MxfParams & p, int t) {
    Unit u;
    if (p.moe) {  // row tiles fastest: neighbouring CTAs share the activation stages
        u.mt         = t % p.m_tiles;
        const int ts = t / p.m_tiles;
        const int e  = p.moe_tile_expert[ts], col0 = p.moe_tile_col0[ts];
        u.ncols      = min(TN, p.moe_count[e] - col0);
        u.n0         = 0;
        u.row_base   = p.moe_offset[e] + col0;
        u.w_off      = (long long) e * p.w_nb_expert;
        u.stage0     = (long long) ts * p.k_chunks;
    } else {      // column tiles fastest: neighbouring CTAs share the weight rows
        const int nt = t % p.n_tiles;
        u.mt         = t / p.n_tiles;
        u.ncols      = min(TN, p.n - nt * TN);
        u.n0         = nt * TN;
        u.row_base   = 0;
        u.w_off      = 0;
        u.stage0     = (long long) nt * p.k_chunks;
    }
    return u;
}

extern __shared__ __align__(1024) uint8_t mxf_smem_raw[];

__global__ void __launch_bounds__(NTHREADS, 1) mmq_mxf_kernel(const MxfParams p) {
    const uint32_t base  = (smem_u32(mxf_smem_raw) + 1023u) & ~1023u;
    uint8_t *      sm    = mxf_smem_raw + (base - smem_u32(mxf_smem_raw));
    const uint32_t sA    = base;
    const uint32_t sB    = sA + A_STAGES * A_STAGE_BYTES;
    const uint32_t sSFA  = sB + B_STAGES * B_STAGE_BYTES;
    const uint32_t sRAW  = sSFA + A_STAGES * SF_BYTES;
    const uint32_t bars  = sRAW + A_STAGES * RAW_BYTES;
    const uint32_t tslot = bars + BAR_COUNT * 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_units = p.moe ? p.moe[0] * p.m_tiles : p.m_tiles * p.n_tiles;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) {
            mbar_init(bars + 8 * (BAR_A_FULL + i), 1);
            mbar_init(bars + 8 * (BAR_A_EMPTY + i), 1);
            mbar_init(bars + 8 * (BAR_B_FULL + i), 1);
            mbar_init(bars + 8 * (BAR_B_EMPTY + i), 1);
            mbar_init(bars + 8 * (BAR_RAW + i), 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bars + 8 * (BAR_ACC_FULL + i), 1);
            mbar_init(bars + 8 * (BAR_ACC_EMPTY + i), 4);
        }
        mbar_fence_init();
    }
    if (warp == WARP_MMA) tmem_alloc(tslot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tslot));

    if (warp >= WARP_WPROD && warp < WARP_WPROD + 4) {
        // ------------------------------------------------------------------ weights: nibble re-order, global -> A stage
        const int      pw    = warp - WARP_WPROD;
        uint8_t *      stage = sm + pw * A_STAGE_BYTES;
        const uint32_t raw_s = sRAW + pw * RAW_BYTES;
        int       g = 0;  // chunk counter of this CTA
        uint32_t  use = 0;
        // the 8 padding bytes of every 16-byte chunk are never written again: clear the stage once
        for (int i = lane; i < A_STAGE_BYTES / 16; i += 32) *(uint4 *) (stage + i * 16) = make_uint4(0, 0, 0, 0);
        const int ib = lane & 3;  // this lane's 32-block within a row piece; its rows are (lane >> 2) + 8 jj, jj = 0 .. 15
        for (int t = blockIdx.x; t < num_units; t += gridDim.x) {
            const Unit      u     = unit_of(p, t);
            const uint8_t * wbase = p.w + u.w_off;
            // fetch: 8 lanes per row (6 of them copy the 16-byte pieces of the row's 96-byte aligned window), 4 rows per
            // instruction, rows (lane >> 3) + 4 it; rows past the end of the matrix are skipped (their results are never stored)
            const int       frow0 = u.mt * TM + (lane >> 3);
            const uint8_t * fptr0 = wbase + (long long) frow0 * p.w_nb1;
            const int       fpc   = lane & 7;
            uint32_t lead0[2] = { 0u, 0u };  // (row pointer & 15) of the 16 rows this lane re-orders, 4 bits each
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int grow = min(u.mt * TM + (lane >> 2) + 8 * jj, p.m - 1);
                const uint32_t l = ((uint32_t) (uintptr_t) wbase + (uint32_t) grow * (uint32_t) p.w_nb1) & 15u;
                lead0[jj >> 3] |= l << (4 * (jj & 7));
            }
            for (int kc = 0; kc < p.k_chunks; ++kc, ++g) {
                if ((int) (g & 3) != pw) continue;
                const int nbk = min(4, (p.k - kc * KC) / 32);  // valid 32-blocks of this chunk
                // raw row pieces -> shared memory (coalesced 16-byte cp.async)
                if (!(p.dbg & 17)) {
                    const uint8_t * rp = fptr0 + (long long) kc * (KC / 32 * 17);
                    uint32_t        ds = raw_s + (uint32_t) ((lane >> 3) * RAW_PITCH + fpc * 16);
#pragma unroll 8
                    for (int it = 0; it < TM / 4; ++it) {
                        if (fpc < 6 && frow0 + 4 * it < p.m) {
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ds), "l"(((uintptr_t) rp & ~(uintptr_t) 15) + fpc * 16) : "memory");
                        }
                        rp += 4 * p.w_nb1;
                        ds += 4 * RAW_PITCH;
                    }
                    asm volatile("cp.async.commit_group;" ::: "memory");
                }
                mbar_wait(bars + 8 * (BAR_A_EMPTY + pw), (use & 1) ^ 1);
                asm volatile("cp.async.wait_group 0;" ::: "memory");
                __syncwarp();
                // 32-bit shared-window addresses and explicit ld.shared / st.shared: through generic pointers this loop took
                // ~180 cycles per item
                const uint32_t stage_s = sA + pw * A_STAGE_BYTES, sfa_s = sSFA + pw * SF_BYTES;
                const bool     ok      = ib < nbk;
#pragma unroll 8
                for (int jj = 0; jj < ((p.dbg & 1) ? 0 : 16); ++jj) {
                    const int      row  = (lane >> 2) + 8 * jj;
                    const uint32_t lead = (((lead0[jj >> 3] >> (4 * (jj & 7))) & 15u) + 4u * (uint32_t) kc) & 15u;
                    const uint32_t bp   = raw_s + (uint32_t) (row * RAW_PITCH) + lead + 17u * (uint32_t) ib;  // exponent byte, then 16 bytes of nibbles
                    const uint32_t e    = lds_u8(bp);
                    const uint32_t wa   = (bp + 1u) & ~3u;
                    const int      sh   = (int) ((bp + 1u) & 3u) * 8;
                    const uint32_t w0 = lds_u32(wa), w1 = lds_u32(wa + 4), w2 = lds_u32(wa + 8), w3 = lds_u32(wa + 12), w4 = lds_u32(wa + 16);
                    const uint32_t qs[4] = { __funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh) };
                    uint32_t lo2[2], hi2[2];
#pragma unroll
                    for (int hv = 0; hv < 2; ++hv) {  // two source words -> one word of 8 consecutive elements
                        const uint32_t a = qs[2 * hv], c = qs[2 * hv + 1];
                        const uint32_t la = a & 0x0F0F0F0Fu, lc = c & 0x0F0F0F0Fu;
                        const uint32_t ha = (a >> 4) & 0x0F0F0F0Fu, hc = (c >> 4) & 0x0F0F0F0Fu;
                        lo2[hv] = __byte_perm(la | (la >> 4), lc | (lc >> 4), 0x6420);
                        hi2[hv] = __byte_perm(ha | (ha >> 4), hc | (hc >> 4), 0x6420);
                    }
                    const uint32_t ra = stage_s + (uint32_t) (row * 128);
                    const uint32_t sw = (uint32_t) (row & 7);
                    sts_v2(ra + (((2u * ib) ^ sw) << 4), ok ? lo2[0] : 0u, ok ? lo2[1] : 0u);       // elements 0 .. 15
                    sts_v2(ra + (((2u * ib + 1u) ^ sw) << 4), ok ? hi2[0] : 0u, ok ? hi2[1] : 0u);  // elements 16 .. 31
                    sts_u8(sfa_s + (uint32_t) ((row & 31) * 16 + (row >> 5) * 4 + ib), ok ? e : 127u);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(bars + 8 * (BAR_A_FULL + pw));
                ++use;
            }
        }
    } else if (warp == WARP_BPROD) {
        // ------------------------------------------------------------------ activations: one bulk copy per chunk
        int      sb = 0;      // running ring position / phase (no runtime division in the issue loops)
        uint32_t bph = 0;
        for (int t = blockIdx.x; t < num_units; t += gridDim.x) {
            const Unit u = unit_of(p, t);
            for (int kc = 0; kc < p.k_chunks; ++kc) {
                mbar_wait(bars + 8 * (BAR_B_EMPTY + sb), bph ^ 1);
                if (elect_one()) {
                    if (p.dbg & 2) {
                        mbar_arrive(bars + 8 * (BAR_B_FULL + sb));
                    } else {
                        mbar_expect_tx(bars + 8 * (BAR_B_FULL + sb), B_STAGE_BYTES);
                        bulk_g2s(sB + sb * B_STAGE_BYTES, p.actp + (u.stage0 + kc) * (long long) B_STAGE_BYTES, B_STAGE_BYTES, bars + 8 * (BAR_B_FULL + sb));
                    }
                }
                __syncwarp();
                if (++sb == B_STAGES) { sb = 0; bph ^= 1; }
            }
        }
    } else if (warp == WARP_MMA) {
        // ------------------------------------------------------------------ scale factors -> TMEM, MMAs
        // The scale-factor copies of chunk g + 1 are issued BEFORE the MMAs of chunk g (two TMEM slots): a tcgen05.cp takes
        // ~1500 cycles until a dependent MMA can start — issued right in front of its own MMAs it made every chunk cost
        // ~2000 cycles instead of 512.
        const int my_units = blockIdx.x < num_units ? (num_units - (int) blockIdx.x + (int) gridDim.x - 1) / (int) gridDim.x : 0;
        const int total    = my_units * p.k_chunks;
        // running ring positions and phases: `c*` for the scale copies (one chunk ahead), `m*` for the MMAs
        int      csa = 0, csb = 0, cslot = 0, msa = 0, msb = 0, mslot = 0;
        uint32_t caph = 0, cbph = 0;
        auto copy_scales = [&]() {
            mbar_wait(bars + 8 * (BAR_A_FULL + csa), caph);
            mbar_wait(bars + 8 * (BAR_B_FULL + csb), cbph);
            tc_fence_after();
            if (!(p.dbg & 4) && elect_one()) {
                const uint32_t tsf = tmem_base + T_SF + (uint32_t) cslot * 16;
                const uint32_t bst = sB + csb * B_STAGE_BYTES;
                tmem_cp_sf(tsf, desc_k_noswizzle(sSFA + csa * SF_BYTES));
                tmem_cp_sf(tsf + 4, desc_k_noswizzle(bst + 2 * B_TILE_BYTES));
                tmem_cp_sf(tsf + 8, desc_k_noswizzle(bst + 2 * B_TILE_BYTES + SF_BYTES));
            }
            __syncwarp();
            if (++csa == A_STAGES) { csa = 0; caph ^= 1; }
            if (++csb == B_STAGES) { csb = 0; cbph ^= 1; }
            cslot ^= 1;
        };
        if (total > 0) copy_scales();
        int      N = 16, ab = 0, ui = 0, kc = 0;
        uint32_t acc = tmem_base;
        for (int g = 0; g < total; ++g) {
            if (kc == 0) {
                const Unit u = unit_of(p, (int) blockIdx.x + ui * (int) gridDim.x);
                ab  = ui & 1;
                N   = max(16, (u.ncols + 15) & ~15);
                acc = tmem_base + (ab ? T_ACC1 : 0);
                mbar_wait(bars + 8 * (BAR_ACC_EMPTY + ab), (uint32_t) ((ui >> 1) & 1) ^ 1);
                tc_fence_after();
            }
            if (g + 1 < total) copy_scales();
            if (elect_one()) {
                const uint32_t tsf = tmem_base + T_SF + (uint32_t) mslot * 16;
                const uint32_t bst = sB + msb * B_STAGE_BYTES;
                const uint64_t ad  = umma_desc_k_sw128(sA + msa * A_STAGE_BYTES);
                const uint64_t bh  = umma_desc_k_sw128(bst), bl = umma_desc_k_sw128(bst + B_TILE_BYTES);
#pragma unroll
                for (int kk = 0; kk < ((p.dbg & 8) ? 0 : 4); ++kk) {
                    const uint32_t id = idesc_mxf(N, kk);
                    umma_mxf(acc, ad + (uint64_t) (kk * 2), bh + (uint64_t) (kk * 2), id, tsf, tsf + 4, (kc | kk) ? 1u : 0u);
                    umma_mxf(acc, ad + (uint64_t) (kk * 2), bl + (uint64_t) (kk * 2), id, tsf, tsf + 8, 1u);
                }
                umma_commit(bars + 8 * (BAR_A_EMPTY + msa));
                umma_commit(bars + 8 * (BAR_B_EMPTY + msb));
                if (kc == p.k_chunks - 1) umma_commit(bars + 8 * (BAR_ACC_FULL + ab));
            }
            __syncwarp();
            if (++msa == A_STAGES) msa = 0;
            if (++msb == B_STAGES) msb = 0;
            mslot ^= 1;
            if (++kc == p.k_chunks) { kc = 0; ++ui; }
        }
    } else if (warp < 4) {
        // ------------------------------------------------------------------ epilogue: TMEM -> dst
        int ui = 0;
        for (int t = blockIdx.x; t < num_units; t += gridDim.x, ++ui) {
            const Unit u   = unit_of(p, t);
            const int  ab  = ui & 1;
            const int  row = u.mt * TM + warp * 32 + lane;
            mbar_wait(bars + 8 * (BAR_ACC_FULL + ab), (uint32_t) ((ui >> 1) & 1));
            tc_fence_after();
            for (int c0 = 0; c0 < u.ncols; c0 += 32) {
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t) (warp * 32) << 16) + (uint32_t) ((ab ? T_ACC1 : 0) + c0), v);
                if (row < p.m) {
                    if (p.moe) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (c0 + j < u.ncols) {
                                const int r  = __ldg(p.moe_rows + u.row_base + c0 + j);
                                const int tk = r / p.n_used, sl = r - tk * p.n_used;
                                *(float *) ((char *) p.dst + (long long) tk * p.d_nb2 + (long long) sl * p.d_nb1 + (long long) row * 4) = v[j];
                            }
                        }
                    } else {
                        char * dp = (char *) p.dst + (long long) (u.n0 + c0) * p.d_nb1 + (long long) row * 4;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (c0 + j < u.ncols) *(float *) dp = v[j];
                            dp += p.d_nb1;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * (BAR_ACC_EMPTY + ab));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == WARP_MMA) tmem_dealloc(tmem_base, TMEM_COLS);
}

constexpr int SMEM_BYTES = 1024 + A_STAGES * A_STAGE_BYTES + B_STAGES * B_STAGE_BYTES + A_STAGES * SF_BYTES + A_STAGES * RAW_BYTES + BAR_COUNT * 8 + 16;
static_assert(SMEM_BYTES <= 227 * 1024, "mmq_mxf: shared memory");

int set_smem_once() {
    static std::mutex               mu;
    static std::map<int, bool>      done;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    if (!done[dev]) {
        if (cudaFuncSetAttribute(mmq_mxf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) return -1;
        done[dev] = true;
    }
    return 0;
}

}  // namespace

bool mxf_supported(int type, int64_t k) {
    static const bool off = getenv("B200Q_NO_MXF") != nullptr;
    return !off && type == T_MXFP4 && k > 0 && k % 32 == 0;
}
size_t mxf_workspace_bytes(int64_t k, int64_t n) {
    return (size_t) ((n + TN - 1) / TN) * (size_t) ((k + KC - 1) / KC) * B_STAGE_BYTES;
}
size_t mxf_moe_workspace_bytes(int64_t k, int64_t n_tokens, int64_t n_used, int64_t n_expert) {
    const int64_t rows = n_tokens * n_used, max_tiles = (rows + TN - 1) / TN + n_expert;
    return mxf_tables(nullptr, (int) n_expert, (int) rows, (int) max_tiles).bytes + (size_t) max_tiles * (size_t) ((k + KC - 1) / KC) * B_STAGE_BYTES;
}

int mxf_launch(const MxfArgs & a, void * workspace, cudaStream_t st, int sms, std::string & err) {
    if (((uintptr_t) workspace & 15) != 0) { err = "mxf: workspace must be 16-byte aligned"; return B200Q_EINVAL; }
    if (set_smem_once()) { err = "mxf: cudaFuncSetAttribute failed"; return B200Q_ECUDA; }
    MxfParams p{};
    p.w = a.w; p.dst = a.dst; p.k = (int) a.k; p.m = (int) a.m; p.n = (int) a.n;
    p.k_chunks = (int) ((a.k + KC - 1) / KC);
    p.m_tiles  = (int) ((a.m + TM - 1) / TM);
    p.w_nb1 = a.w_nb1; p.d_nb1 = a.d_nb1; p.d_nb2 = a.d_nb2;
    {
        static const int dbg = getenv("B200Q_MXF_DBG") ? atoi(getenv("B200Q_MXF_DBG")) : 0;
        p.dbg = dbg;
    }
    int units;
    if (a.ids) {
        if (a.n_expert > 1024) { err = "mxf: more than 1024 experts"; return B200Q_ENOTSUP; }
        const int rows = (int) (a.n_tokens * a.n_used), max_tiles = (rows + TN - 1) / TN + (int) a.n_expert;
        MxfTables mt = mxf_tables(workspace, (int) a.n_expert, rows, max_tiles);
        RouteParams rp{ a.ids, a.ids_nb1, (int) a.n_tokens, (int) a.n_used, (int) a.n_expert, mt };
        mxf_route<<<1, 1024, ((size_t) a.n_expert + (size_t) std::min(rows, ROUTE_SMEM_IDS)) * sizeof(int), st>>>(rp);
        MoePackParams mp{ a.act, (uint8_t *) workspace + mt.bytes, mt, (int) a.k, p.k_chunks, (int) a.n_used, a.b_bcast ? 1 : 0, a.a_nb1, a.a_nb2 };
        mxf_moe_pack_act<<<sms * 16, 256, 0, st>>>(mp);
        p.actp = (const uint8_t *) workspace + mt.bytes;
        p.moe = mt.total; p.moe_tile_expert = mt.tile_expert; p.moe_tile_col0 = mt.tile_col0; p.moe_count = mt.count;
        p.moe_offset = mt.offset; p.moe_rows = mt.rows; p.n_used = (int) a.n_used; p.w_nb_expert = a.w_nb2;
        p.n_tiles = 0;
        units = max_tiles * p.m_tiles;  // upper bound: the kernel reads the real tile count on the device
    } else {
        p.n_tiles = (int) ((a.n + TN - 1) / TN);
        PackParams pp{ a.act, (uint8_t *) workspace, (int) a.k, (int) a.n, p.n_tiles, p.k_chunks, a.a_nb1 };
        const long long total = (long long) p.n_tiles * p.k_chunks * TN * 4 * 8;
        mxf_pack_act<<<(int) std::min<long long>((total + 255) / 256, (long long) sms * 16), 256, 0, st>>>(pp);
        p.actp = (const uint8_t *) workspace;
        units  = p.m_tiles * p.n_tiles;
    }
    const int grid = std::max(1, std::min(units, sms));
    mmq_mxf_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("mxf: launch failed: ") + cudaGetErrorString(e); return B200Q_ECUDA; }
    return 0;
}

}  // namespace b200q
