// gemv.cu — decode GEMV (n <= 8 activation columns) for GGUF block formats on sm_100a.
//
// Replaces the reference's quantize_q8_1 + mul_mat_vec_q pair (ggml-cuda/quantize.cu:4-48,
// ggml-cuda/mmvq.cu:139-226, host side mmvq.cu:500-570) and the ids-indexed GEMV used for MoE decode
// (mmvq.cu:163-166, ggml-cuda.cu:2101-2108) with ONE kernel designed around B200's HBM3e:
//
//   * the weight matrix is treated as a raw byte stream.  One elected producer thread per CTA moves
//     16-byte-aligned windows of whole rows HBM -> shared memory with cp.async.bulk (the 1-D TMA path,
//     SASS UBLKCP) into a ring of mbarrier-guarded stages, evict-first in L2.  17/18/34/210-byte blocks
//     never see a misaligned global access; lanes decode at arbitrary byte offsets from smem.
//   * the q8 activation quantization (per 32 values: int8 + fp32 scale + fp32 scale*sum) is fused into the
//     prologue: each CTA quantizes the activation columns once into shared memory (16-byte chunks XOR-swizzled
//     so the strided per-lane reads are bank-conflict free), no second launch, no global scratch buffer.
//   * 8 consumer warps run integer dots with dp4a (units.cuh), two rows per lane so every activation word
//     read from shared memory is used twice; fixed summation order (bit-reproducible, which the row-split
//     parity test relies on).  Short rows put several rows in one warp (lpr lanes per row), long rows put
//     several warps on one row pair (wpr), medium rows split the warps into groups that own alternate stages.
//   * one CTA per SM and <= ~100 KB of shared memory, launched with programmatic stream serialization:
//     griddepcontrol.launch_dependents is issued at once, the NEXT GEMV's CTAs become co-resident and — when
//     the caller marks the weights immutable — start streaming their weights BEFORE griddepcontrol.wait, so
//     the HBM pipe stays busy across the back-to-back GEMVs of a decode step.
//
// Work decomposition: the (channel, row) space is cut into contiguous ranges, one per CTA; a channel is one
// (i12, i13) batch element of MUL_MAT or one (token, slot) pair of MUL_MAT_ID (expert id read on the device
// from `ids`, no host round trip).  All shared-memory addressing is by 32-bit offset into b200q_smem.
#include "gemv.cuh"
#include "units.cuh"

#include <cuda_runtime.h>

namespace b200q {

namespace {

#ifndef B200Q_GEMV_NW
#define B200Q_GEMV_NW 16
#endif
constexpr int NW         = B200Q_GEMV_NW;  // consumer warps
constexpr int RB         = 2;  // rows per lane
constexpr int NTHREADS   = (NW + 1) * 32;
constexpr int MAX_STAGES = 8;
// shared memory map (byte offsets)
constexpr uint32_t OFF_FULL  = 0;     // MAX_STAGES mbarriers
constexpr uint32_t OFF_EMPTY = 64;    // MAX_STAGES mbarriers
constexpr uint32_t OFF_ARDY  = 128;   // mbarrier: the consumers have issued their first activation loads
constexpr uint32_t OFF_RED   = 144;   // [2][NW][RB*8] floats for the cross-warp reduction
constexpr uint32_t OFF_ACT   = 144 + 2 * NW * RB * 8 * 4;  // quantized activations

__device__ __forceinline__ uint32_t smem_base_u32() { return (uint32_t) __cvta_generic_to_shared(b200q_smem); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// 1-D bulk copy global -> shared, completion signalled on an mbarrier (TMA, UBLKCP in SASS)
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void * src_gmem, uint32_t bytes, uint32_t bar,
                                         uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(dst_smem),
        "l"(src_gmem), "r"(bytes), "r"(bar), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void named_bar(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void stamp(const GemvParams & p, int slot) {
    if (p.ts) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.ts[blockIdx.x * 8 + slot] = t;
    }
}

struct ChannelOff {
    long long w, a, d;
};

__device__ __forceinline__ ChannelOff channel_offsets(const GemvParams & p, int c) {
    ChannelOff o;
    if (p.mode == GEMV_MODE_MOE) {
        const int t = c / p.n_used;
        const int s = c - t * p.n_used;
        int       e = *(const int32_t *) ((const char *) p.ids + (long long) t * p.ids_nb1 + (long long) s * 4);
        e           = e < 0 ? 0 : (e >= p.n_expert ? p.n_expert - 1 : e);
        o.w         = (long long) e * p.w_nb2;
        o.a         = (long long) t * p.a_nb2 + (p.b_bcast ? 0 : (long long) s * p.a_nb1);
        o.d         = (long long) t * p.d_nb2 + (long long) s * p.d_nb1;
    } else if (p.nch == 1) {
        o.w = o.a = o.d = 0;
    } else {
        const int i13 = c / p.ne12;
        const int i12 = c - i13 * p.ne12;
        o.w           = (long long) (i12 / p.r2) * p.w_nb2 + (long long) (i13 / p.r3) * p.w_nb3;
        o.a           = (long long) i12 * p.a_nb2 + (long long) i13 * p.a_nb3;
        o.d           = (long long) i12 * p.d_nb2 + (long long) i13 * p.d_nb3;
    }
    return o;
}

// physical word index of logical activation word `widx` inside one column: 16-byte chunks are XOR-swizzled
// within aligned groups of 8 so that lanes reading chunks at a stride of 2 or 4 hit 8 different bank groups
__device__ __forceinline__ uint32_t act_swz_chunk(uint32_t chunk) { return chunk ^ ((chunk >> 3) & 7u); }

// quantize NCOLS activation columns of length k (f32, global) into shared memory:
//   int8x4 words at OFF_ACT + j * act_col_bytes (swizzled), (d, d*sum(q)) float2 at ds_off + (j*k/32 + blk)*8
// 8 lanes share one 32-value block (float4 each); executed by the NW consumer warps only.
template <int NCOLS>
__device__ __forceinline__ void quantize_activations(const GemvParams & p, const char * act_base, int tid, bool signal) {
    constexpr int QB         = 4;         // iterations whose global loads are issued together (latency paid once)
    const int chunks_per_col = p.k >> 2;  // float4 chunks == int8x4 words
    const int total          = NCOLS * chunks_per_col;
    const int nblk           = p.k >> 5;
    for (int base = 0; base < total; base += QB * NW * 32) {
        float4 v[QB];
        int    jj[QB], pp[QB];
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            const int  q      = base + b * NW * 32 + tid;
            const bool active = q < total;
            int        j      = 0;
            int        pos    = active ? q : 0;
            if (NCOLS > 1) {
                j   = pos / chunks_per_col;
                pos = pos - j * chunks_per_col;
            }
            jj[b] = active ? j : -1;
            pp[b] = pos;
            v[b]  = make_float4(0.f, 0.f, 0.f, 0.f);
            if (active) {
                const float * src = (const float *) (act_base + (long long) j * p.a_nb1) + 4 * pos;
                if (p.act_vec4) {
                    v[b] = __ldg((const float4 *) src);
                } else {
                    v[b].x = __ldg(src);
                    v[b].y = __ldg(src + 1);
                    v[b].z = __ldg(src + 2);
                    v[b].w = __ldg(src + 3);
                }
            }
        }
        if (signal && base == 0 && (tid & 31) == 0) {
            // the first activation loads are in flight: the producer may now queue the rest of the weight stream
            mbar_arrive(smem_base_u32() + OFF_ARDY);
        }
        if (p.ts && tid == 0 && base == 0) {
            // debug: time at which the first activation value is usable
            if (v[0].x == 123456.0f) stamp(p, 7);
            stamp(p, 5);
        }
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            if (base + b * NW * 32 >= total) break;  // warp-uniform
            float amax = fmaxf(fmaxf(fabsf(v[b].x), fabsf(v[b].y)), fmaxf(fabsf(v[b].z), fabsf(v[b].w)));
            amax       = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax       = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            amax       = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
            const float d  = amax * (1.0f / 127.0f);
            const float id = amax > 0.0f ? 127.0f / amax : 0.0f;
            const int   q0 = __float2int_rn(v[b].x * id);
            const int   q1 = __float2int_rn(v[b].y * id);
            const int   q2 = __float2int_rn(v[b].z * id);
            const int   q3 = __float2int_rn(v[b].w * id);
            int         s  = q0 + q1 + q2 + q3;
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            if (jj[b] >= 0) {
                const int      pos  = pp[b];
                const uint32_t word = (act_swz_chunk((uint32_t) pos >> 2) << 2) | ((uint32_t) pos & 3u);
                *(int *) (b200q_smem + OFF_ACT + (uint32_t) jj[b] * p.act_col_bytes + word * 4) =
                    (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | ((q3 & 0xff) << 24);
                if ((pos & 7) == 0) {
                    *(float2 *) (b200q_smem + p.ds_off + (uint32_t) (jj[b] * nblk + (pos >> 3)) * 8) =
                        make_float2(d, d * (float) s);
                }
            }
        }
    }
}

// the stage walk shared by producer and consumers: (channel, first row, row count) of consecutive stages
struct StageWalk {
    long long cur, end;
    int       c, r, m, stage_rows;
    __device__ __forceinline__ void init(long long R0, long long R1, int m_, int stage_rows_) {
        cur = R0; end = R1; m = m_; stage_rows = stage_rows_;
        c = (int) (R0 / m_);
        r = (int) (R0 - (long long) c * m_);
    }
    __device__ __forceinline__ bool done() const { return cur >= end; }
    __device__ __forceinline__ int  nrows() const {
        int       n    = m - r < stage_rows ? m - r : stage_rows;
        long long left = end - cur;
        return left < n ? (int) left : n;
    }
    __device__ __forceinline__ void advance(int n) {
        cur += n;
        r += n;
        if (r >= m) { r = 0; ++c; }
    }
};

template <int T, int NCOLS>
__global__ void __launch_bounds__(NTHREADS, NW <= 8 ? 2 : 1) gemv_stream_kernel(const GemvParams p) {
    using U = Unit<T>;
    const uint32_t sb   = smem_base_u32();
    const int      warp = threadIdx.x >> 5;
    const int      lane = threadIdx.x & 31;
    const int      S    = p.nstages;
    const int      G    = p.groups;
    const int      WS   = NW / G;  // warps per group

    const long long rows_total = (long long) p.nch * p.m;
    const long long R0         = rows_total * blockIdx.x / gridDim.x;
    const long long R1         = rows_total * (blockIdx.x + 1) / gridDim.x;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(sb + OFF_FULL + 8 * s, 1);
            mbar_init(sb + OFF_EMPTY + 8 * s, WS);
        }
        mbar_init(sb + OFF_ARDY, NW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) stamp(p, 0);
    pdl_launch_dependents();

    StageWalk walk;
    walk.init(R0, R1, p.m, p.stage_rows);

    if (warp == NW) {
        // ------------------------------------------------------------------ producer: one elected lane
        if (lane == 0) {
            if (!p.early_prefetch) {
                pdl_wait();
            }
            const uint64_t pol = l2_evict_first_policy();
            int            s = 0, round = 0, issued = 0;
            while (!walk.done()) {
                const int nrows = walk.nrows();
                if (issued == p.early_stages) {
                    // only `early_stages` stages are queued ahead of the activation loads (a deeper queue would
                    // delay them by the time the SM needs to ingest it); the rest follows once those are issued
                    pdl_wait();
                    mbar_wait(sb + OFF_ARDY, 0);
                }
                ++issued;
                if (round > 0) {
                    mbar_wait(sb + OFF_EMPTY + 8 * s, (round & 1) ^ 1);
                }
                const ChannelOff o    = channel_offsets(p, walk.c);
                const uint8_t *  src  = p.w + o.w + (long long) walk.r * p.w_nb1;
                const uint32_t   dst  = sb + p.stage_off + (uint32_t) s * p.stage_bytes;
                const uint32_t   full = sb + OFF_FULL + 8 * s;
                if (p.contiguous) {
                    const uint32_t lead  = (uint32_t) ((uintptr_t) src & 15);
                    const uint32_t bytes = (lead + (uint32_t) nrows * (uint32_t) p.w_nb1 + 15u) & ~15u;
                    mbar_expect_tx(full, bytes);
                    bulk_g2s(dst, src - lead, bytes, full, pol);
                } else {
                    uint32_t total = 0;
                    for (int rr = 0; rr < nrows; ++rr) {
                        const uint32_t lead = (uint32_t) ((uintptr_t) (src + (long long) rr * p.w_nb1) & 15);
                        total += (lead + (uint32_t) p.row_bytes + 15u) & ~15u;
                    }
                    mbar_expect_tx(full, total);
                    for (int rr = 0; rr < nrows; ++rr) {
                        const uint8_t * a     = src + (long long) rr * p.w_nb1;
                        const uint32_t  lead  = (uint32_t) ((uintptr_t) a & 15);
                        const uint32_t  bytes = (lead + (uint32_t) p.row_bytes + 15u) & ~15u;
                        bulk_g2s(dst + (uint32_t) rr * p.smem_pitch, a - lead, bytes, full, pol);
                    }
                }
                walk.advance(nrows);
                if (++s == S) {
                    s = 0;
                    ++round;
                }
            }
        }
        return;
    }

    // ---------------------------------------------------------------------- consumers
    if (p.mode == GEMV_MODE_BATCH && !walk.done()) {
        // warm the TLB / L2 path of the activation vector while the previous kernel is still running
        // (L2 is the point of coherence, so a line fetched early is still correct after griddepcontrol.wait)
        const ChannelOff o0  = channel_offsets(p, walk.c);
        const int        idx = threadIdx.x * 32;  // one 128-byte line per thread
        if (idx < p.k) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *) p.act + o0.a + (long long) idx * 4));
        }
    }
    pdl_wait();  // activations (and dst) belong to the previous kernels of the stream
    if (threadIdx.x == 0) stamp(p, 1);

    const int tid   = threadIdx.x;       // 0 .. NW*32-1
    const int gi    = warp / WS;         // group
    const int wi    = warp - gi * WS;    // warp inside the group
    const int slot  = wi / p.wpr;        // row-pair slot inside a pass
    const int part  = wi - slot * p.wpr; // which slice of the row this warp sums
    const int lpr   = p.lpr;
    const int RW    = 32 / lpr;          // rows per warp step (per rr)
    const int sub   = lane & (lpr - 1);
    const int lrow  = lane / lpr;
    const int nu    = p.k / U::ELEMS;
    const int ustep = lpr * p.wpr;
    const int pass_rows = (WS / p.wpr) * RW * RB;
    const int nblk  = p.k >> 5;

    int       cur_ch  = -1;
    long long d_off   = 0;
    uint32_t  lead0   = 0;
    int       red_buf = 0;
    int       s = 0, round = 0, gsel = 0;
    while (!walk.done()) {
        const int nrows = walk.nrows();
        const int c = walk.c, r = walk.r;
        if (c != cur_ch) {
            const ChannelOff o = channel_offsets(p, c);
            if (cur_ch >= 0) {
                named_bar(9, NW * 32);  // everyone is done reading the previous channel's activations
            }
            quantize_activations<NCOLS>(p, (const char *) p.act + o.a, tid, cur_ch < 0);
            if (threadIdx.x == 0 && cur_ch < 0) stamp(p, 6);
            named_bar(9, NW * 32);
            if (threadIdx.x == 0 && cur_ch < 0) stamp(p, 2);
            cur_ch = c;
            d_off  = o.d;
            lead0  = (uint32_t) ((uintptr_t) (p.w + o.w) & 15);
        }
        if (gsel == gi) {
            mbar_wait(sb + OFF_FULL + 8 * s, round & 1);
            if (threadIdx.x == 0 && round == 0 && s == 0) stamp(p, 3);
            const uint32_t stage = p.stage_off + (uint32_t) s * p.stage_bytes;
            // byte offset (inside b200q_smem) of row `row` of this stage
            const uint32_t row_lead = (lead0 + (uint32_t) (((long long) r * p.w_nb1) & 15)) & 15u;

            for (int pass = 0; pass * pass_rows < nrows; ++pass) {
                const int row_a = pass * pass_rows + slot * (RW * RB) + lrow;  // rr = 0; rr = 1 is RW rows further
                float     acc[RB][NCOLS];
#pragma unroll
                for (int rr = 0; rr < RB; ++rr) {
#pragma unroll
                    for (int j = 0; j < NCOLS; ++j) {
                        acc[rr][j] = 0.0f;
                    }
                }
                const bool v0 = row_a < nrows;
                const bool v1 = row_a + RW < nrows;
                uint32_t   rowp0, rowp1;
                if (p.contiguous) {
                    rowp0 = stage + row_lead + (uint32_t) row_a * (uint32_t) p.smem_pitch;
                    rowp1 = rowp0 + (uint32_t) RW * (uint32_t) p.smem_pitch;
                } else {
                    const uint32_t l0 = (lead0 + (uint32_t) (((long long) (r + row_a) * p.w_nb1) & 15)) & 15u;
                    const uint32_t l1 = (lead0 + (uint32_t) (((long long) (r + row_a + RW) * p.w_nb1) & 15)) & 15u;
                    rowp0             = stage + (uint32_t) row_a * (uint32_t) p.smem_pitch + l0;
                    rowp1             = stage + (uint32_t) (row_a + RW) * (uint32_t) p.smem_pitch + l1;
                }
                if (v0) {
                    for (int u = part * lpr + sub; u < nu; u += ustep) {
                        int   wq[RB][U::NG * U::GW];
                        float ws[RB][U::NG], wm[RB][U::NG];
                        U::decode(rowp0, u, wq[0], ws[0], wm[0]);
                        if (v1) {
                            U::decode(rowp1, u, wq[1], ws[1], wm[1]);
                        } else {
#pragma unroll
                            for (int i = 0; i < U::NG * U::GW; ++i) {
                                wq[1][i] = 0;
                            }
#pragma unroll
                            for (int i = 0; i < U::NG; ++i) {
                                ws[1][i] = 0.0f;
                                wm[1][i] = 0.0f;
                            }
                        }
#pragma unroll
                        for (int j = 0; j < NCOLS; ++j) {
                            const uint32_t colb = OFF_ACT + (uint32_t) j * p.act_col_bytes;
#pragma unroll
                            for (int g = 0; g < U::NG; ++g) {
                                const uint32_t chunk0 = (uint32_t) U::act_word(u, g) >> 2;
                                int            aw[U::GW];
#pragma unroll
                                for (int i4 = 0; i4 < U::GW / 4; ++i4) {
                                    const int4 t   = lds128(colb + act_swz_chunk(chunk0 + i4) * 16);
                                    aw[4 * i4 + 0] = t.x;
                                    aw[4 * i4 + 1] = t.y;
                                    aw[4 * i4 + 2] = t.z;
                                    aw[4 * i4 + 3] = t.w;
                                }
                                float2 x = lds64f(p.ds_off + (uint32_t) (j * nblk + U::act_blk(u, g)) * 8);
                                if (U::MIN16) {  // the group's min applies to 16 activations: d_x * sum of this group's q_x
                                    int asum = 0;
#pragma unroll
                                    for (int i = 0; i < U::GW; ++i) asum = dp4a_(0x01010101, aw[i], asum);
                                    x.y = x.x * (float) asum;
                                }
#pragma unroll
                                for (int rr = 0; rr < RB; ++rr) {
                                    int s0 = 0, s1 = 0;
#pragma unroll
                                    for (int i = 0; i < U::GW; i += 2) {
                                        if (U::UNSIGNED_W) {
                                            s0 = dp4a_us((uint32_t) wq[rr][g * U::GW + i], aw[i], s0);
                                            s1 = dp4a_us((uint32_t) wq[rr][g * U::GW + i + 1], aw[i + 1], s1);
                                        } else {
                                            s0 = dp4a_(wq[rr][g * U::GW + i], aw[i], s0);
                                            s1 = dp4a_(wq[rr][g * U::GW + i + 1], aw[i + 1], s1);
                                        }
                                    }
                                    acc[rr][j] += ws[rr][g] * (x.x * (float) (s0 + s1)) + wm[rr][g] * x.y;
                                }
                            }
                        }
                    }
                }
                // reduce over the lpr lanes of each row
#pragma unroll
                for (int rr = 0; rr < RB; ++rr) {
#pragma unroll
                    for (int j = 0; j < NCOLS; ++j) {
                        float v = acc[rr][j];
                        if (lpr == 32) {
                            v += __shfl_xor_sync(0xffffffffu, v, 16);
                            v += __shfl_xor_sync(0xffffffffu, v, 8);
                            v += __shfl_xor_sync(0xffffffffu, v, 4);
                            v += __shfl_xor_sync(0xffffffffu, v, 2);
                            v += __shfl_xor_sync(0xffffffffu, v, 1);
                        } else {
                            for (int o = lpr >> 1; o > 0; o >>= 1) {
                                v += __shfl_xor_sync(0xffffffffu, v, o);
                            }
                        }
                        acc[rr][j] = v;
                    }
                }
                if (p.wpr == 1) {
                    if (sub == 0) {
#pragma unroll
                        for (int rr = 0; rr < RB; ++rr) {
                            const int row = row_a + rr * RW;
                            if (row < nrows) {
#pragma unroll
                                for (int j = 0; j < NCOLS; ++j) {
                                    float * out = (float *) ((char *) p.dst + d_off + (long long) j * p.d_nb1) + (r + row);
                                    *out        = acc[rr][j];
                                }
                            }
                        }
                    }
                } else {
                    // lpr == 32: lane 0 holds the warp's partial sums; the group's warps exchange them through smem
                    float * rb = (float *) (b200q_smem + OFF_RED) + red_buf * (NW * RB * 8);
                    if (lane == 0) {
#pragma unroll
                        for (int rr = 0; rr < RB; ++rr) {
#pragma unroll
                            for (int j = 0; j < NCOLS; ++j) {
                                rb[warp * (RB * 8) + rr * 8 + j] = acc[rr][j];
                            }
                        }
                    }
                    named_bar(1 + gi, WS * 32);
                    if (part == 0 && lane < RB * 8) {
                        const int rr  = lane >> 3;
                        const int j   = lane & 7;
                        const int row = row_a + rr * RW;
                        if (j < NCOLS && row < nrows) {
                            float v = 0.0f;
                            for (int pp = 0; pp < p.wpr; ++pp) {
                                v += rb[(warp + pp) * (RB * 8) + lane];
                            }
                            float * out = (float *) ((char *) p.dst + d_off + (long long) j * p.d_nb1) + (r + row);
                            *out        = v;
                        }
                    }
                    red_buf ^= 1;
                }
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sb + OFF_EMPTY + 8 * s);
            }
        }
        walk.advance(nrows);
        if (++s == S) {
            s = 0;
            ++round;
        }
        if (++gsel == G) {
            gsel = 0;
        }
    }
    if (threadIdx.x == 0) stamp(p, 4);
}

template <int T> gemv_kernel_t pick_ncols(int ncols) {
    switch (ncols) {
        case 1: return gemv_stream_kernel<T, 1>;
        case 2: return gemv_stream_kernel<T, 2>;
        case 3: return gemv_stream_kernel<T, 3>;
        case 4: return gemv_stream_kernel<T, 4>;
        case 5: return gemv_stream_kernel<T, 5>;
        case 6: return gemv_stream_kernel<T, 6>;
        case 7: return gemv_stream_kernel<T, 7>;
        case 8: return gemv_stream_kernel<T, 8>;
        default: return nullptr;
    }
}

}  // namespace

gemv_kernel_t gemv_kernel_for(int type, int ncols) {
    switch (type) {
        case T_Q4_0:  return pick_ncols<T_Q4_0>(ncols);
        case T_Q4_1:  return pick_ncols<T_Q4_1>(ncols);
        case T_Q5_0:  return pick_ncols<T_Q5_0>(ncols);
        case T_Q5_1:  return pick_ncols<T_Q5_1>(ncols);
        case T_Q8_0:  return pick_ncols<T_Q8_0>(ncols);
        case T_MXFP4: return pick_ncols<T_MXFP4>(ncols);
        case T_Q4_K:  return pick_ncols<T_Q4_K>(ncols);
        case T_Q5_K:  return pick_ncols<T_Q5_K>(ncols);
        case T_Q6_K:  return pick_ncols<T_Q6_K>(ncols);
        case T_Q2_K:  return pick_ncols<T_Q2_K>(ncols);
        case T_Q3_K:  return pick_ncols<T_Q3_K>(ncols);
        case T_IQ4_NL: return pick_ncols<T_IQ4_NL>(ncols);
        case T_IQ4_XS: return pick_ncols<T_IQ4_XS>(ncols);
        default: return nullptr;
    }
}

int gemv_threads() { return NTHREADS; }
int gemv_consumer_warps() { return NW; }
int gemv_act_off() { return (int) OFF_ACT; }
int gemv_unit_elems(int type) {
    switch (type) {
        case T_Q4_K: case T_Q5_K: case T_Q6_K: case T_Q2_K: case T_Q3_K: return 64;
        default: return 32;
    }
}

}  // namespace b200q
