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
//     prologue: each CTA quantizes the activation columns once into shared memory, no second launch,
//     no global scratch buffer.
//   * consumer warps run integer dots with dp4a (units.cuh), two rows per warp pass so every activation
//     word read from shared memory is used twice, warp-shuffle reduction, fixed summation order
//     (bit-reproducible, which the row-split parity test relies on).
//   * launched with programmatic stream serialization: griddepcontrol.launch_dependents is issued at once
//     and, when the caller marks the weights immutable, the producer starts streaming weights BEFORE
//     griddepcontrol.wait, so the HBM pipe stays busy across back-to-back GEMVs of a decode step.
//
// Work decomposition: the (channel, row) space is cut into contiguous ranges, one per CTA (grid = a multiple
// of the SM count); a channel is one (i12, i13) batch element of MUL_MAT or one (token, slot) pair of
// MUL_MAT_ID (expert id read on the device from `ids`, no host round trip).
#include "gemv.cuh"
#include "units.cuh"

#include <cuda_runtime.h>

namespace b200q {

namespace {

constexpr int NW          = 8;   // consumer warps
constexpr int RB          = 2;   // rows per warp pass
constexpr int NTHREADS    = (NW + 1) * 32;
constexpr int MAX_STAGES  = 8;
constexpr int RED_FLOATS  = 2 * NW * RB * 8;

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t * bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t * bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t * bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(a),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// 1-D bulk copy global -> shared, completion signalled on an mbarrier (TMA, UBLKCP in SASS)
__device__ __forceinline__ void bulk_g2s(void * dst_smem, const void * src_gmem, uint32_t bytes, uint64_t * bar,
                                         uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(NW * 32) : "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

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
    } else {
        const int i13 = c / p.ne12;
        const int i12 = c - i13 * p.ne12;
        o.w           = (long long) (i12 / p.r2) * p.w_nb2 + (long long) (i13 / p.r3) * p.w_nb3;
        o.a           = (long long) i12 * p.a_nb2 + (long long) i13 * p.a_nb3;
        o.d           = (long long) i12 * p.d_nb2 + (long long) i13 * p.d_nb3;
    }
    return o;
}

// quantize NCOLS activation columns of length k (f32, global) into shared memory:
//   qa  int8  [NCOLS][k]          (as int8x4 words)
//   ds  float2[NCOLS][k/32]       (d, d * sum(q))
// 8 lanes share one 32-value block (float4 each); executed by the NW consumer warps only.
template <int NCOLS>
__device__ __forceinline__ void quantize_activations(const GemvParams & p, const char * act_base, int * qa, float2 * ds,
                                                     int tid) {
    const int chunks_per_col = p.k >> 2;
    const int total          = NCOLS * chunks_per_col;
    for (int base = 0; base < total; base += NW * 32) {
        const int  q      = base + tid;
        const bool active = q < total;
        const int  j      = active ? q / chunks_per_col : 0;
        const int  pos    = active ? q - j * chunks_per_col : 0;
        float4     v      = make_float4(0.f, 0.f, 0.f, 0.f);
        if (active) {
            const float * src = (const float *) (act_base + (long long) j * p.a_nb1) + 4 * pos;
            if (p.act_vec4) {
                v = *(const float4 *) src;
            } else {
                v.x = src[0];
                v.y = src[1];
                v.z = src[2];
                v.w = src[3];
            }
        }
        float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
        amax       = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax       = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        amax       = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
        const float d  = amax * (1.0f / 127.0f);
        const float id = amax > 0.0f ? 127.0f / amax : 0.0f;
        const int   q0 = __float2int_rn(v.x * id);
        const int   q1 = __float2int_rn(v.y * id);
        const int   q2 = __float2int_rn(v.z * id);
        const int   q3 = __float2int_rn(v.w * id);
        int         s  = q0 + q1 + q2 + q3;
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        if (active) {
            qa[q] = (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | ((q3 & 0xff) << 24);
            if ((pos & 7) == 0) {
                ds[j * (p.k >> 5) + (pos >> 3)] = make_float2(d, d * (float) s);
            }
        }
    }
}

template <int T, int NCOLS>
__global__ void __launch_bounds__(NTHREADS, NCOLS <= 4 ? 2 : 1) gemv_stream_kernel(const GemvParams p) {
    using U = Unit<T>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t * full   = (uint64_t *) smem;
    uint64_t * empty  = full + MAX_STAGES;
    float *    red    = (float *) (smem + 128);
    int *      qa     = (int *) (smem + 128 + RED_FLOATS * 4);
    float2 *   ds     = (float2 *) ((uint8_t *) qa + (size_t) NCOLS * p.k);
    uint8_t *  stages = smem + p.stage_off;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const long long rows_total = (long long) p.nch * p.m;
    const long long R0         = rows_total * blockIdx.x / gridDim.x;
    const long long R1         = rows_total * (blockIdx.x + 1) / gridDim.x;
    const int       pass_rows  = (NW * RB) / p.wpr;
    const int       stage_rows = pass_rows * p.passes;
    const int       S          = p.nstages;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NW);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    pdl_launch_dependents();

    if (warp == NW) {
        // ------------------------------------------------------------------ producer: one elected lane
        if (lane == 0) {
            if (!p.early_prefetch) {
                pdl_wait();
            }
            const uint64_t pol = l2_evict_first_policy();
            long long      cur = R0;
            for (int it = 0; cur < R1; ++it) {
                const int c     = (int) (cur / p.m);
                const int r     = (int) (cur - (long long) c * p.m);
                long long left  = R1 - cur;
                int       nrows = p.m - r < stage_rows ? p.m - r : stage_rows;
                nrows           = left < nrows ? (int) left : nrows;
                const int s     = it % S;
                if (it >= S) {
                    mbar_wait(&empty[s], ((it / S) & 1) ^ 1);
                }
                const ChannelOff o   = channel_offsets(p, c);
                const uint8_t *  src = p.w + o.w + (long long) r * p.w_nb1;
                uint8_t *        dst = stages + (size_t) s * p.stage_bytes;
                if (p.contiguous) {
                    const uint32_t lead  = (uint32_t) ((uintptr_t) src & 15);
                    const uint32_t bytes = (lead + (uint32_t) nrows * (uint32_t) p.w_nb1 + 15u) & ~15u;
                    mbar_expect_tx(&full[s], bytes);
                    bulk_g2s(dst, src - lead, bytes, &full[s], pol);
                } else {
                    uint32_t total = 0;
                    for (int rr = 0; rr < nrows; ++rr) {
                        const uint32_t lead = (uint32_t) ((uintptr_t) (src + (long long) rr * p.w_nb1) & 15);
                        total += (lead + (uint32_t) p.row_bytes + 15u) & ~15u;
                    }
                    mbar_expect_tx(&full[s], total);
                    for (int rr = 0; rr < nrows; ++rr) {
                        const uint8_t * a     = src + (long long) rr * p.w_nb1;
                        const uint32_t  lead  = (uint32_t) ((uintptr_t) a & 15);
                        const uint32_t  bytes = (lead + (uint32_t) p.row_bytes + 15u) & ~15u;
                        bulk_g2s(dst + (size_t) rr * p.smem_pitch, a - lead, bytes, &full[s], pol);
                    }
                }
                cur += nrows;
            }
        }
        return;
    }

    // ---------------------------------------------------------------------- consumers
    pdl_wait();  // activations (and dst) belong to the previous kernels of the stream

    const int tid  = threadIdx.x;  // 0 .. NW*32-1
    const int g    = warp / p.wpr;
    const int part = warp - g * p.wpr;
    const int nu   = p.k / U::ELEMS;

    int       cur_ch = -1;
    long long d_off  = 0;
    int       red_buf = 0;
    long long cur     = R0;
    for (int it = 0; cur < R1; ++it) {
        const int c     = (int) (cur / p.m);
        const int r     = (int) (cur - (long long) c * p.m);
        long long left  = R1 - cur;
        int       nrows = p.m - r < stage_rows ? p.m - r : stage_rows;
        nrows           = left < nrows ? (int) left : nrows;
        const int s     = it % S;

        const ChannelOff o = channel_offsets(p, c);
        if (c != cur_ch) {
            if (cur_ch >= 0) {
                consumer_bar();  // everyone is done reading the previous channel's activations
            }
            quantize_activations<NCOLS>(p, (const char *) p.act + o.a, qa, ds, tid);
            consumer_bar();
            cur_ch = c;
            d_off  = o.d;
        }

        mbar_wait(&full[s], (it / S) & 1);

        const uint8_t * src0  = p.w + o.w + (long long) r * p.w_nb1;
        const uint8_t * sbase = stages + (size_t) s * p.stage_bytes;

        for (int pass = 0; pass * pass_rows < nrows; ++pass) {
            const int row0 = pass * pass_rows + g * RB;
            float     acc[RB][NCOLS];
#pragma unroll
            for (int rr = 0; rr < RB; ++rr) {
#pragma unroll
                for (int j = 0; j < NCOLS; ++j) {
                    acc[rr][j] = 0.0f;
                }
            }
            const uint8_t * rowp[RB];
            bool            valid[RB];
#pragma unroll
            for (int rr = 0; rr < RB; ++rr) {
                const int row = row0 + rr;
                valid[rr]     = row < nrows;
                if (p.contiguous) {
                    rowp[rr] = sbase + ((uintptr_t) src0 & 15) + (size_t) row * p.w_nb1;
                } else {
                    rowp[rr] = sbase + (size_t) row * p.smem_pitch + ((uintptr_t) (src0 + (long long) row * p.w_nb1) & 15);
                }
            }
            if (valid[0]) {
                for (int u = part * 32 + lane; u < nu; u += 32 * p.wpr) {
                    int   wq[RB][U::NG * U::GW];
                    float ws[RB][U::NG], wm[RB][U::NG];
                    U::decode(rowp[0], u, wq[0], ws[0], wm[0]);
                    if (valid[1]) {
                        U::decode(rowp[1], u, wq[1], ws[1], wm[1]);
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
                        const int *    qaj = qa + j * (p.k >> 2);
                        const float2 * dsj = ds + j * (p.k >> 5);
#pragma unroll
                        for (int gi = 0; gi < U::NG; ++gi) {
                            const int * a = qaj + U::act_word(u, gi);
                            int         aw[U::GW];
#pragma unroll
                            for (int i4 = 0; i4 < U::GW / 4; ++i4) {
                                const int4 t  = *(const int4 *) (a + 4 * i4);
                                aw[4 * i4 + 0] = t.x;
                                aw[4 * i4 + 1] = t.y;
                                aw[4 * i4 + 2] = t.z;
                                aw[4 * i4 + 3] = t.w;
                            }
                            const float2 x = dsj[U::act_blk(u, gi)];
#pragma unroll
                            for (int rr = 0; rr < RB; ++rr) {
                                int sumi = 0;
#pragma unroll
                                for (int i = 0; i < U::GW; ++i) {
                                    sumi = dp4a_(wq[rr][gi * U::GW + i], aw[i], sumi);
                                }
                                acc[rr][j] += ws[rr][gi] * (x.x * (float) sumi) + wm[rr][gi] * x.y;
                            }
                        }
                    }
                }
            }
            // warp all-reduce, then lane (rr*NCOLS + j) owns output (row0+rr, column j)
            float mine = 0.0f;
#pragma unroll
            for (int rr = 0; rr < RB; ++rr) {
#pragma unroll
                for (int j = 0; j < NCOLS; ++j) {
                    float v = acc[rr][j];
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    if (lane == rr * NCOLS + j) {
                        mine = v;
                    }
                }
            }
            const int my_rr = lane / NCOLS;
            const int my_j  = lane - my_rr * NCOLS;
            if (p.wpr == 1) {
                if (lane < RB * NCOLS && row0 + my_rr < nrows) {
                    float * out = (float *) ((char *) p.dst + d_off + (long long) my_j * p.d_nb1) + (r + row0 + my_rr);
                    *out        = mine;
                }
            } else {
                float * rb = red + red_buf * (NW * RB * 8);
                if (lane < RB * NCOLS) {
                    rb[warp * (RB * 8) + lane] = mine;
                }
                consumer_bar();
                if (part == 0 && lane < RB * NCOLS && row0 + my_rr < nrows) {
                    float v = 0.0f;
                    for (int pp = 0; pp < p.wpr; ++pp) {
                        v += rb[(warp + pp) * (RB * 8) + lane];
                    }
                    float * out = (float *) ((char *) p.dst + d_off + (long long) my_j * p.d_nb1) + (r + row0 + my_rr);
                    *out        = v;
                }
                red_buf ^= 1;
            }
        }
        __syncwarp();
        if (lane == 0) {
            mbar_arrive(&empty[s]);
        }
        cur += nrows;
    }
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
        default: return nullptr;
    }
}

int gemv_threads() { return NTHREADS; }
int gemv_fixed_smem() { return 128 + RED_FLOATS * 4; }
int gemv_pass_rows(int wpr) { return (NW * RB) / wpr; }

}  // namespace b200q
