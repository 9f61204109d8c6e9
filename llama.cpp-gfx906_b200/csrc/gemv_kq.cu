// gemv_kq.cu — decode GEMV (n = 1): K-quants (q4_K, q6_K — the formats that carry a Q4_K_M model) and the 32-element
// block formats (q4_0 q4_1 q5_0 q5_1 q8_0 mxfp4); MUL_MAT_ID for one token; row-split peer stores.
//
// Same role as gemv.cu (it replaces the reference's quantize_q8_1 + mul_mat_vec_q pair, ggml-cuda/quantize.cu:4-48,
// mmvq.cu:139-226), redesigned around what the B200 timeline of gemv.cu showed (profiles/r01_gemv_timeline_v1.txt):
//
//   * 8 warps, <= 110 registers, <= ~110 KB of shared memory per CTA, one CTA per SM per launch, so that TWO launches
//     are co-resident on every SM under programmatic dependent launch: while kernel i computes, kernel i+1 has already
//     filled its ring with ~100 KB of weights per SM, i.e. the HBM pipe keeps streaming across the launch boundary.
//   * activations are quantized per 256-element super-block (one fp32 scale + int16 partial sums per 16 / 32 values),
//     the arithmetic ggml-cpu itself uses for these formats (quantize_row_q8_K + ggml_vec_dot_q4_K_q8_K /
//     _q6_K_q8_K, ggml-cpu/quants.c:550-623, :705-758): the 6-bit sub-block scales then multiply INTEGER sums
//     (IMAD / dp2a) instead of costing an I2F + FFMA per 32 values, and the kernel needs ~0.7 (q4_K) / ~1.6 (q6_K)
//     instructions per weight instead of ~2.3.
//   * a lane owns whole super-blocks (LDS.128 for q4_K; q6_K's 210-byte blocks through aligned words + funnel shift),
//     lanes of a row read consecutive 16-byte activation chunks (chunk-major layout: conflict-free), one row per lane
//     group, one warp pass = one ring stage; every warp owns its ring slots and refills them itself (no producer warp,
//     no CTA-wide barrier in the main loop).
//   * up to 4 matrices that share the activation vector (q/k/v, gate/up) run in ONE launch ("segments"): one
//     activation load + quantization, one launch hop.
// Fixed summation order: results are bit-reproducible and independent of the grid (row-split relies on it).
#include "gemv_kq.cuh"

#include "blocks.cuh"
#include "kq_dot.cuh"
#include "sm100.cuh"

#include <cuda_runtime.h>

namespace b200q {

namespace {

using namespace sm100;
using namespace kqdot;

constexpr int NTHREADS = NW * 32;
constexpr int MAX_S    = 32;             // ring depth limit
constexpr uint32_t OFF_FULL = 0;  // full-barriers (8 B each, MAX_S of them); data regions start at 384

extern __shared__ __align__(128) uint8_t kq_smem[];

// development aid, compiled only with -DB200Q_DEBUG_HANG: a host-mapped buffer that stuck waiters report into
// (b200q_debug_hang_buffer); the release kernel neither reads the pointer nor spins with a counter
#ifdef B200Q_DEBUG_HANG
__device__ unsigned long long * g_kq_hang = nullptr;
__device__ __forceinline__ void mbar_wait_dbg(uint32_t bar, uint32_t parity, int tag, int j) {
    unsigned long long * hb = g_kq_hang;
    if (!hb) {
        mbar_wait(bar, parity);
        return;
    }
    unsigned spins = 0;
    for (;;) {
        uint32_t ok;
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
        if (++spins == (1u << 20)) {
            const unsigned long long idx = atomicAdd(hb, 1ull);
            if (idx < 500) {
                unsigned long long st;
                asm volatile("ld.shared.b64 %0, [%1];" : "=l"(st) : "r"(bar));
                hb[1 + idx * 4 + 0] = ((unsigned long long) blockIdx.x << 32) | threadIdx.x;
                hb[1 + idx * 4 + 1] = ((unsigned long long) tag << 32) | (unsigned) j;
                hb[1 + idx * 4 + 2] = parity;
                hb[1 + idx * 4 + 3] = st;
                __threadfence_system();
            }
        }
    }
}
#define KQ_HANG_COUNT(i) do { if (g_kq_hang && threadIdx.x == 0) atomicAdd(g_kq_hang + (i), 1ull); } while (0)
#else
__device__ __forceinline__ void mbar_wait_dbg(uint32_t bar, uint32_t parity, int, int) { mbar_wait(bar, parity); }
#define KQ_HANG_COUNT(i) do { } while (0)
#endif

template <int T> struct IsKQuant { static constexpr bool value = T == T_Q4_K || T == T_Q5_K || T == T_Q6_K; };

struct StageInfo {
    int g, r0, nrows;
};
template <int NSEG> __device__ __forceinline__ StageInfo stage_info(const KqParams & p, int t) {
    StageInfo si;
    si.g = 0;
#pragma unroll
    for (int i = 1; i < NSEG; ++i) {
        if (i < p.nseg && t >= p.seg[i].stage0) si.g = i;
    }
    si.r0    = (t - p.seg[si.g].stage0) * p.rows_per_stage;
    si.nrows = min(p.rows_per_stage, p.seg[si.g].m - si.r0);
    return si;
}

__device__ __forceinline__ const uint8_t * seg_ptr(const uint8_t * const (&w)[KQ_MAX_SEG], int g) {
    const uint8_t * r = w[0];
#pragma unroll
    for (int i = 1; i < KQ_MAX_SEG; ++i) {
        if (g == i) r = w[i];
    }
    return r;
}

// MOE = false: up to KQ_PLAIN_SEG matrices sharing one activation vector; MOE = true: up to KQ_MAX_SEG experts chosen by
// ids[], optionally one activation vector per expert (kept out of the plain instantiation: its prologue is on the
// latency chain of every decode launch)
constexpr int KQ_PLAIN_SEG = 4;
template <int T, bool MOE> __global__ void __launch_bounds__(NTHREADS, 2) gemv_kq_kernel(const KqParams p) {
    constexpr int NSEG = MOE ? KQ_MAX_SEG : KQ_PLAIN_SEG;
    const uint32_t sb   = smem_u32(kq_smem);
    const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int      S    = p.nstages;
    const int      NA   = min(NW, S);  // warps that own slots (very long rows: fewer slots than warps fit)
    // ring depth and first slot of this warp (the S % NA spare slots go to the first warps)
    const int      D     = warp < NA ? S / NA + (warp < S % NA ? 1 : 0) : 1;
    const int      slot0 = warp * (S / NA) + min(warp, S % NA);
    // this CTA's contiguous range of stages
    const int T0 = (int) ((long long) p.total_stages * blockIdx.x / gridDim.x);
    const int T1 = (int) ((long long) p.total_stages * (blockIdx.x + 1) / gridDim.x);
    KQ_HANG_COUNT(2041);  // CTAs started

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(sb + OFF_FULL + 8 * s, 1);
        }
        mbar_fence_init();
    }
    __syncthreads();
    pdl_launch_dependents();  // the next launch may become co-resident and start prefetching its weights

    // Every warp feeds itself: lane 0 bulk-copies the warp's own next stages into the warp's own D slots (slot ownership:
    // warp w owns D consecutive slots), so successive uses of a slot's full-barrier are waited on by ONE warp in order.
    // (A ring shared by all warps let a fast warp test a slot's barrier one phase early — parity ABA — when bulk copies
    // completed out of order: stale data, a desynchronised ring, rare hangs / launch failures.)  No empty-barriers and no
    // producer hop: the warp that finishes a slot refills it at once.
    // MoE: the experts are chosen by ids[] (written by an earlier kernel of the stream: read only after the wait)
    const uint8_t * segw[KQ_MAX_SEG];
    long long       eoff[KQ_MAX_SEG];  // MoE with a bias: the segment's expert * bias_stride
    if (MOE) {
#pragma unroll
        for (int g = 0; g < KQ_MAX_SEG; ++g) {
            segw[g] = p.seg[g].w;
            eoff[g] = 0;
        }
    }
    auto seg_w = [&](int g) -> const uint8_t * { return MOE ? seg_ptr(segw, g) : p.seg[g].w; };
    const uint64_t pol    = policy_evict_first();
    const int      nmine  = warp < NA && T0 + warp < T1 ? (T1 - T0 - warp + NA - 1) / NA : 0;  // stages of this warp
    auto           issue  = [&](int i) {
        const int       t     = T0 + warp + i * NA, slot = slot0 + i % D;
        const StageInfo si    = stage_info<NSEG>(p, t);
        // rows of the 32-element block formats need not be 16-byte multiples: copy the aligned window around the rows
        // (the over-read stays inside the tensor or its allocation slack, include/b200q.h)
        const uint8_t * src   = seg_w(si.g) + (size_t) si.r0 * p.row_bytes;
        const uint32_t  lead  = (uint32_t) ((uintptr_t) src & 15);
        const uint32_t  bytes = (lead + (uint32_t) si.nrows * (uint32_t) p.row_bytes + 15u) & ~15u;
        const uint32_t  full  = sb + OFF_FULL + 8 * slot;
        mbar_expect_tx(full, bytes);
        bulk_g2s_hint(sb + p.off_stage + (uint32_t) slot * p.stage_bytes, src - lead, bytes, full, pol);
    };
    if (p.early_prefetch && lane == 0) {
        for (int i = 0; i < min(D, nmine); ++i) issue(i);  // immutable weights: stream them under the previous kernel's tail
    }
    pdl_wait();  // the activations (and dst) belong to the previous kernels of the stream
    if (MOE && p.ids) {
#pragma unroll
        for (int g = 0; g < KQ_MAX_SEG; ++g) {
            if (g < p.nseg) {
                const long long e = (long long) __ldg(p.ids + g);
                segw[g] = p.seg[g].w + e * p.expert_stride;
                eoff[g] = e * p.bias_stride;
            }
        }
    }
    if (!p.early_prefetch && lane == 0) {
        for (int i = 0; i < min(D, nmine); ++i) issue(i);
    }
    for (int g = 0; g < (MOE && p.act_smem_stride ? p.nseg : 1); ++g) {
        const float * av = p.act + (size_t) g * p.act_seg_stride;
        if (IsKQuant<T>::value) quantize_act(p, av, sb + (uint32_t) (g * p.act_smem_stride), warp, lane);
        else quantize_act32<T == T_MXFP4>(p, av, sb + (uint32_t) (g * p.act_smem_stride), warp, lane);
    }
    __syncthreads();

    const int lpr  = p.lpr;
    const int sub  = lane & (lpr - 1);
    const int lrow = lane / lpr;
    for (int i = 0; i < nmine; ++i) {
        const int       t = T0 + warp + i * NA, slot = slot0 + i % D, use = i / D;
        const StageInfo si = stage_info<NSEG>(p, t);
        mbar_wait_dbg(sb + OFF_FULL + 8 * slot, use & 1, 2, i);
        const uint32_t lead = (uint32_t) ((uintptr_t) (seg_w(si.g) + (size_t) si.r0 * p.row_bytes) & 15);
        const uint32_t sba  = MOE ? sb + (uint32_t) (si.g * p.act_smem_stride) : sb;  // this segment's quantized activations
        const uint32_t row  = sb + p.off_stage + (uint32_t) slot * p.stage_bytes + lead + (uint32_t) lrow * (uint32_t) p.row_bytes;
        float acc = 0.0f;
        if (lrow < si.nrows) {
            for (int s = sub; s < p.nsb; s += lpr) {
                if constexpr (IsKQuant<T>::value) acc += SbDot<T>::run(p, sba, row, s);
                else acc += B32Dot<T>::run(p, sba, row, s);
            }
        }
        __syncwarp();  // every lane is done reading the slot
        if (lane == 0 && i + D < nmine) {
            fence_proxy_async_smem();  // order the generic-proxy reads before the async-proxy refill
            issue(i + D);
        }
        for (int o = lpr >> 1; o > 0; o >>= 1) {
            acc += __shfl_xor_sync(0xffffffffu, acc, o);
        }
        if (sub == 0 && lrow < si.nrows) {
            float * d = p.seg[si.g].dst + si.r0 + lrow;
            if (MOE) {
                if (p.bias) {  // the expert's bias row (ggml_add_id behind ggml_mul_mat_id): one f32 add, as the separate operator
                    long long eo = eoff[0];
#pragma unroll
                    for (int q = 1; q < KQ_MAX_SEG; ++q) {
                        if (si.g == q) eo = eoff[q];
                    }
                    acc = acc + __ldg(p.bias + eo + si.r0 + lrow);
                }
            }
            *d = acc;
            for (int r = 0; r < p.n_peers; ++r) {  // fire-and-forget stores over NVLink (visible at kernel completion)
                *(float *) ((char *) d + p.peer_delta[r]) = acc;
            }
        }
    }
    KQ_HANG_COUNT(2042);  // warp 0 of a CTA done
}

}  // namespace

template <bool MOE> static kq_kernel_t kq_kernel_sel(int type) {
    switch (type) {
        case T_Q4_K: return gemv_kq_kernel<T_Q4_K, MOE>;
        case T_Q5_K: return gemv_kq_kernel<T_Q5_K, MOE>;
        case T_Q6_K: return gemv_kq_kernel<T_Q6_K, MOE>;
        case T_Q4_0: return gemv_kq_kernel<T_Q4_0, MOE>;
        case T_Q4_1: return gemv_kq_kernel<T_Q4_1, MOE>;
        case T_Q5_0: return gemv_kq_kernel<T_Q5_0, MOE>;
        case T_Q5_1: return gemv_kq_kernel<T_Q5_1, MOE>;
        case T_Q8_0: return gemv_kq_kernel<T_Q8_0, MOE>;
        case T_MXFP4: return gemv_kq_kernel<T_MXFP4, MOE>;
        default: return nullptr;
    }
}
kq_kernel_t kq_kernel_for(int type, bool moe) { return moe ? kq_kernel_sel<true>(type) : kq_kernel_sel<false>(type); }
int         kq_max_segments(bool moe) { return moe ? KQ_MAX_SEG : KQ_PLAIN_SEG; }
void kq_set_hang_buffer(unsigned long long * host_mapped) {
#ifdef B200Q_DEBUG_HANG
    cudaMemcpyToSymbol(g_kq_hang, &host_mapped, sizeof(host_mapped));
#else
    (void) host_mapped;  // release build: the kernels carry no hang reporting
#endif
}
int kq_threads() { return NTHREADS; }
int kq_consumer_warps() { return kqdot::NW; }

}  // namespace b200q
