// gemv_mma.cu — decode GEMV for 2..8 activation columns (q4_K q5_K q6_K, q4_0 q4_1 q5_0 q5_1 q8_0): the weights stream
// through shared memory once, a warp contracts a 16-row x 8-column tile per super-block with mma.sync (u8 x s8 -> s32),
// so 8 columns cost the instructions of one (gemv_mma.cuh has the arithmetic and its CPU selftest).
// Replaces the reference's quantize_q8_1 + mul_mat_vec_q<ncols_dst = 2..8> pair (ggml-cuda/quantize.cu:4-48,
// mmvq.cu:139-226).
//
//   * 16 warps (8 when the ring cannot give each of 16 a slot), one CTA per SM with the whole shared memory: quantized
//     activation images of all columns (int8, permuted into B-fragment order) + a ring of stages.  A stage = 16 rows x kc
//     super-blocks (or blocks); every row of a stage is its own bulk copy (TMA 1-D) into a padded pitch (== 32 or 96 mod
//     128 for LDS.64 fragments, an odd multiple of 16 for LDS.32 ones), which makes the per-lane A-fragment loads (rows 8
//     apart) bank-conflict-free without touching the GGUF layout in HBM.
//   * every warp feeds itself (slot ownership, as in gemv_kq.cu): one elected lane issues the 16 row copies of the warp's
//     next stage, from warp-uniform code, into the slot the warp has just finished.
//   * CTA pairs (2-CTA clusters) split the activation quantization and store their halves into both shared memories.
//   * a 16-row group is shared by kp warps along k ("team", kp = 1, 2, 4, 8 chosen by the planner so that short
//     matrices still occupy every warp); partial tiles meet in shared memory behind a named barrier and are summed in
//     fixed order: results are bit-reproducible and independent of the grid.
#include "gemv_mma.cuh"

#include "sm100.cuh"

#include <cuda_runtime.h>

namespace b200q {

namespace {

using namespace sm100;
using namespace mma;

constexpr int MAX_WARPS   = 16;
constexpr int MAX_THREADS = MAX_WARPS * 32;

extern __shared__ __align__(128) uint8_t mma_smem[];

struct DevExec {
    __host__ __device__ __forceinline__ void mma32(int, int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#ifdef __CUDA_ARCH__
        asm volatile(
            "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
            : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(0));
#endif
    }
    __host__ __device__ __forceinline__ void mma32s(int, int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#ifdef __CUDA_ARCH__
        asm volatile(
            "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
            : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
            : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(0));
#endif
    }
    __host__ __device__ __forceinline__ void mma16(int, int (&c)[4], const uint32_t (&a)[2], uint32_t b0) {
#ifdef __CUDA_ARCH__
        asm volatile(
            "mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%7, %7, %7, %7};"
            : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
            : "r"(a[0]), "r"(a[1]), "r"(b0), "r"(0));
#endif
    }
};

// ---- 2-CTA clusters share the activation quantization: each CTA quantizes half of the (column, unit) pairs and stores
// the result into its own shared memory AND its peer's (distributed shared memory), halving the prologue every CTA pays
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
struct Peer {  // peer's image of this CTA's shared window: remote address = rb + (local address - sb)
    uint32_t sb, rb;
    bool     on;
};
__device__ __forceinline__ void st_u32(const Peer & pr, uint32_t a, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
    if (pr.on) asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(pr.rb + (a - pr.sb)), "r"(v) : "memory");
}
__device__ __forceinline__ void st_u16(const Peer & pr, uint32_t a, short v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(v) : "memory");
    if (pr.on) asm volatile("st.shared::cluster.u16 [%0], %1;" ::"r"(pr.rb + (a - pr.sb)), "h"(v) : "memory");
}
__device__ __forceinline__ void st_f32(const Peer & pr, uint32_t a, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
    if (pr.on) asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(pr.rb + (a - pr.sb)), "f"(v) : "memory");
}
__device__ __forceinline__ void st_v2u32(const Peer & pr, uint32_t a, uint32_t v0, uint32_t v1) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(v0), "r"(v1) : "memory");
    if (pr.on) asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(pr.rb + (a - pr.sb)), "r"(v0), "r"(v1) : "memory");
}
__device__ __forceinline__ void st_v2f32(const Peer & pr, uint32_t a, float v0, float v1) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(v0), "f"(v1) : "memory");
    if (pr.on) asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(pr.rb + (a - pr.sb)), "f"(v0), "f"(v1) : "memory");
}

// ---- activation quantization: f32 column -> int8 image in B-fragment order, scale, int16 partial sums -------------
// the arithmetic of gemv_kq.cu's quantize_act (per 256 values: d = amax / 127, q = round(x * 127 / amax)); one warp
// per super-block, lane l holds values 8l .. 8l+7
template <int T> __device__ __forceinline__ void quantize_cols(const MmaParams & p, uint32_t sb, int warp, int lane, int nw, const Peer & pr) {
    // warp = this warp's index among the nw warps of the cluster (or CTA)
    constexpr int BSB = T == T_Q6_K ? 32 : 16;
    constexpr int QB  = 8;  // units in flight per warp: the global loads of a batch are issued together
    const int     total = p.n * p.nsb;
    for (int u0 = warp; u0 < total; u0 += nw * QB) {
        float4 x[QB][2];
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            const int u = u0 + b * nw;
            if (u < total) {
                const int      col = u / p.nsb, s = u - col * p.nsb;
                const float4 * src = (const float4 *) ((const char *) p.act + (long long) col * p.a_nb1) + (size_t) s * 64 + lane * 2;
                x[b][0] = __ldg(src);
                x[b][1] = __ldg(src + 1);
            }
        }
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            const int u = u0 + b * nw;
            if (u >= total) break;  // warp-uniform
            const int   col = u / p.nsb, s = u - col * p.nsb;
            const float v[8] = { x[b][0].x, x[b][0].y, x[b][0].z, x[b][0].w, x[b][1].x, x[b][1].y, x[b][1].z, x[b][1].w };
            float amax = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
            const float d  = amax * (1.0f / 127.0f);
            const float id = amax > 0.0f ? 127.0f / amax : 0.0f;
            int q[8], sum = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                q[i] = __float2int_rn(v[i] * id);
                sum += q[i];
            }
            const uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t) q[3] << 24);
            const uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t) q[7] << 24);
            const uint32_t qa = sb + p.off_q + (uint32_t) col * (uint32_t) p.col_pitch + (uint32_t) s * 256u;
            st_u32(pr, qa + (uint32_t) act_word_off<T>(lane * 8), w0);
            st_u32(pr, qa + (uint32_t) act_word_off<T>(lane * 8 + 4), w1);
            const int      s16 = sum + __shfl_xor_sync(0xffffffffu, sum, 1);
            const int      s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 2);
            const uint32_t bsa = sb + p.off_bs + (uint32_t) (col * p.bs_pitch + s * BSB);
            if (T == T_Q6_K) {
                if ((lane & 1) == 0) st_u16(pr, bsa + (uint32_t) (lane >> 1) * 2, (short) s16);
            } else {
                if ((lane & 3) == 0) st_u16(pr, bsa + (uint32_t) (lane >> 2) * 2, (short) s32);
            }
            if (lane == 0) st_f32(pr, sb + p.off_dx + (uint32_t) (col * p.dx_pitch + s * 4), d);
        }
    }
}

// 32-element block formats: per-32 scales, natural element order, (d, d * sum q) per block — quantize_row_q8_1's numbers
// (gemv_kq.cu quantize_act32).  One warp per 256 values per step; lane l holds values 8l .. 8l+7, 4 lanes per block.
__device__ __forceinline__ void quantize_cols32(const MmaParams & p, uint32_t sb, int warp, int lane, int nw, const Peer & pr) {
    constexpr int QB = 8;
    const int nb = p.nsb, ng = (nb + 7) / 8, total = p.n * ng;
    for (int u0 = warp; u0 < total; u0 += nw * QB) {
        float4 x[QB][2];
#pragma unroll
        for (int bi = 0; bi < QB; ++bi) {
            const int u = u0 + bi * nw;
            x[bi][0] = x[bi][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (u < total) {
                const int col = u / ng, g = u - col * ng;
                if (g * 8 + (lane >> 2) < nb) {
                    const float4 * src = (const float4 *) ((const char *) p.act + (long long) col * p.a_nb1) + (size_t) g * 64 + lane * 2;
                    x[bi][0] = __ldg(src);
                    x[bi][1] = __ldg(src + 1);
                }
            }
        }
#pragma unroll
        for (int bi = 0; bi < QB; ++bi) {
            const int u = u0 + bi * nw;
            if (u >= total) break;  // warp-uniform
            const int   col = u / ng, g = u - col * ng;
            const int   b     = g * 8 + (lane >> 2);
            const bool  valid = b < nb;  // uniform over the 4 lanes of a block
            const float v[8] = { x[bi][0].x, x[bi][0].y, x[bi][0].z, x[bi][0].w, x[bi][1].x, x[bi][1].y, x[bi][1].z, x[bi][1].w };
            float amax = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d  = amax * (1.0f / 127.0f);
            const float id = amax > 0.0f ? 127.0f / amax : 0.0f;
            int q[8], sum = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                q[i] = __float2int_rn(v[i] * id);
                sum += q[i];
            }
            sum += __shfl_xor_sync(0xffffffffu, sum, 1);
            sum += __shfl_xor_sync(0xffffffffu, sum, 2);
            if (valid) {
                const uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t) q[3] << 24);
                const uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t) q[7] << 24);
                const uint32_t qa = sb + p.off_q + (uint32_t) col * (uint32_t) p.col_pitch + (uint32_t) (g * 256 + lane * 8);
                st_v2u32(pr, qa, w0, w1);
                if ((lane & 3) == 0) st_v2f32(pr, sb + p.off_dx + (uint32_t) (col * p.dx_pitch + b * 8), d, d * (float) sum);
            }
        }
    }
}

template <int T> __global__ void __launch_bounds__(MAX_THREADS, 1) gemv_mma_kernel(const MmaParams p) {
    constexpr int BB   = BlockInfo<T>::BYTES;
    const uint32_t sb   = smem_u32(mma_smem);
    const int      NW   = (int) (blockDim.x >> 5);                                // 8 or 16 warps (planner)
    const int      warp = __shfl_sync(0xffffffffu, (int) (threadIdx.x >> 5), 0);  // warp-uniform for the compiler
    const int      lane = threadIdx.x & 31;
    const int      S    = p.nstages;
    // ring slots of this warp (the S % NW spare slots go to the first warps)
    const int D     = S / NW + (warp < S % NW ? 1 : 0);
    const int slot0 = warp * (S / NW) + min(warp, S % NW);
    // this CTA's contiguous range of 16-row groups; team t of the CTA takes groups G0 + t, G0 + t + teams, ...
    const int G0 = (int) ((long long) p.groups * blockIdx.x / gridDim.x);
    const int G1 = (int) ((long long) p.groups * (blockIdx.x + 1) / gridDim.x);
    const int kp = p.kp, teams = NW / kp, team = warp / kp, member = warp - team * kp;
    const int ngrp  = G1 - G0 > team ? (G1 - G0 - team + teams - 1) / teams : 0;
    const int cpm   = p.cpm;
    const int nmine = ngrp * cpm;  // stages of this warp

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; ++s) mbar_init(sb + 8 * s, 1);
        mbar_fence_init();
    }
    const bool clustered = p.cluster > 1;
    Peer       pr{ sb, 0u, clustered };
    if (clustered) {
        cluster_sync_all();  // the peer CTA is running: its shared memory may be written from here on
        pr.rb = mapa_u32(sb, cluster_ctarank() ^ 1u);
    } else {
        __syncthreads();
    }
    pdl_launch_dependents();

    const uint64_t pol   = policy_evict_first();
    // whole warp, warp-uniform operands; one elected lane issues (code under `if (lane == 0)` would make the compiler
    // wrap every bulk copy in an ELECT / R2UR.BROADCAST loop).  Stage i of this warp = (group G0 + team + it * teams,
    // chunk member * cpm + ci) with i = it * cpm + ci, in ring slot slot0 + i % D: running cursors, no divisions.
    int  p_g = G0 + team, p_ci = 0, p_slot = 0, issued = 0;  // next stage to request
    auto issue_next = [&]() {
        const int       chunk = member * cpm + p_ci;
        const int       nrows = min(16, p.m - 16 * p_g);
        const long long coff  = (long long) chunk * p.chunk_bytes;
        const uint32_t  lead  = (uint32_t) (coff & 15);
        const uint32_t  win   = (lead + (uint32_t) p.chunk_bytes + 15u) & ~15u;
        const uint32_t  full  = sb + 8 * (slot0 + p_slot);
        const uint8_t * src   = p.w + (long long) (16 * p_g) * p.row_bytes + coff - lead;
        uint32_t        dst   = sb + p.off_stage + (uint32_t) (slot0 + p_slot) * p.stage_bytes;
        if (elect_one()) mbar_expect_tx(full, win * (uint32_t) nrows);
        __syncwarp();
        for (int r = 0; r < nrows; ++r) {
            if (elect_one()) bulk_g2s_hint(dst, src, win, full, pol);
            src += p.row_bytes;
            dst += p.pitch;
        }
        if (++p_ci == cpm) { p_ci = 0; p_g += teams; }
        if (++p_slot == D) p_slot = 0;
        ++issued;
    };
    // immutable weights may stream before the predecessor has finished: this kernel owns the whole shared memory, so a CTA
    // only starts once its SM is free, but SMs whose predecessor CTA had less work are free early
    if (p.early_prefetch) {
        while (issued < min(D, nmine)) issue_next();
    }
    pdl_wait();  // the activations (and dst) belong to the previous kernels of the stream
    while (issued < min(D, nmine)) issue_next();
    {
        const int qw = clustered ? (int) cluster_ctarank() * NW + warp : warp, qn = clustered ? 2 * NW : NW;
        if (!MmaIsKQuant<T>::value) quantize_cols32(p, sb, qw, lane, qn, pr);
        else if (T == T_Q6_K) quantize_cols<T_Q6_K>(p, sb, qw, lane, qn, pr);
        else quantize_cols<T_Q4_K>(p, sb, qw, lane, qn, pr);  // q4_K and q5_K share the layout
    }
    if (clustered) cluster_sync_all();  // both halves of the images have landed in both CTAs
    else __syncthreads();

    const TileCtx t{ sb + (uint32_t) p.off_q, sb + (uint32_t) p.off_dx, sb + (uint32_t) p.off_bs, p.col_pitch, p.dx_pitch, p.bs_pitch, p.n };
    DevExec       ex;
    const int     r = lane >> 2, c = lane & 3;
    float         facc[4] = { 0.f, 0.f, 0.f, 0.f };
    int           g = G0 + team, ci = 0, slot = 0, it = 0;  // the stage being consumed
    uint32_t      parity = 0;
    for (int i = 0; i < nmine; ++i) {
        const int chunk = member * cpm + ci;
        mbar_wait(sb + 8 * (slot0 + slot), parity);
        const uint32_t lead = (uint32_t) (((long long) chunk * p.chunk_bytes) & 15);
        const uint32_t blk  = sb + p.off_stage + (uint32_t) (slot0 + slot) * p.stage_bytes + lead + (uint32_t) r * p.pitch;
        if constexpr (MmaIsKQuant<T>::value) {
            for (int sbi = 0; sbi < p.kc; ++sbi) {
                TileSb<T, DevExec>::run(ex, t, lane, chunk * p.kc + sbi, blk + (uint32_t) sbi * BB, blk + (uint32_t) sbi * BB + 8u * p.pitch, facc);
            }
        } else {
#pragma unroll 4
            for (int sbi = 0; sbi < p.kc; ++sbi) {
                TileSb<T, DevExec>::run(ex, t, lane, chunk * p.kc + sbi, blk + (uint32_t) sbi * BB, blk + (uint32_t) sbi * BB + 8u * p.pitch, facc);
            }
        }
        __syncwarp();              // every lane is done reading the slot
        fence_proxy_async_smem();  // order the generic-proxy reads before the async-proxy refill
        if (issued < nmine) issue_next();  // stage i + D goes into the slot just finished
        if (++slot == D) { slot = 0; parity ^= 1; }
        if (++ci == cpm) {  // this warp's share of group g is complete
            ci = 0;
            if (kp > 1) {
                const uint32_t part = sb + p.off_part + (uint32_t) (((team * 2 + (it & 1)) * kp) * 32 + lane) * 16u;
                if (member != 0) {
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(part + (uint32_t) member * 512u), "f"(facc[0]), "f"(facc[1]),
                                 "f"(facc[2]), "f"(facc[3]) : "memory");
                }
                asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(kp * 32) : "memory");
                if (member == 0) {
                    for (int mb = 1; mb < kp; ++mb) {
                        float4 o;
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w) : "r"(part + (uint32_t) mb * 512u) : "memory");
                        facc[0] += o.x; facc[1] += o.y; facc[2] += o.z; facc[3] += o.w;
                    }
                }
            }
            if (member == 0) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = 16 * g + r + 8 * (e >> 1), col = 2 * c + (e & 1);
                    if (row < p.m && col < p.n) *(float *) ((char *) p.dst + (long long) col * p.d_nb1 + (long long) row * 4) = facc[e];
                }
            }
            facc[0] = facc[1] = facc[2] = facc[3] = 0.f;
            g += teams;
            ++it;
        }
    }
}

}  // namespace

mma_kernel_t mma_kernel_for(int type) {
    switch (type) {
        case T_Q4_K: return gemv_mma_kernel<T_Q4_K>;
        case T_Q5_K: return gemv_mma_kernel<T_Q5_K>;
        case T_Q6_K: return gemv_mma_kernel<T_Q6_K>;
        case T_Q4_0: return gemv_mma_kernel<T_Q4_0>;
        case T_Q8_0: return gemv_mma_kernel<T_Q8_0>;
        case T_Q4_1: return gemv_mma_kernel<T_Q4_1>;
        case T_Q5_0: return gemv_mma_kernel<T_Q5_0>;
        case T_Q5_1: return gemv_mma_kernel<T_Q5_1>;
        default: return nullptr;
    }
}
int mma_max_warps() { return MAX_WARPS; }
int mma_bs_bytes(int type) { return type == T_Q6_K ? 32 : type == T_Q4_K || type == T_Q5_K ? 16 : 0; }

}  // namespace b200q
