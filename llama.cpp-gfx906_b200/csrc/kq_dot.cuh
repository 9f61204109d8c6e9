// kq_dot.cuh — the device arithmetic shared by the decode GEMV kernels (gemv_kq.cu: one launch per matmul group;
// gemv_list.cu: one persistent launch per matmul list): activation quantization into shared memory and the per-unit
// dot products on raw GGUF block bytes.  P is any parameter struct with the shared-memory layout fields
// off_q, off_dx, off_bs16, off_bs32 and nsb (units per row).
#pragma once
#include "blocks.cuh"
#include "sm100.cuh"

#include <cuda_runtime.h>
#include <stdint.h>

namespace b200q {
namespace kqdot {

using namespace sm100;

constexpr int NW = 8;  // warps that share one quantization pass / own ring slots

__device__ __forceinline__ int4 lds128(uint32_t a) {
    int4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ float lds32f(uint32_t a) { return __uint_as_float(lds32(a)); }
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {  // a: 4 unsigned bytes, b: 4 signed bytes
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_lo_su(int a, uint32_t b, int c) {  // a: 2 x s16, b: bytes 0,1 unsigned
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(int a, uint32_t b, int c) {  // bytes 2,3 of b
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_lo_ss(int a, int b, int c) {
    int d;
    asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_ss(int a, int b, int c) {
    int d;
    asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// ---- activation quantization: f32 [k] -> per super-block int8 (chunk-major), scale, int16 sums -----------------
//   q   : 16-byte chunk c (0..15) of super-block s at off_q + (c * nsb + s) * 16
//   dx  : float at off_dx + 4 s                      (x ~= dx * q)
//   bs16: 16 x int16 at off_bs16 + 32 s              (sums of q over 16 values)
//   bs32:  8 x int16 at off_bs32 + 16 s              (sums over 32 values)
// One warp per super-block per step; lane l holds values 8l .. 8l+7 (coalesced 1 KB per warp).
// where the eight values of a lane come from: the vector itself, or silu(a) * b of two vectors (the SWIGLU in front of ffn_down
// folded into the quantization, same expression as ops.cu's glu_kernel)
struct ActPlain {
    const float * a;
    __device__ __forceinline__ void operator()(size_t i, float4 & x0, float4 & x1) const {
        const float4 * src = (const float4 *) (a + i);
        x0 = __ldcg(src);
        x1 = __ldcg(src + 1);
    }
};
struct ActSwiglu {
    const float * a, * b;
    // every CTA evaluates the whole vector with its four quantizer warps, nothing to overlap with behind the barrier point: the
    // IEEE division and expf cost 9.5 us per 14336 values, the approximate forms (2^-21 relative, far below the int8 step) 2 us
    static __device__ __forceinline__ float f(float x, float g) { return __fdividef(x, 1.0f + __expf(-x)) * g; }
    __device__ __forceinline__ void operator()(size_t i, float4 & x0, float4 & x1) const {
        const float4 * sa = (const float4 *) (a + i), * sg = (const float4 *) (b + i);
        const float4   a0 = __ldcg(sa), a1 = __ldcg(sa + 1), g0 = __ldcg(sg), g1 = __ldcg(sg + 1);
        x0 = make_float4(f(a0.x, g0.x), f(a0.y, g0.y), f(a0.z, g0.z), f(a0.w, g0.w));
        x1 = make_float4(f(a1.x, g1.x), f(a1.y, g1.y), f(a1.z, g1.z), f(a1.w, g1.w));
    }
};
template <int NWQ = NW, class P, class L> __device__ __forceinline__ void quantize_act_from(const P & p, const L & load, uint32_t sb, int warp, int lane);
template <int NWQ = NW, class P> __device__ __forceinline__ void quantize_act(const P & p, const float * act, uint32_t sb, int warp, int lane) {
    quantize_act_from<NWQ>(p, ActPlain{ act }, sb, warp, lane);
}
template <int NWQ, class P, class L> __device__ __forceinline__ void quantize_act_from(const P & p, const L & load, uint32_t sb, int warp, int lane) {
    constexpr int QB = 4;
    for (int s0 = warp; s0 < p.nsb; s0 += NWQ * QB) {
        float4 x[QB][2];
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            const int s = s0 + b * NWQ;
            if (s < p.nsb) {
                load((size_t) s * 256 + lane * 8, x[b][0], x[b][1]);
            } else {
                x[b][0] = x[b][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        // the QB reductions advance in lockstep (the loop over shuffle distances is the outer one): a warp that quantizes
        // alone on its scheduler has no other warp to hide the shuffle latency behind
        float amax[QB];
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            amax[b] = fmaxf(fmaxf(fmaxf(fabsf(x[b][0].x), fabsf(x[b][0].y)), fmaxf(fabsf(x[b][0].z), fabsf(x[b][0].w))),
                            fmaxf(fmaxf(fabsf(x[b][1].x), fabsf(x[b][1].y)), fmaxf(fabsf(x[b][1].z), fabsf(x[b][1].w))));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int b = 0; b < QB; ++b) amax[b] = fmaxf(amax[b], __shfl_xor_sync(0xffffffffu, amax[b], o));
        }
        uint32_t w0[QB], w1[QB];
        int      sum[QB];
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            const float v[8] = { x[b][0].x, x[b][0].y, x[b][0].z, x[b][0].w, x[b][1].x, x[b][1].y, x[b][1].z, x[b][1].w };
            const float id   = amax[b] > 0.0f ? 127.0f / amax[b] : 0.0f;
            int q[8];
            sum[b] = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                q[i] = __float2int_rn(v[i] * id);
                sum[b] += q[i];
            }
            w0[b] = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t) q[3] << 24);
            w1[b] = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t) q[7] << 24);
        }
        int s16[QB], s32[QB];
#pragma unroll
        for (int b = 0; b < QB; ++b) s16[b] = sum[b] + __shfl_xor_sync(0xffffffffu, sum[b], 1);
#pragma unroll
        for (int b = 0; b < QB; ++b) s32[b] = s16[b] + __shfl_xor_sync(0xffffffffu, s16[b], 2);
#pragma unroll
        for (int b = 0; b < QB; ++b) {
            const int s = s0 + b * NWQ;
            if (s >= p.nsb) break;  // warp-uniform
            const uint32_t qa = sb + p.off_q + (uint32_t) (((lane >> 1) * p.nsb + s) * 16 + (lane & 1) * 8);
            asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(qa), "r"(w0[b]), "r"(w1[b]) : "memory");
            if ((lane & 1) == 0) {
                asm volatile("st.shared.u16 [%0], %1;" ::"r"(sb + p.off_bs16 + (uint32_t) (s * 32 + (lane >> 1) * 2)), "h"((short) s16[b]) : "memory");
            }
            if ((lane & 3) == 0) {
                asm volatile("st.shared.u16 [%0], %1;" ::"r"(sb + p.off_bs32 + (uint32_t) (s * 16 + (lane >> 2) * 2)), "h"((short) s32[b]) : "memory");
            }
            if (lane == 0) {
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(sb + p.off_dx + (uint32_t) s * 4), "f"(amax[b] * (1.0f / 127.0f)) : "memory");
            }
        }
    }
}

// ---- where a super-block's quantized activations come from -----------------------------------------------------------
// shared memory (the image quantize_act wrote) ...
template <class P> struct SmemAct {
    const P & p;
    uint32_t  sb;
    int       s;
    __device__ __forceinline__ int4 chunk(int c) const { return lds128(sb + p.off_q + (uint32_t) s * 16u + (uint32_t) c * ((uint32_t) p.nsb * 16u)); }
    __device__ __forceinline__ int4 bs32() const { return lds128(sb + p.off_bs32 + (uint32_t) s * 16u); }
    __device__ __forceinline__ int4 bs16(int i) const { return lds128(sb + p.off_bs16 + (uint32_t) s * 32u + 16u * i); }
    __device__ __forceinline__ float dx() const { return lds32f(sb + p.off_dx + (uint32_t) s * 4u); }
};
// ... or registers: a lane that keeps ONE super-block column for many rows loads it once (gemv_list.cu: the activation
// re-reads were 60 % of the shared-memory wavefronts of the decode kernels, profiles/r02_list_timeline_v1.txt)
struct RegAct {
    int4  a[16];
    int4  s16[2];  // 16 x int16 sums over 16 values (q6_K)
    int4  s32;     // 8 x int16 sums over 32 values (q4_K)
    float d;
    __device__ __forceinline__ int4 chunk(int c) const { return a[c]; }
    __device__ __forceinline__ int4 bs32() const { return s32; }
    __device__ __forceinline__ int4 bs16(int i) const { return s16[i]; }
    __device__ __forceinline__ float dx() const { return d; }
    template <class P> __device__ __forceinline__ void load(const P & p, uint32_t sb, int s) {
        const SmemAct<P> m{ p, sb, s };
#pragma unroll
        for (int c = 0; c < 16; ++c) a[c] = m.chunk(c);
        s16[0] = m.bs16(0);
        s16[1] = m.bs16(1);
        s32    = m.bs32();
        d      = m.dx();
    }
};

// ---- super-block dots ---------------------------------------------------------------------------------------------
// returns  sum_{i in super-block s} w_i * x_i   for the row that starts at shared address `row`
template <int T> struct SbDot;

template <> struct SbDot<T_Q4_K> {
    static constexpr int SB_BYTES = 144;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int s) {
        return core(row + (uint32_t) s * 144u, SmemAct<P>{ p, sb, s });
    }
    // b: shared address of the super-block; A: where the quantized activations of this super-block come from
    template <class A> static __device__ __forceinline__ float core(uint32_t b, const A & act) {
        const int4     h = lds128(b);  // [d dmin] [scales x3]
        // get_scale_min_k4 (ggml-quants.c:703-710) on whole words: bytes of sc_lo/mn_lo = sub-blocks 0..3, *_hi = 4..7
        const uint32_t sc_lo = (uint32_t) h.y & 0x3F3F3F3Fu, mn_lo = (uint32_t) h.z & 0x3F3F3F3Fu;
        const uint32_t sc_hi = ((uint32_t) h.w & 0x0F0F0F0Fu) | (((uint32_t) h.y >> 2) & 0x30303030u);
        const uint32_t mn_hi = (((uint32_t) h.w >> 4) & 0x0F0F0F0Fu) | (((uint32_t) h.z >> 2) & 0x30303030u);
        int acc = 0;
#pragma unroll
        for (int pq = 0; pq < 4; ++pq) {
            const int4 q0 = lds128(b + 16 + 32 * pq), q1 = lds128(b + 32 + 32 * pq);
            const int4 a0 = act.chunk(4 * pq + 0), a1 = act.chunk(4 * pq + 1);
            const int4 a2 = act.chunk(4 * pq + 2), a3 = act.chunk(4 * pq + 3);
            int lo = 0, lo2 = 0, hi = 0, hi2 = 0;  // hi sums carry a factor 16 (high nibbles are not shifted down)
            lo  = dp4a_us((uint32_t) q0.x & 0x0F0F0F0Fu, a0.x, lo);
            lo2 = dp4a_us((uint32_t) q0.y & 0x0F0F0F0Fu, a0.y, lo2);
            lo  = dp4a_us((uint32_t) q0.z & 0x0F0F0F0Fu, a0.z, lo);
            lo2 = dp4a_us((uint32_t) q0.w & 0x0F0F0F0Fu, a0.w, lo2);
            lo  = dp4a_us((uint32_t) q1.x & 0x0F0F0F0Fu, a1.x, lo);
            lo2 = dp4a_us((uint32_t) q1.y & 0x0F0F0F0Fu, a1.y, lo2);
            lo  = dp4a_us((uint32_t) q1.z & 0x0F0F0F0Fu, a1.z, lo);
            lo2 = dp4a_us((uint32_t) q1.w & 0x0F0F0F0Fu, a1.w, lo2);
            hi  = dp4a_us((uint32_t) q0.x & 0xF0F0F0F0u, a2.x, hi);
            hi2 = dp4a_us((uint32_t) q0.y & 0xF0F0F0F0u, a2.y, hi2);
            hi  = dp4a_us((uint32_t) q0.z & 0xF0F0F0F0u, a2.z, hi);
            hi2 = dp4a_us((uint32_t) q0.w & 0xF0F0F0F0u, a2.w, hi2);
            hi  = dp4a_us((uint32_t) q1.x & 0xF0F0F0F0u, a3.x, hi);
            hi2 = dp4a_us((uint32_t) q1.y & 0xF0F0F0F0u, a3.y, hi2);
            hi  = dp4a_us((uint32_t) q1.z & 0xF0F0F0F0u, a3.z, hi);
            hi2 = dp4a_us((uint32_t) q1.w & 0xF0F0F0F0u, a3.w, hi2);
            const uint32_t scw = pq < 2 ? sc_lo : sc_hi;
            const int      sh  = 16 * (pq & 1);
            acc += (int) ((scw >> sh) & 0xFFu) * (lo + lo2) + (int) ((scw >> (sh + 8)) & 0xFFu) * ((hi + hi2) >> 4);
        }
        const int4 bs = act.bs32();  // 8 x int16 sums over 32 activations
        int mins = dp2a_lo_su(bs.x, mn_lo, 0);
        mins     = dp2a_hi_su(bs.y, mn_lo, mins);
        mins     = dp2a_lo_su(bs.z, mn_hi, mins);
        mins     = dp2a_hi_su(bs.w, mn_hi, mins);
        const float d    = half_bits_to_float((uint16_t) ((uint32_t) h.x & 0xffffu));
        const float dmin = half_bits_to_float((uint16_t) ((uint32_t) h.x >> 16));
        const float dx   = act.dx();
        return dx * (d * (float) acc - dmin * (float) mins);
    }
};

// q5_K (ggml-common.h block_q5_K, dequantize_row_q5_K ggml-quants.c:1477-1513): q4_K's layout with 32 bytes of fifth bits behind
// the header — bit 2j of qh[l] lifts element l of sub-block 2j by 16, bit 2j + 1 that of sub-block 2j + 1.  Both nibbles are
// brought down to bytes 0..31 so that the fifth bit fits (q4_K leaves the high nibbles in place and shifts the sum).
template <> struct SbDot<T_Q5_K> {
    static constexpr int SB_BYTES = 176;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int s) {
        return core(row + (uint32_t) s * 176u, SmemAct<P>{ p, sb, s });
    }
    template <class A> static __device__ __forceinline__ float core(uint32_t b, const A & act) {
        const int4     h = lds128(b);  // [d dmin] [scales x3]
        const uint32_t sc_lo = (uint32_t) h.y & 0x3F3F3F3Fu, mn_lo = (uint32_t) h.z & 0x3F3F3F3Fu;
        const uint32_t sc_hi = ((uint32_t) h.w & 0x0F0F0F0Fu) | (((uint32_t) h.y >> 2) & 0x30303030u);
        const uint32_t mn_hi = (((uint32_t) h.w >> 4) & 0x0F0F0F0Fu) | (((uint32_t) h.z >> 2) & 0x30303030u);
        const int4     f0 = lds128(b + 16), f1 = lds128(b + 32);  // qh[0..15], qh[16..31]
        const uint32_t fh[8] = { (uint32_t) f0.x, (uint32_t) f0.y, (uint32_t) f0.z, (uint32_t) f0.w,
                                 (uint32_t) f1.x, (uint32_t) f1.y, (uint32_t) f1.z, (uint32_t) f1.w };
        int acc = 0;
#pragma unroll
        for (int pq = 0; pq < 4; ++pq) {
            const int4 q0 = lds128(b + 48 + 32 * pq), q1 = lds128(b + 64 + 32 * pq);
            const uint32_t qw[8] = { (uint32_t) q0.x, (uint32_t) q0.y, (uint32_t) q0.z, (uint32_t) q0.w,
                                     (uint32_t) q1.x, (uint32_t) q1.y, (uint32_t) q1.z, (uint32_t) q1.w };
            const int4 a0 = act.chunk(4 * pq + 0), a1 = act.chunk(4 * pq + 1);
            const int4 a2 = act.chunk(4 * pq + 2), a3 = act.chunk(4 * pq + 3);
            const int  al[8] = { a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w };
            const int  ah[8] = { a2.x, a2.y, a2.z, a2.w, a3.x, a3.y, a3.z, a3.w };
            int lo = 0, hi = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t vl = (qw[i] & 0x0F0F0F0Fu) | (((fh[i] >> (2 * pq)) & 0x01010101u) << 4);
                const uint32_t vh = ((qw[i] >> 4) & 0x0F0F0F0Fu) | (((fh[i] >> (2 * pq + 1)) & 0x01010101u) << 4);
                lo = dp4a_us(vl, al[i], lo);
                hi = dp4a_us(vh, ah[i], hi);
            }
            const uint32_t scw = pq < 2 ? sc_lo : sc_hi;
            const int      sh  = 16 * (pq & 1);
            acc += (int) ((scw >> sh) & 0xFFu) * lo + (int) ((scw >> (sh + 8)) & 0xFFu) * hi;
        }
        const int4 bs = act.bs32();  // 8 x int16 sums over 32 activations
        int mins = dp2a_lo_su(bs.x, mn_lo, 0);
        mins     = dp2a_hi_su(bs.y, mn_lo, mins);
        mins     = dp2a_lo_su(bs.z, mn_hi, mins);
        mins     = dp2a_hi_su(bs.w, mn_hi, mins);
        const float d    = half_bits_to_float((uint16_t) ((uint32_t) h.x & 0xffffu));
        const float dmin = half_bits_to_float((uint16_t) ((uint32_t) h.x >> 16));
        const float dx   = act.dx();
        return dx * (d * (float) acc - dmin * (float) mins);
    }
};

template <> struct SbDot<T_Q6_K> {
    static constexpr int SB_BYTES = 210;
    // 210-byte super-blocks are only 2-byte aligned: aligned words + funnel shift (shift is 0 or 16, constant per lane)
    template <int N> static __device__ __forceinline__ void ldw(uint32_t b4, int sh, int first, uint32_t (&out)[N]) {
        uint32_t prev = lds32(b4 + 4u * first);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const uint32_t next = lds32(b4 + 4u * (first + i + 1));
            out[i]              = __funnelshift_r(prev, next, sh);
            prev                = next;
        }
    }
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int s) {
        return core(row + (uint32_t) s * 210u, SmemAct<P>{ p, sb, s });
    }
    template <class A> static __device__ __forceinline__ float core(uint32_t b, const A & act) {
        const uint32_t b4    = b & ~3u;
        const int      sh    = (int) (b & 2u) * 8;
        uint32_t scw[5];  // [sc0..3] [sc4..7] [sc8..11] [sc12..15] [d | next]
        ldw<5>(b4, sh, 48, scw);
        int acc = 0;
#pragma unroll
        for (int n = 0; n < 2; ++n) {  // 128-element halves
            uint32_t ql[16], qh[8];
            ldw<16>(b4, sh, 16 * n, ql);      // ql[l], l = 0..63 of this half
            ldw<8>(b4, sh, 32 + 8 * n, qh);   // qh[l], l = 0..31
#pragma unroll
            for (int quad = 0; quad < 4; ++quad) {  // elements 128 n + 32 quad + l
                const int4 a0 = act.chunk(8 * n + 2 * quad);
                const int4 a1 = act.chunk(8 * n + 2 * quad + 1);
                const int  av[8] = { a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w };
                int        s0 = 0, s1 = 0;
#pragma unroll
                for (int wi = 0; wi < 8; ++wi) {
                    const uint32_t lw = ql[wi + 8 * (quad & 1)];
                    const uint32_t lo = (quad < 2 ? lw : lw >> 4) & 0x0F0F0F0Fu;
                    const uint32_t h2 = quad == 0 ? qh[wi] << 4 : quad == 1 ? qh[wi] << 2 : quad == 2 ? qh[wi] : qh[wi] >> 2;
                    const uint32_t v  = lo | (h2 & 0x30303030u);  // 6-bit value 0..63 (offset -32 applied through bs16)
                    if (wi < 4) s0 = dp4a_us(v, av[wi], s0); else s1 = dp4a_us(v, av[wi], s1);
                }
                const uint32_t w  = scw[2 * n + (quad >> 1)];
                const int      sa = (int) (int8_t) (w >> (16 * (quad & 1)));
                const int      sb_ = (int) (int8_t) (w >> (16 * (quad & 1) + 8));
                acc += sa * s0 + sb_ * s1;
            }
        }
        // - 32 * sum_g sc_g * bs16_g
        const int4 bs0 = act.bs16(0), bs1 = act.bs16(1);
        int off = dp2a_lo_ss(bs0.x, (int) scw[0], 0);
        off     = dp2a_hi_ss(bs0.y, (int) scw[0], off);
        off     = dp2a_lo_ss(bs0.z, (int) scw[1], off);
        off     = dp2a_hi_ss(bs0.w, (int) scw[1], off);
        off     = dp2a_lo_ss(bs1.x, (int) scw[2], off);
        off     = dp2a_hi_ss(bs1.y, (int) scw[2], off);
        off     = dp2a_lo_ss(bs1.z, (int) scw[3], off);
        off     = dp2a_hi_ss(bs1.w, (int) scw[3], off);
        const float d  = half_bits_to_float((uint16_t) (scw[4] & 0xffffu));
        const float dx = act.dx();
        return dx * d * (float) (acc - 32 * off);
    }
};

// ---- 32-element block formats (q4_0 q4_1 q5_0 q5_1 q8_0 mxfp4) --------------------------------------------------
// Activations are quantized per 32 values, the arithmetic of ggml-cpu's q8_0 / q8_1 partners (quantize_row_q8_1,
// ggml-cpu/quants.c; vec_dot_q4_0_q8_0 ... vec_dot_mxfp4_q8_0):
//   q  : 16-byte chunk c (0..1) of block b at off_q + (c * nb + b) * 16     (chunk-major: lanes on consecutive blocks
//        read consecutive 16-byte chunks, conflict-free)
//   ds : float2 (d, d * sum q) at off_dx + 8 b
// One warp per 256 values per step; lane l holds values 8l .. 8l+7, 4 lanes per block.
// ILV (mxfp4): the 8 words of a block are stored as (e, e+16, e+1, e+17) for e = 0, 2, .. 14 — the order in which a 16-bit
// slice of the packed nibbles (lo(qs[e]), hi(qs[e]), lo(qs[e+1]), hi(qs[e+1])) indexes a PRMT table look-up
template <bool ILV, int NWQ = NW, class P, class L> __device__ __forceinline__ void quantize_act32_from(const P & p, const L & load, uint32_t sb, int warp, int lane);
template <bool ILV, int NWQ = NW, class P> __device__ __forceinline__ void quantize_act32(const P & p, const float * act, uint32_t sb, int warp, int lane) {
    quantize_act32_from<ILV, NWQ>(p, ActPlain{ act }, sb, warp, lane);
}
template <bool ILV, int NWQ, class P, class L> __device__ __forceinline__ void quantize_act32_from(const P & p, const L & load, uint32_t sb, int warp, int lane) {
    const int nb = p.nsb;  // 32-element blocks per row
    for (int g = warp; g * 8 < nb; g += NWQ) {
        const int  b     = g * 8 + (lane >> 2);
        const bool valid = b < nb;  // uniform over the 4 lanes of a block
        float4     x0 = make_float4(0.f, 0.f, 0.f, 0.f), x1 = x0;
        if (valid) load((size_t) g * 256 + lane * 8, x0, x1);
        const float v[8] = { x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w };
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
        if (ILV) {
            const uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t) q[3] << 24);
            const uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t) q[7] << 24);
            // lanes 0,1 of a block hold elements 0..15, lanes 2,3 elements 16..31: fetch the partner half
            const uint32_t h0 = __shfl_xor_sync(0xffffffffu, w0, 2), h1 = __shfl_xor_sync(0xffffffffu, w1, 2);
            if (valid && (lane & 2) == 0) {
                const uint32_t o0 = __byte_perm(w0, h0, 0x5140), o1 = __byte_perm(w0, h0, 0x7362);
                const uint32_t o2 = __byte_perm(w1, h1, 0x5140), o3 = __byte_perm(w1, h1, 0x7362);
                const uint32_t qa = sb + p.off_q + (uint32_t) (((lane & 1) * nb + b) * 16);
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(qa), "r"(o0), "r"(o1), "r"(o2), "r"(o3) : "memory");
            }
        }
        if (valid) {
            const uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t) q[3] << 24);
            const uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t) q[7] << 24);
            if (!ILV) {
                const int      c  = (lane & 3) >> 1;
                const uint32_t qa = sb + p.off_q + (uint32_t) ((c * nb + b) * 16 + (lane & 1) * 8);
                asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(qa), "r"(w0), "r"(w1) : "memory");
            }
            if ((lane & 3) == 0) {
                asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(sb + p.off_dx + (uint32_t) b * 8), "f"(d), "f"(d * (float) sum) : "memory");
            }
        }
    }
}

// N consecutive little-endian words from ANY byte address of shared memory: N+1 aligned loads + funnel shifts
template <int N> __device__ __forceinline__ void ldsu_n(uint32_t a, uint32_t (&out)[N]) {
    const uint32_t w    = a & ~3u;
    const int      sh   = (int) (a & 3u) * 8;
    uint32_t       prev = lds32(w);
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const uint32_t next = lds32(w + 4 * (i + 1));
        out[i]              = __funnelshift_r(prev, next, sh);
        prev                = next;
    }
}
__device__ __forceinline__ int dp4a_ss(int a, int b, int c) {
    int d;
    asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// bit i of b -> bit 8 i (the 5th bits of q5_0 / q5_1)
__device__ __forceinline__ uint32_t spread4_(uint32_t b) { return ((b & 0xFu) * 0x00204081u) & 0x01010101u; }

// 8 nibbles of one word -> int8x4 words of the (doubled) E2M1 code points (ggml-common.h kvalues_mxfp4), low / high nibbles
__device__ __forceinline__ void mxfp4_words(uint32_t q, uint32_t & lo, uint32_t & hi) {
    const uint32_t t0 = 0x03020100u, t1 = 0x0C080604u, t2 = 0xFDFEFF00u, t3 = 0xF4F8FAFCu;
    uint32_t       r[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t n   = (q >> (4 * h)) & 0x0F0F0F0Fu;
        const uint32_t sel = n & 0x07070707u;
        const uint32_t s   = (sel & 0xF) | ((sel >> 4) & 0xF0) | ((sel >> 8) & 0xF00) | ((sel >> 12) & 0xF000);
        const uint32_t pos = __byte_perm(t0, t1, s), neg = __byte_perm(t2, t3, s);
        const uint32_t m   = ((n >> 3) & 0x01010101u) * 0xFFu;
        r[h]               = (pos & ~m) | (neg & m);
    }
    lo = r[0];
    hi = r[1];
}

// dot of block b of the row at shared address `row` (any byte alignment) with the quantized activations:
//   returns  sum_{i in block} w_i * x_i  =  d_w * (d_x * sum q_w q_x)  +  m_w * (d_x * sum q_x)
template <int T> struct B32Dot;
template <int T, class P> __device__ __forceinline__ void act_block(const P & p, uint32_t sb, int b, int (&a)[8], float2 & ds) {
    const int4 a0 = lds128(sb + p.off_q + (uint32_t) b * 16u), a1 = lds128(sb + p.off_q + (uint32_t) (p.nsb + b) * 16u);
    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(ds.x), "=f"(ds.y) : "r"(sb + p.off_dx + (uint32_t) b * 8u));
}
template <> struct B32Dot<T_Q4_0> {
    static constexpr int BLK = 18;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int b) {
        uint32_t w[5];  // [d | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        ldsu_n<5>(row + (uint32_t) b * BLK, w);
        int a[8]; float2 ds;
        act_block<T_Q4_0>(p, sb, b, a, ds);
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = __funnelshift_r(w[i], w[i + 1], 16);
            sumi = dp4a_us(q & 0x0F0F0F0Fu, a[i], sumi);
            sumi = dp4a_us((q >> 4) & 0x0F0F0F0Fu, a[4 + i], sumi);
        }
        const float d = half_bits_to_float((uint16_t) (w[0] & 0xffffu));
        return d * (ds.x * (float) sumi) + (-8.0f * d) * ds.y;
    }
};
template <> struct B32Dot<T_Q4_1> {
    static constexpr int BLK = 20;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int b) {
        uint32_t w[5];  // [d m] [qs x4]
        ldsu_n<5>(row + (uint32_t) b * BLK, w);
        int a[8]; float2 ds;
        act_block<T_Q4_1>(p, sb, b, a, ds);
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            sumi = dp4a_us(w[1 + i] & 0x0F0F0F0Fu, a[i], sumi);
            sumi = dp4a_us((w[1 + i] >> 4) & 0x0F0F0F0Fu, a[4 + i], sumi);
        }
        return half_bits_to_float((uint16_t) (w[0] & 0xffffu)) * (ds.x * (float) sumi) + half_bits_to_float((uint16_t) (w[0] >> 16)) * ds.y;
    }
};
template <> struct B32Dot<T_Q5_0> {
    static constexpr int BLK = 22;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int b) {
        uint32_t w[6];  // [d | qh0 qh1] [qh2 qh3 | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        ldsu_n<6>(row + (uint32_t) b * BLK, w);
        int a[8]; float2 ds;
        act_block<T_Q5_0>(p, sb, b, a, ds);
        const uint32_t qh = __funnelshift_r(w[0], w[1], 16);
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = __funnelshift_r(w[1 + i], w[2 + i], 16);
            sumi = dp4a_us((q & 0x0F0F0F0Fu) | (spread4_(qh >> (4 * i)) << 4), a[i], sumi);
            sumi = dp4a_us(((q >> 4) & 0x0F0F0F0Fu) | (spread4_(qh >> (16 + 4 * i)) << 4), a[4 + i], sumi);
        }
        const float d = half_bits_to_float((uint16_t) (w[0] & 0xffffu));
        return d * (ds.x * (float) sumi) + (-16.0f * d) * ds.y;
    }
};
template <> struct B32Dot<T_Q5_1> {
    static constexpr int BLK = 24;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int b) {
        uint32_t w[6];  // [d m] [qh] [qs x4]
        ldsu_n<6>(row + (uint32_t) b * BLK, w);
        int a[8]; float2 ds;
        act_block<T_Q5_1>(p, sb, b, a, ds);
        const uint32_t qh = w[1];
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            sumi = dp4a_us((w[2 + i] & 0x0F0F0F0Fu) | (spread4_(qh >> (4 * i)) << 4), a[i], sumi);
            sumi = dp4a_us(((w[2 + i] >> 4) & 0x0F0F0F0Fu) | (spread4_(qh >> (16 + 4 * i)) << 4), a[4 + i], sumi);
        }
        return half_bits_to_float((uint16_t) (w[0] & 0xffffu)) * (ds.x * (float) sumi) + half_bits_to_float((uint16_t) (w[0] >> 16)) * ds.y;
    }
};
template <> struct B32Dot<T_Q8_0> {
    static constexpr int BLK = 34;
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int b) {
        uint32_t w[9];  // [d | q0 q1] ... [q30 q31 | next]
        ldsu_n<9>(row + (uint32_t) b * BLK, w);
        int a[8]; float2 ds;
        act_block<T_Q8_0>(p, sb, b, a, ds);
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) sumi = dp4a_ss((int) __funnelshift_r(w[i], w[i + 1], 16), a[i], sumi);
        return half_bits_to_float((uint16_t) (w[0] & 0xffffu)) * (ds.x * (float) sumi);
    }
};
__device__ __forceinline__ uint32_t prmt_(uint32_t a, uint32_t b, uint32_t sel) {  // bit 3 of a selector nibble: sign-replicate
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}
template <> struct B32Dot<T_MXFP4> {
    static constexpr int BLK = 17;
    // Each 16-bit slice of the nibble words IS a PRMT selector: the low 3 bits of a nibble index the magnitude table
    // (doubled E2M1 code points, ggml-common.h kvalues_mxfp4), bit 3 (the fp4 sign) turns the look-up into a sign
    // replication that yields the negative-lane mask.  Magnitudes split into a positive and a negative word, two
    // unsigned dp4a each against activations stored in the matching interleaved order (quantize_act32<true>).
    template <class P> static __device__ __forceinline__ float run(const P & p, uint32_t sb, uint32_t row, int b) {
        uint32_t w[5];  // [e | qs0 qs1 qs2] ... [qs15 | next]
        ldsu_n<5>(row + (uint32_t) b * BLK, w);
        int a[8]; float2 ds;
        act_block<T_MXFP4>(p, sb, b, a, ds);
        const uint32_t t0 = 0x03020100u, t1 = 0x0C080604u, sg = 0x80808080u;
        int sp = 0, sn = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q  = __funnelshift_r(w[i], w[i + 1], 8);
            const uint32_t q7 = q & 0x77777777u;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t mag = prmt_(t0, t1, q7 >> (16 * h));
                const uint32_t neg = mag & prmt_(sg, sg, q >> (16 * h)) & 0x7F7F7F7Fu;  // 0xFF where the sign is set, 0x80 elsewhere
                sp = dp4a_us(mag ^ neg, a[2 * i + h], sp);
                sn = dp4a_us(neg, a[2 * i + h], sn);
            }
        }
        return e8m0_half_to_float((uint8_t) (w[0] & 0xff)) * (ds.x * (float) (sp - sn));
    }
};

}  // namespace kqdot
}  // namespace b200q
