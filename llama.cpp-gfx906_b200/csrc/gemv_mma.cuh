// gemv_mma.cuh — decode GEMV for 2..8 activation columns (speculative / batched decode) on K-quant and 32-element block
// weights: launch parameters and the per-super-block / per-block tile arithmetic of gemv_mma.cu.
//
// Role in the reference: mul_mat_vec_q with ncols_dst = 2..8 (ggml-cuda/mmvq.cu:139-226, the ncols switch :237-498),
// which re-runs the per-column dp4a dot for every column.  Here one warp owns a 16-row x 8-column tile and contracts
// it with `mma.sync.m16n8k32` (u8 weights x s8 activations, s32): the weights are unpacked ONCE into A fragments and
// the instruction count per weight no longer grows with the number of columns (8 columns cost what 1 costs).
// The 6-bit sub-block scales multiply the INTEGER MMA results (the arithmetic of ggml-cpu's vec_dot_q4_K_q8_K /
// _q5_K_q8_K / _q6_K_q8_K, ggml-cpu/quants.c:550-758), activations are quantized per 256 values as in
// quantize_row_q8_K — the same numbers the n = 1 kernel (gemv_kq.cu) produces.
//
// The tile arithmetic is __host__ __device__ and templated on the MMA executor, so csrc/selftest_mma.cu runs the very
// same code on the CPU with an emulated MMA (fragment layouts of the PTX ISA) against oracle/oracle.c.
#pragma once
#include "blocks.cuh"

#include <stdint.h>
#ifndef __CUDA_ARCH__
#include <assert.h>
#include <string.h>
#endif

namespace b200q {

struct MmaParams {
    const uint8_t * w;    // [m][row_bytes] quantized rows, dense, 16-byte aligned
    const float *   act;  // column j at act + j * a_nb1 bytes (16-byte aligned)
    float *         dst;  // column j at dst + j * d_nb1 bytes, rows contiguous
    long long       a_nb1, d_nb1;
    int             m, k, n;  // n = 2..8 columns
    int             nsb;        // units per row: 256-element super-blocks (K-quants) | 32-element blocks
    int             row_bytes;  // nsb * block bytes (a multiple of 16)
    int             kc;         // units per ring stage (one k-chunk of 16 rows)
    int             chunk_bytes;  // kc * block bytes
    int             chunks;     // nsb / kc
    int             kp;         // warps that share one 16-row group along k (1, 2, 4 or 8)
    int             cpm;        // chunks per warp and group = chunks / kp
    int             groups;     // ceil(m / 16)
    int             pitch;      // shared-memory distance between the rows of a stage (== 32 or 96 mod 128: conflict-free)
    int             stage_bytes;  // 16 * pitch
    int             nstages;    // ring depth (>= number of warps)
    int             nwarps;     // 8 or 16 (block size = 32 * nwarps)
    int             early_prefetch;  // weights immutable: request the first stages before griddepcontrol.wait
    int             cluster;    // 1, or 2: CTA pairs share the activation quantization through distributed shared memory
    int             col_pitch;  // bytes between the int8 images of two columns (k + 64 | k + 16: conflict-free fragment loads)
    int             dx_pitch, bs_pitch;  // bytes between the columns' scales / partial sums (one unit of padding each)
    int             off_q, off_dx, off_bs, off_part, off_stage;  // shared-memory offsets
};

typedef void (*mma_kernel_t)(const MmaParams);
mma_kernel_t mma_kernel_for(int type);
int          mma_max_warps();
int          mma_bs_bytes(int type);  // bytes of activation partial sums per (column, super-block)

namespace mma {

struct U2 { uint32_t x, y; };
struct U4 { uint32_t x, y, z, w; };

// ---- shared memory access: 32-bit shared-window addresses on the device, offsets into host_smem in the selftest ----
#ifndef __CUDA_ARCH__
extern uint8_t * host_smem;
#endif
B200Q_HD uint32_t lds32(uint32_t a) {
    uint32_t v;
#ifdef __CUDA_ARCH__
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
#else
    assert(a % 4 == 0);
    memcpy(&v, host_smem + a, 4);
#endif
    return v;
}
B200Q_HD uint32_t lds16(uint32_t a) {
#ifdef __CUDA_ARCH__
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
#else
    uint16_t v;
    assert(a % 2 == 0);
    memcpy(&v, host_smem + a, 2);
    return v;
#endif
}
B200Q_HD U2 lds64(uint32_t a) {
    U2 v;
#ifdef __CUDA_ARCH__
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
#else
    assert(a % 8 == 0);
    memcpy(&v, host_smem + a, 8);
#endif
    return v;
}
B200Q_HD U4 lds128(uint32_t a) {
    U4 v;
#ifdef __CUDA_ARCH__
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
#else
    assert(a % 16 == 0);
    memcpy(&v, host_smem + a, 16);
#endif
    return v;
}
// one little-endian word from a 2-byte aligned address: aligned word (pair) + funnel shift
B200Q_HD uint32_t ldu32(uint32_t a) {
    const uint32_t w = a & ~3u;
    if ((a & 3u) == 0) return lds32(w);
    const uint32_t lo = lds32(w), hi = lds32(w + 4);
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, 16);
#else
    return (lo >> 16) | (hi << 16);
#endif
}

B200Q_HD float __uint_as_float_hd(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

// ---- integer helpers --------------------------------------------------------------------------------------------
B200Q_HD int dp2a_lo_su(int a, uint32_t b, int c) {  // a: 2 x s16, b: bytes 0, 1 unsigned
#ifdef __CUDA_ARCH__
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    return c + (int) (int16_t) (a & 0xffff) * (int) (b & 0xff) + (int) (int16_t) ((uint32_t) a >> 16) * (int) ((b >> 8) & 0xff);
#endif
}
B200Q_HD int dp2a_hi_su(int a, uint32_t b, int c) {  // bytes 2, 3 of b
#ifdef __CUDA_ARCH__
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    return c + (int) (int16_t) (a & 0xffff) * (int) ((b >> 16) & 0xff) + (int) (int16_t) ((uint32_t) a >> 16) * (int) ((b >> 24) & 0xff);
#endif
}
B200Q_HD int dp2a_lo_ss(int a, uint32_t b, int c) {
#ifdef __CUDA_ARCH__
    int d;
    asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    return c + (int) (int16_t) (a & 0xffff) * (int) (int8_t) (b & 0xff) + (int) (int16_t) ((uint32_t) a >> 16) * (int) (int8_t) ((b >> 8) & 0xff);
#endif
}
B200Q_HD int dp2a_hi_ss(int a, uint32_t b, int c) {
#ifdef __CUDA_ARCH__
    int d;
    asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    return c + (int) (int16_t) (a & 0xffff) * (int) (int8_t) ((b >> 16) & 0xff) + (int) (int16_t) ((uint32_t) a >> 16) * (int) (int8_t) ((b >> 24) & 0xff);
#endif
}

// byte idx of w, zero- / sign-extended: one PRMT
B200Q_HD int ubyte_of(uint32_t w, int idx) {
#ifdef __CUDA_ARCH__
    return (int) __byte_perm(w, 0, 0x4440 | idx);
#else
    return (int) ((w >> (8 * idx)) & 0xffu);
#endif
}
B200Q_HD int sbyte_of(uint32_t w, int idx) {
#ifdef __CUDA_ARCH__
    // bit 3 of a selector nibble replicates the sign of the selected byte (__byte_perm only honours 3 bits: raw prmt)
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(0u), "r"((uint32_t) (((idx | 8) * 0x1110) | idx)));
    return (int) d;
#else
    return (int) (int8_t) ((w >> (8 * idx)) & 0xffu);
#endif
}

// ---- layout of the quantized activations in shared memory --------------------------------------------------------
// A column's int8 image keeps 256 bytes per super-block, but permuted so that every lane of the MMA finds the B
// fragments of a whole super-block in four 16-byte chunks (LDS.128, conflict-free with col_pitch == 64 mod 128).
// act_word_off<T>(e): byte offset inside the super-block of the 4-byte word holding elements e .. e+3 (e % 4 == 0).
//   q4_K / q5_K (k32 MMA per 32-element sub-block j, lane column c owns elements 8c .. 8c+7 of it):
//       (j / 2) * 64 + c * 16 + (j & 1) * 8 + half * 4
//   q6_K (k16 MMA per 16-element scale group g = e / 16, lane column c owns elements 4c .. 4c+3 of it):
//       (g / 4) * 64 + c * 16 + (g & 3) * 4
template <int T> B200Q_HD int act_word_off(int e) {
    if (T == T_Q6_K) {
        const int g = e >> 4, c = (e & 15) >> 2;
        return (g >> 2) * 64 + c * 16 + (g & 3) * 4;
    }
    const int j = e >> 5, c = (e & 31) >> 3, h = (e & 7) >> 2;
    return (j >> 1) * 64 + c * 16 + (j & 1) * 8 + h * 4;
}

// where a warp finds things in shared memory (byte addresses)
struct TileCtx {
    uint32_t q, dx, bs;  // column 0 of: int8 images, scales [col][nsb], int16 partial sums [col][nsb][8 | 16]
    int      col_pitch, dx_pitch, bs_pitch;  // bytes between columns (each one unit longer than a column: the 4 columns a warp reads
                                             // at once then sit in different banks)
    int      n;
};

// ---- one super-block of a 16-row x 8-column tile -------------------------------------------------------------------
// lane = 4 * r + c : A fragments of rows r and r + 8 (shared addresses blk0 / blk1 of their blocks), B fragment of
// column r, C fragment = rows {r, r, r+8, r+8} x columns {2c, 2c+1, 2c, 2c+1}.  facc[i] += this super-block's share.
// E::mma32(id, C, A, b0, b1): C = A(16x32, u8) * B(32x8, s8);  E::mma32s: A is s8;  E::mma16: the 16x16 * 16x8 form.
template <int T, class E> struct TileSb;

// get_scale_min_k4 (ggml-quants.c:703-710) on whole words: bytes of *_lo = sub-blocks 0..3, *_hi = 4..7
struct K4Hdr {
    uint32_t sc_lo, sc_hi, mn_lo, mn_hi;
    float    d, dmin;
};
B200Q_HD K4Hdr k4_hdr(uint32_t blk) {
    const U4 h = lds128(blk);  // [d dmin] [scales x 12 bytes]
    K4Hdr    r;
    r.sc_lo = h.y & 0x3F3F3F3Fu;
    r.mn_lo = h.z & 0x3F3F3F3Fu;
    r.sc_hi = (h.w & 0x0F0F0F0Fu) | ((h.y >> 2) & 0x30303030u);
    r.mn_hi = ((h.w >> 4) & 0x0F0F0F0Fu) | ((h.z >> 2) & 0x30303030u);
    r.d     = half_bits_to_float((uint16_t) (h.x & 0xffffu));
    r.dmin  = half_bits_to_float((uint16_t) (h.x >> 16));
    return r;
}

template <int T, class E> struct TileSbK4 {  // q4_K, q5_K
    static B200Q_HD void run(E & ex, const TileCtx & t, int lane, int s, uint32_t blk0, uint32_t blk1, float (&facc)[4]) {
        constexpr int  QS = T == T_Q5_K ? 48 : 16;
        constexpr bool X16 = T == T_Q4_K;  // high nibbles stay in place (values x 16 still fit u8): one AND instead of SHR + AND
        const int     r = lane >> 2, c = lane & 3;
        const K4Hdr   h0 = k4_hdr(blk0), h1 = k4_hdr(blk1);
        // B fragment column: columns >= n read column n - 1 (their results are never stored)
        const uint32_t qb = t.q + (uint32_t) (r < t.n ? r : t.n - 1) * (uint32_t) t.col_pitch + (uint32_t) s * 256u + (uint32_t) c * 16u;
        U2 qh0 = { 0, 0 }, qh1 = { 0, 0 };
        if (T == T_Q5_K) {  // bit j of qh byte l: the 5th bit of element l of sub-block j
            qh0 = lds64(blk0 + 16 + 8 * c);
            qh1 = lds64(blk1 + 16 + 8 * c);
        }
        int acc[4] = { 0, 0, 0, 0 }, acch[4] = { 0, 0, 0, 0 };
#pragma unroll
        for (int pq = 0; pq < 4; ++pq) {
            const U2 w0 = lds64(blk0 + QS + 32 * pq + 8 * c), w1 = lds64(blk1 + QS + 32 * pq + 8 * c);
            const U4 b  = lds128(qb + 64 * pq);
            uint32_t alo[4] = { w0.x & 0x0F0F0F0Fu, w1.x & 0x0F0F0F0Fu, w0.y & 0x0F0F0F0Fu, w1.y & 0x0F0F0F0Fu };
            uint32_t ahi[4];
            if (X16) {
                ahi[0] = w0.x & 0xF0F0F0F0u; ahi[1] = w1.x & 0xF0F0F0F0u; ahi[2] = w0.y & 0xF0F0F0F0u; ahi[3] = w1.y & 0xF0F0F0F0u;
            } else {
                ahi[0] = (w0.x >> 4) & 0x0F0F0F0Fu; ahi[1] = (w1.x >> 4) & 0x0F0F0F0Fu;
                ahi[2] = (w0.y >> 4) & 0x0F0F0F0Fu; ahi[3] = (w1.y >> 4) & 0x0F0F0F0Fu;
            }
            if (T == T_Q5_K) {
                alo[0] |= ((qh0.x >> (2 * pq)) & 0x01010101u) << 4;
                alo[1] |= ((qh1.x >> (2 * pq)) & 0x01010101u) << 4;
                alo[2] |= ((qh0.y >> (2 * pq)) & 0x01010101u) << 4;
                alo[3] |= ((qh1.y >> (2 * pq)) & 0x01010101u) << 4;
                ahi[0] |= ((qh0.x >> (2 * pq + 1)) & 0x01010101u) << 4;
                ahi[1] |= ((qh1.x >> (2 * pq + 1)) & 0x01010101u) << 4;
                ahi[2] |= ((qh0.y >> (2 * pq + 1)) & 0x01010101u) << 4;
                ahi[3] |= ((qh1.y >> (2 * pq + 1)) & 0x01010101u) << 4;
            }
            int clo[4], chi[4];
            ex.mma32(2 * pq, clo, alo, b.x, b.y);
            ex.mma32(2 * pq + 1, chi, ahi, b.z, b.w);
            const uint32_t s0 = pq < 2 ? h0.sc_lo : h0.sc_hi, s1 = pq < 2 ? h1.sc_lo : h1.sc_hi;
            const int      s0a = ubyte_of(s0, 2 * (pq & 1)), s0b = ubyte_of(s0, 2 * (pq & 1) + 1);
            const int      s1a = ubyte_of(s1, 2 * (pq & 1)), s1b = ubyte_of(s1, 2 * (pq & 1) + 1);
            acc[0] += s0a * clo[0];
            acc[1] += s0a * clo[1];
            acc[2] += s1a * clo[2];
            acc[3] += s1a * clo[3];
            acch[0] += s0b * chi[0];
            acch[1] += s0b * chi[1];
            acch[2] += s1b * chi[2];
            acch[3] += s1b * chi[3];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i] += X16 ? acch[i] >> 4 : acch[i];  // exact: every term is a multiple of 16
        // - dmin * sum_j m_j * (sum of the activations of sub-block j), dx * d scaling: as ggml_vec_dot_q4_K_q8_K
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int      col = 2 * c + cc < t.n ? 2 * c + cc : t.n - 1;
            const U4       bs  = lds128(t.bs + (uint32_t) (col * t.bs_pitch + s * 16));  // 8 x int16 sums over 32 activations
            const float    dx  = __uint_as_float_hd(lds32(t.dx + (uint32_t) (col * t.dx_pitch + s * 4)));
            int m0 = dp2a_lo_su((int) bs.x, h0.mn_lo, 0);
            m0     = dp2a_hi_su((int) bs.y, h0.mn_lo, m0);
            m0     = dp2a_lo_su((int) bs.z, h0.mn_hi, m0);
            m0     = dp2a_hi_su((int) bs.w, h0.mn_hi, m0);
            int m1 = dp2a_lo_su((int) bs.x, h1.mn_lo, 0);
            m1     = dp2a_hi_su((int) bs.y, h1.mn_lo, m1);
            m1     = dp2a_lo_su((int) bs.z, h1.mn_hi, m1);
            m1     = dp2a_hi_su((int) bs.w, h1.mn_hi, m1);
            facc[cc]     += dx * (h0.d * (float) acc[cc] - h0.dmin * (float) m0);
            facc[2 + cc] += dx * (h1.d * (float) acc[2 + cc] - h1.dmin * (float) m1);
        }
    }
};

template <class E> struct TileSbQ6 {  // q6_K: blocks are only 2-byte aligned
    static B200Q_HD void run(E & ex, const TileCtx & t, int lane, int s, uint32_t blk0, uint32_t blk1, float (&facc)[4]) {
        const int      r = lane >> 2, c = lane & 3;
        const uint32_t qb = t.q + (uint32_t) (r < t.n ? r : t.n - 1) * (uint32_t) t.col_pitch + (uint32_t) s * 256u + (uint32_t) c * 16u;
        uint32_t sc0[4], sc1[4];  // 16 int8 scales of rows r, r + 8
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            sc0[i] = ldu32(blk0 + 192 + 4 * i);
            sc1[i] = ldu32(blk1 + 192 + 4 * i);
        }
        int acc[4] = { 0, 0, 0, 0 };
#pragma unroll
        for (int n = 0; n < 2; ++n) {  // 128-element halves
            uint32_t qh0[2], qh1[2];        // qh[32 n + 16 hl + 4c ..]
            uint32_t ql0[2][2], ql1[2][2];  // ql[64 n + 32 qq + 16 hl + 4c ..]: low nibbles -> quad qq, high nibbles -> quad qq + 2
#pragma unroll
            for (int hl = 0; hl < 2; ++hl) {
                qh0[hl] = ldu32(blk0 + 128 + 32 * n + 16 * hl + 4 * c);
                qh1[hl] = ldu32(blk1 + 128 + 32 * n + 16 * hl + 4 * c);
#pragma unroll
                for (int qq = 0; qq < 2; ++qq) {
                    ql0[qq][hl] = ldu32(blk0 + 64 * n + 32 * qq + 16 * hl + 4 * c);
                    ql1[qq][hl] = ldu32(blk1 + 64 * n + 32 * qq + 16 * hl + 4 * c);
                }
            }
#pragma unroll
            for (int up = 0; up < 2; ++up) {
                // this 16-byte chunk holds the scale groups g = 8 n + 4 up + {0, 1, 2, 3} = (quad q = 2 up + qq, half hl)
                const U4 b = lds128(qb + 64 * (2 * n + up));
#pragma unroll
                for (int qq = 0; qq < 2; ++qq) {
                    const int q = 2 * up + qq;  // elements 128 n + 32 q + l
#pragma unroll
                    for (int hl = 0; hl < 2; ++hl) {
                        uint32_t a[2];
                        a[0] = ((ql0[qq][hl] >> (4 * up)) & 0x0F0F0F0Fu) | (((qh0[hl] >> (2 * q)) & 0x03030303u) << 4);
                        a[1] = ((ql1[qq][hl] >> (4 * up)) & 0x0F0F0F0Fu) | (((qh1[hl] >> (2 * q)) & 0x03030303u) << 4);
                        const uint32_t bw = qq == 0 ? (hl == 0 ? b.x : b.y) : (hl == 0 ? b.z : b.w);
                        const int      g  = 8 * n + 2 * q + hl;
                        int            cf[4];
                        ex.mma16(g, cf, a, bw);
                        const int sa = sbyte_of(sc0[g >> 2], g & 3);
                        const int sb = sbyte_of(sc1[g >> 2], g & 3);
                        acc[0] += sa * cf[0];
                        acc[1] += sa * cf[1];
                        acc[2] += sb * cf[2];
                        acc[3] += sb * cf[3];
                    }
                }
            }
        }
        const float d0 = half_bits_to_float((uint16_t) lds16(blk0 + 208)), d1 = half_bits_to_float((uint16_t) lds16(blk1 + 208));
        // values are stored with +32: subtract 32 * sum_g sc_g * (sum of the activations of group g)
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int      col = 2 * c + cc < t.n ? 2 * c + cc : t.n - 1;
            const uint32_t ba  = t.bs + (uint32_t) (col * t.bs_pitch + s * 32);
            const U4       b0  = lds128(ba), b1 = lds128(ba + 16u);  // 16 x int16 sums over 16
            const float    dx  = __uint_as_float_hd(lds32(t.dx + (uint32_t) (col * t.dx_pitch + s * 4)));
            int o0 = dp2a_lo_ss((int) b0.x, sc0[0], 0), o1 = dp2a_lo_ss((int) b0.x, sc1[0], 0);
            o0 = dp2a_hi_ss((int) b0.y, sc0[0], o0); o1 = dp2a_hi_ss((int) b0.y, sc1[0], o1);
            o0 = dp2a_lo_ss((int) b0.z, sc0[1], o0); o1 = dp2a_lo_ss((int) b0.z, sc1[1], o1);
            o0 = dp2a_hi_ss((int) b0.w, sc0[1], o0); o1 = dp2a_hi_ss((int) b0.w, sc1[1], o1);
            o0 = dp2a_lo_ss((int) b1.x, sc0[2], o0); o1 = dp2a_lo_ss((int) b1.x, sc1[2], o1);
            o0 = dp2a_hi_ss((int) b1.y, sc0[2], o0); o1 = dp2a_hi_ss((int) b1.y, sc1[2], o1);
            o0 = dp2a_lo_ss((int) b1.z, sc0[3], o0); o1 = dp2a_lo_ss((int) b1.z, sc1[3], o1);
            o0 = dp2a_hi_ss((int) b1.w, sc0[3], o0); o1 = dp2a_hi_ss((int) b1.w, sc1[3], o1);
            facc[cc]     += dx * d0 * (float) (acc[cc] - 32 * o0);
            facc[2 + cc] += dx * d1 * (float) (acc[2 + cc] - 32 * o1);
        }
    }
};

// ---- 32-element block formats: one k32 MMA per block, float scales applied to the s32 tile ---------------------------
// activations: natural-order int8 image per column (col_pitch = k + 16: the LDS.32 of 8 columns x 4 lanes are
// conflict-free), (d, d * sum q) per (column, block) at t.dx — the arithmetic of quantize_row_q8_1 + vec_dot_q4_0_q8_0 /
// vec_dot_q8_0_q8_0 (ggml-cpu/quants.c:115-149, :305-333), as in gemv_kq.cu's B32Dot.
// lane = 4 r + c owns elements 4c .. 4c+3 and 16+4c .. 16+4c+3 of the block (the MMA's own k order).
B200Q_HD U2 lds64f(uint32_t a) { return lds64(a); }
template <int T, class E> struct TileB32;
template <class E> struct TileB32<T_Q4_0, E> {
    static B200Q_HD void run(E & ex, const TileCtx & t, int lane, int b, uint32_t blk0, uint32_t blk1, float (&facc)[4]) {
        const int      r = lane >> 2, c = lane & 3;
        const uint32_t qb = t.q + (uint32_t) (r < t.n ? r : t.n - 1) * (uint32_t) t.col_pitch + (uint32_t) b * 32u + (uint32_t) c * 4u;
        const uint32_t w0 = ldu32(blk0 + 2 + 4 * c), w1 = ldu32(blk1 + 2 + 4 * c);  // qs[4c .. 4c+3]: low nibbles = elements 4c.., high = 16+4c..
        const uint32_t a[4] = { w0 & 0x0F0F0F0Fu, w1 & 0x0F0F0F0Fu, (w0 >> 4) & 0x0F0F0F0Fu, (w1 >> 4) & 0x0F0F0F0Fu };
        int            cf[4];
        ex.mma32(0, cf, a, lds32(qb), lds32(qb + 16));
        const float d0 = half_bits_to_float((uint16_t) lds16(blk0)), d1 = half_bits_to_float((uint16_t) lds16(blk1));
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int col = 2 * c + cc < t.n ? 2 * c + cc : t.n - 1;
            const U2  ds  = lds64f(t.dx + (uint32_t) (col * t.dx_pitch + b * 8));
            const float dx = __uint_as_float_hd(ds.x), dxs = __uint_as_float_hd(ds.y);
            facc[cc]     += d0 * (dx * (float) cf[cc]) + (-8.0f * d0) * dxs;
            facc[2 + cc] += d1 * (dx * (float) cf[2 + cc]) + (-8.0f * d1) * dxs;
        }
    }
};
template <class E> struct TileB32<T_Q8_0, E> {
    static B200Q_HD void run(E & ex, const TileCtx & t, int lane, int b, uint32_t blk0, uint32_t blk1, float (&facc)[4]) {
        const int      r = lane >> 2, c = lane & 3;
        const uint32_t qb = t.q + (uint32_t) (r < t.n ? r : t.n - 1) * (uint32_t) t.col_pitch + (uint32_t) b * 32u + (uint32_t) c * 4u;
        const uint32_t a[4] = { ldu32(blk0 + 2 + 4 * c), ldu32(blk1 + 2 + 4 * c), ldu32(blk0 + 18 + 4 * c), ldu32(blk1 + 18 + 4 * c) };
        int            cf[4];
        ex.mma32s(0, cf, a, lds32(qb), lds32(qb + 16));
        const float d0 = half_bits_to_float((uint16_t) lds16(blk0)), d1 = half_bits_to_float((uint16_t) lds16(blk1));
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int   col = 2 * c + cc < t.n ? 2 * c + cc : t.n - 1;
            const float dx  = __uint_as_float_hd(lds32(t.dx + (uint32_t) (col * t.dx_pitch + b * 8)));
            facc[cc]     += d0 * (dx * (float) cf[cc]);
            facc[2 + cc] += d1 * (dx * (float) cf[2 + cc]);
        }
    }
};

// bit i of b (i = 0..3) -> bit 8 i: the 5th bits of q5_0 / q5_1 for 4 consecutive elements
B200Q_HD uint32_t spread4(uint32_t b) { return ((b & 0xFu) * 0x00204081u) & 0x01010101u; }

// q4_1 / q5_0 / q5_1: the q4_0 tile with (d, m) headers and / or a 5th-bit plane (vec_dot_q4_1_q8_1, _q5_0_q8_0,
// _q5_1_q8_1, ggml-cpu/quants.c:152-303).  HDR = bytes before qs, QH = offset of the 32 5th bits (or -1), M = has a min
template <int T, class E, int QS, int QH, bool M> struct TileB32N {
    static B200Q_HD void run(E & ex, const TileCtx & t, int lane, int b, uint32_t blk0, uint32_t blk1, float (&facc)[4]) {
        const int      r = lane >> 2, c = lane & 3;
        const uint32_t qb = t.q + (uint32_t) (r < t.n ? r : t.n - 1) * (uint32_t) t.col_pitch + (uint32_t) b * 32u + (uint32_t) c * 4u;
        const uint32_t w0 = ldu32(blk0 + QS + 4 * c), w1 = ldu32(blk1 + QS + 4 * c);
        uint32_t       a[4] = { w0 & 0x0F0F0F0Fu, w1 & 0x0F0F0F0Fu, (w0 >> 4) & 0x0F0F0F0Fu, (w1 >> 4) & 0x0F0F0F0Fu };
        if constexpr (QH >= 0) {
            const uint32_t h0 = ldu32(blk0 + (uint32_t) QH), h1 = ldu32(blk1 + (uint32_t) QH);
            a[0] |= spread4(h0 >> (4 * c)) << 4;
            a[1] |= spread4(h1 >> (4 * c)) << 4;
            a[2] |= spread4(h0 >> (16 + 4 * c)) << 4;
            a[3] |= spread4(h1 >> (16 + 4 * c)) << 4;
        }
        int cf[4];
        ex.mma32(0, cf, a, lds32(qb), lds32(qb + 16));
        const float d0 = half_bits_to_float((uint16_t) lds16(blk0)), d1 = half_bits_to_float((uint16_t) lds16(blk1));
        // the offset term: m * (d_x * sum q_x) for the formats with a min, else -(8 | 16) * d * (d_x * sum q_x)
        const float m0 = M ? half_bits_to_float((uint16_t) lds16(blk0 + 2)) : (QH >= 0 ? -16.0f : -8.0f) * d0;
        const float m1 = M ? half_bits_to_float((uint16_t) lds16(blk1 + 2)) : (QH >= 0 ? -16.0f : -8.0f) * d1;
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
            const int   col = 2 * c + cc < t.n ? 2 * c + cc : t.n - 1;
            const U2    ds  = lds64f(t.dx + (uint32_t) (col * t.dx_pitch + b * 8));
            const float dx = __uint_as_float_hd(ds.x), dxs = __uint_as_float_hd(ds.y);
            facc[cc]     += d0 * (dx * (float) cf[cc]) + m0 * dxs;
            facc[2 + cc] += d1 * (dx * (float) cf[2 + cc]) + m1 * dxs;
        }
    }
};

template <int T> struct MmaIsKQuant { static constexpr bool value = T == T_Q4_K || T == T_Q5_K || T == T_Q6_K; };

template <class E> struct TileSb<T_Q4_K, E> : TileSbK4<T_Q4_K, E> {};
template <class E> struct TileSb<T_Q5_K, E> : TileSbK4<T_Q5_K, E> {};
template <class E> struct TileSb<T_Q6_K, E> : TileSbQ6<E> {};
template <class E> struct TileSb<T_Q4_0, E> : TileB32<T_Q4_0, E> {};
template <class E> struct TileSb<T_Q8_0, E> : TileB32<T_Q8_0, E> {};
template <class E> struct TileSb<T_Q4_1, E> : TileB32N<T_Q4_1, E, 4, -1, true> {};
template <class E> struct TileSb<T_Q5_0, E> : TileB32N<T_Q5_0, E, 6, 2, false> {};
template <class E> struct TileSb<T_Q5_1, E> : TileB32N<T_Q5_1, E, 8, 4, true> {};

}  // namespace mma
}  // namespace b200q
