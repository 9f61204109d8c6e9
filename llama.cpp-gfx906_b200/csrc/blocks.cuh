// blocks.cuh — GGUF block formats as the kernels see them, plus the scalar decoders.
//
// The storage layouts are those of the reference's ggml/src/ggml-common.h (block_q4_0 :170-175,
// block_q4_1 :178-189, block_mxfp4 :190-195, block_q5_0 :198-203, block_q5_1 :206-218, block_q8_0 :219-224,
// block_q2_K :265-276, block_q3_K :282-288, block_q4_K :295-306, block_q5_K :312-324, block_q6_K :330-336, block_iq4_nl
// :416-420, block_iq4_xs :422-428) and are reused UNCHANGED so existing
// GGUF files work untouched.  We never dereference them as structs on the device (rows are only 1- or
// 2-byte aligned at block granularity) — every access goes through byte offsets below.
//
// Everything here is __host__ __device__ so the decoders and the integer-dot "unit" functions can be
// exercised on the CPU (csrc/selftest.cu) before any GPU time is spent.
#pragma once

#include <cuda_fp16.h>
#include <stdint.h>

#define B200Q_HD __host__ __device__ __forceinline__

namespace b200q {

enum : int {
    T_F32   = 0,
    T_F16   = 1,     // unquantized weights take the tensor-core kernel as "blocks" of 32 plain values
    T_BF16  = 30,
    T_F32W  = 1000,  // internal: f32 as a MUL_MAT weight type (32 values = 128 bytes); T_F32 itself keeps block size 1
    T_Q4_0  = 2,
    T_Q4_1  = 3,
    T_Q5_0  = 6,
    T_Q5_1  = 7,
    T_Q8_0  = 8,
    T_Q2_K  = 10,
    T_Q3_K  = 11,
    T_Q4_K  = 12,
    T_Q5_K  = 13,
    T_Q6_K  = 14,
    T_IQ4_NL = 20,
    T_IQ4_XS = 23,
    T_MXFP4 = 39,
};

template <int T> struct BlockInfo;  // ELEMS per block, BYTES per block, byte offsets of the fields
template <> struct BlockInfo<T_Q4_0>  { static constexpr int ELEMS = 32,  BYTES = 18;  };  // d:0  qs:2
template <> struct BlockInfo<T_Q4_1>  { static constexpr int ELEMS = 32,  BYTES = 20;  };  // d:0 m:2 qs:4
template <> struct BlockInfo<T_Q5_0>  { static constexpr int ELEMS = 32,  BYTES = 22;  };  // d:0 qh:2 qs:6
template <> struct BlockInfo<T_Q5_1>  { static constexpr int ELEMS = 32,  BYTES = 24;  };  // d:0 m:2 qh:4 qs:8
template <> struct BlockInfo<T_Q8_0>  { static constexpr int ELEMS = 32,  BYTES = 34;  };  // d:0 qs:2
template <> struct BlockInfo<T_MXFP4> { static constexpr int ELEMS = 32,  BYTES = 17;  };  // e:0 qs:1
template <> struct BlockInfo<T_IQ4_NL> { static constexpr int ELEMS = 32, BYTES = 18;  };  // d:0 qs:2 (code points kvalues_iq4nl)
template <> struct BlockInfo<T_Q2_K>  { static constexpr int ELEMS = 256, BYTES = 84;  };  // scales:0 qs:16 d:80 dmin:82
template <> struct BlockInfo<T_Q3_K>  { static constexpr int ELEMS = 256, BYTES = 110; };  // hmask:0 qs:32 scales:96 d:108
template <> struct BlockInfo<T_IQ4_XS> { static constexpr int ELEMS = 256, BYTES = 136; }; // d:0 scales_h:2 scales_l:4 qs:8
template <> struct BlockInfo<T_Q4_K>  { static constexpr int ELEMS = 256, BYTES = 144; };  // d:0 dmin:2 scales:4 qs:16
template <> struct BlockInfo<T_Q5_K>  { static constexpr int ELEMS = 256, BYTES = 176; };  // d:0 dmin:2 scales:4 qh:16 qs:48
template <> struct BlockInfo<T_Q6_K>  { static constexpr int ELEMS = 256, BYTES = 210; };  // ql:0 qh:128 scales:192 d:208

inline __host__ __device__ int blck_elems(int type) {
    switch (type) {
        case T_Q4_0: case T_Q4_1: case T_Q5_0: case T_Q5_1: case T_Q8_0: case T_MXFP4: case T_IQ4_NL: return 32;
        case T_Q4_K: case T_Q5_K: case T_Q6_K: case T_Q2_K: case T_Q3_K: case T_IQ4_XS: return 256;
        case T_F32: return 1;
        case T_F16: case T_BF16: case T_F32W: return 32;
        default: return 0;
    }
}
inline __host__ __device__ int blck_bytes(int type) {
    switch (type) {
        case T_Q4_0: return 18;  case T_Q4_1: return 20;  case T_Q5_0: return 22;  case T_Q5_1: return 24;
        case T_Q8_0: return 34;  case T_MXFP4: return 17; case T_Q4_K: return 144; case T_Q5_K: return 176;
        case T_Q6_K: return 210; case T_F32: return 4;
        case T_Q2_K: return 84;  case T_Q3_K: return 110; case T_IQ4_NL: return 18; case T_IQ4_XS: return 136;
        case T_F16: case T_BF16: return 64; case T_F32W: return 128;
        default: return 0;
    }
}

// ---------------------------------------------------------------------------------------------------------
// scalar helpers
// ---------------------------------------------------------------------------------------------------------
B200Q_HD float half_bits_to_float(uint16_t h) {
#ifdef __CUDA_ARCH__
    return __half2float(__ushort_as_half(h));
#else
    // exact IEEE binary16 -> binary32 (same values as GGML_FP16_TO_FP32, ggml-impl.h)
    uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    uint32_t exp  = (h >> 10) & 0x1f;
    uint32_t man  = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else {
            int e = -1;
            do { man <<= 1; ++e; } while (!(man & 0x400u));
            bits = sign | ((uint32_t) (127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
#endif
}

B200Q_HD uint16_t ld_u16(const uint8_t * p) { return (uint16_t) (p[0] | (p[1] << 8)); }
B200Q_HD uint32_t ld_u32_bytes(const uint8_t * p) {
    return (uint32_t) p[0] | ((uint32_t) p[1] << 8) | ((uint32_t) p[2] << 16) | ((uint32_t) p[3] << 24);
}

// E8M0 exponent -> 2^(e-128)  (ggml_e8m0_to_fp32_half, ggml-impl.h:451-469; subnormal for e < 2)
B200Q_HD float e8m0_half_to_float(uint8_t e) {
    uint32_t bits = e < 2 ? (0x00200000u << e) : ((uint32_t) (e - 1) << 23);
#ifdef __CUDA_ARCH__
    return __uint_as_float(bits);
#else
    float f;
    memcpy(&f, &bits, 4);
    return f;
#endif
}

// doubled E2M1 code points (kvalues_mxfp4, ggml-common.h:1094-1096)
B200Q_HD int mxfp4_value(int nib) {
    // 0, 1, 2, 3, 4, 6, 8, 12, 0, -1, -2, -3, -4, -6, -8, -12
    const int mag = nib & 7;
    const int v   = mag < 4 ? mag : (mag == 4 ? 4 : (mag == 5 ? 6 : (mag == 6 ? 8 : 12)));
    return (nib & 8) ? -v : v;
}

// rounding-exact fp32 arithmetic: the oracle (libggml-base) is built without FMA contraction, so
// `a*b - c` is a rounded multiply followed by a rounded subtract.
B200Q_HD float mul_rn(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fmul_rn(a, b);
#else
    volatile float r = a * b;
    return r;
#endif
}
B200Q_HD float add_rn(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fadd_rn(a, b);
#else
    volatile float r = a + b;
    return r;
#endif
}
B200Q_HD float sub_rn(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fsub_rn(a, b);
#else
    volatile float r = a - b;
    return r;
#endif
}

// 6-bit (scale, min) pair j of a K-quant super-block (get_scale_min_k4, ggml-quants.c:703-710)
B200Q_HD void k4_scale_min(int j, const uint8_t * q, int & sc, int & mn) {
    if (j < 4) {
        sc = q[j] & 63;
        mn = q[j + 4] & 63;
    } else {
        sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

// ---------------------------------------------------------------------------------------------------------
// decode_elem<T>(row, i): value of element i of a quantized row, bit-identical to dequantize_row_<T>
// (ggml-quants.c: q4_0 :307, q4_1 :327, q5_0 :348, q5_1 :374, q8_0 :401, mxfp4 :417, q4_K :1352,
//  q5_K :1554, q6_K :1762).  One fp32 rounding per arithmetic step, same order as the C code.
// ---------------------------------------------------------------------------------------------------------
template <int T> B200Q_HD float decode_elem(const uint8_t * row, int64_t i);

template <> B200Q_HD float decode_elem<T_Q4_0>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 5) * 18;
    const int       j = (int) (i & 31);
    const float     d = half_bits_to_float(ld_u16(b));
    const int       q = j < 16 ? (b[2 + j] & 0xF) : (b[2 + j - 16] >> 4);
    return mul_rn((float) (q - 8), d);
}
template <> B200Q_HD float decode_elem<T_Q4_1>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 5) * 20;
    const int       j = (int) (i & 31);
    const float     d = half_bits_to_float(ld_u16(b));
    const float     m = half_bits_to_float(ld_u16(b + 2));
    const int       q = j < 16 ? (b[4 + j] & 0xF) : (b[4 + j - 16] >> 4);
    return add_rn(mul_rn((float) q, d), m);
}
template <> B200Q_HD float decode_elem<T_Q5_0>(const uint8_t * row, int64_t i) {
    const uint8_t * b  = row + (i >> 5) * 22;
    const int       j  = (int) (i & 31);
    const float     d  = half_bits_to_float(ld_u16(b));
    const uint32_t  qh = ld_u32_bytes(b + 2);
    int             q;
    if (j < 16) {
        q = (b[6 + j] & 0xF) | (((qh >> j) << 4) & 0x10);
    } else {
        q = (b[6 + j - 16] >> 4) | ((qh >> (j - 16 + 12)) & 0x10);
    }
    return mul_rn((float) (q - 16), d);
}
template <> B200Q_HD float decode_elem<T_Q5_1>(const uint8_t * row, int64_t i) {
    const uint8_t * b  = row + (i >> 5) * 24;
    const int       j  = (int) (i & 31);
    const float     d  = half_bits_to_float(ld_u16(b));
    const float     m  = half_bits_to_float(ld_u16(b + 2));
    const uint32_t  qh = ld_u32_bytes(b + 4);
    int             q;
    if (j < 16) {
        q = (b[8 + j] & 0xF) | (((qh >> j) << 4) & 0x10);
    } else {
        q = (b[8 + j - 16] >> 4) | ((qh >> (j - 16 + 12)) & 0x10);
    }
    return add_rn(mul_rn((float) q, d), m);
}
template <> B200Q_HD float decode_elem<T_Q8_0>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 5) * 34;
    const float     d = half_bits_to_float(ld_u16(b));
    return mul_rn((float) (int8_t) b[2 + (i & 31)], d);
}
template <> B200Q_HD float decode_elem<T_MXFP4>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 5) * 17;
    const int       j = (int) (i & 31);
    const float     d = e8m0_half_to_float(b[0]);
    const int       q = j < 16 ? (b[1 + j] & 0xF) : (b[1 + j - 16] >> 4);
    return mul_rn((float) mxfp4_value(q), d);
}
template <> B200Q_HD float decode_elem<T_Q4_K>(const uint8_t * row, int64_t i) {
    const uint8_t * b    = row + (i >> 8) * 144;
    const int       r    = (int) (i & 255);
    const int       j64  = r >> 6;
    const int       hi   = (r >> 5) & 1;
    const int       l    = r & 31;
    const float     d    = half_bits_to_float(ld_u16(b));
    const float     dmin = half_bits_to_float(ld_u16(b + 2));
    int             sc, mn;
    k4_scale_min(2 * j64 + hi, b + 4, sc, mn);
    const float   d1 = mul_rn(d, (float) sc);
    const float   m1 = mul_rn(dmin, (float) mn);
    const uint8_t qb = b[16 + 32 * j64 + l];
    const int     q  = hi ? (qb >> 4) : (qb & 0xF);
    return sub_rn(mul_rn(d1, (float) q), m1);
}
template <> B200Q_HD float decode_elem<T_Q5_K>(const uint8_t * row, int64_t i) {
    const uint8_t * b    = row + (i >> 8) * 176;
    const int       r    = (int) (i & 255);
    const int       j64  = r >> 6;
    const int       hi   = (r >> 5) & 1;
    const int       l    = r & 31;
    const float     d    = half_bits_to_float(ld_u16(b));
    const float     dmin = half_bits_to_float(ld_u16(b + 2));
    int             sc, mn;
    k4_scale_min(2 * j64 + hi, b + 4, sc, mn);
    const float   d1 = mul_rn(d, (float) sc);
    const float   m1 = mul_rn(dmin, (float) mn);
    const uint8_t qb = b[48 + 32 * j64 + l];
    const int     hb = (b[16 + l] >> (2 * j64 + hi)) & 1;
    const int     q  = (hi ? (qb >> 4) : (qb & 0xF)) + (hb ? 16 : 0);
    return sub_rn(mul_rn(d1, (float) q), m1);
}
template <> B200Q_HD float decode_elem<T_Q6_K>(const uint8_t * row, int64_t i) {
    const uint8_t * b    = row + (i >> 8) * 210;
    const int       r    = (int) (i & 255);
    const int       n128 = r >> 7;
    const int       quad = (r >> 5) & 3;
    const int       l    = r & 31;
    const float     d    = half_bits_to_float(ld_u16(b + 208));
    const uint8_t   lb   = b[64 * n128 + l + ((quad & 1) ? 32 : 0)];
    const int       nib  = quad >= 2 ? (lb >> 4) : (lb & 0xF);
    const int       hb   = (b[128 + 32 * n128 + l] >> (2 * quad)) & 3;
    const int       q    = (int) (int8_t) (nib | (hb << 4)) - 32;
    const int       sc   = (int8_t) b[192 + 8 * n128 + (l >> 4) + 2 * quad];
    return mul_rn(mul_rn(d, (float) sc), (float) q);
}


inline __host__ __device__ bool dense_type(int type) { return type == T_F16 || type == T_BF16 || type == T_F32W; }

template <> B200Q_HD float decode_elem<T_F16>(const uint8_t * row, int64_t i) { return half_bits_to_float(ld_u16(row + 2 * i)); }
template <> B200Q_HD float decode_elem<T_BF16>(const uint8_t * row, int64_t i) {
    const uint32_t bits = (uint32_t) ld_u16(row + 2 * i) << 16;  // GGML_BF16_TO_FP32, ggml-impl.h
#ifdef __CUDA_ARCH__
    return __uint_as_float(bits);
#else
    float f;
    memcpy(&f, &bits, 4);
    return f;
#endif
}

// non-linear 4-bit code points (kvalues_iq4nl, ggml-common.h:1088-1090)
B200Q_HD int iq4nl_value(int nib) {
    // -127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113 packed as bytes of four words
    const uint32_t t[4] = { 0xBFAD9881u, 0xF6EADDCFu, 0x26190D01u, 0x71594535u };
    return (int) (int8_t) (t[(nib >> 2) & 3] >> (8 * (nib & 3)));
}
// 6-bit scale j (0..15) of a q3_K super-block from its 12 packed bytes (the aux[] shuffle of dequantize_row_q3_K,
// ggml-quants.c:1147-1152): low 4 bits in bytes 0..7 (low / high nibble), high 2 bits in bytes 8..11
B200Q_HD int q3k_scale(const uint8_t * sc, int j) {
    const int lo = j < 8 ? (sc[j] & 0xF) : (sc[j - 8] >> 4);
    const int hi = (sc[8 + (j & 3)] >> (2 * (j >> 2))) & 3;
    return (lo | (hi << 4)) - 32;
}

// dequantize_row_q2_K ggml-quants.c:784-814, _q3_K :1128-1176, _iq4_nl :2512-2528, _iq4_xs :2530-2550
template <> B200Q_HD float decode_elem<T_Q2_K>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 8) * 84;
    const int       r = (int) (i & 255);
    const int       n = r >> 7, j = (r >> 5) & 3, h = (r >> 4) & 1, l = r & 15;
    const float     d    = half_bits_to_float(ld_u16(b + 80));
    const float     dmin = half_bits_to_float(ld_u16(b + 82));
    const uint8_t   sc   = b[8 * n + 2 * j + h];
    const float     dl   = mul_rn(d, (float) (sc & 0xF));
    const float     ml   = mul_rn(dmin, (float) (sc >> 4));
    const int       q    = (b[16 + 32 * n + 16 * h + l] >> (2 * j)) & 3;
    return sub_rn(mul_rn(dl, (float) q), ml);
}
template <> B200Q_HD float decode_elem<T_Q3_K>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 8) * 110;
    const int       r = (int) (i & 255);
    const int       n = r >> 7, j = (r >> 5) & 3, h = (r >> 4) & 1, l = r & 15;
    const float     d  = half_bits_to_float(ld_u16(b + 108));
    const float     dl = mul_rn(d, (float) q3k_scale(b + 96, 8 * n + 2 * j + h));
    const int       q  = (b[32 + 32 * n + 16 * h + l] >> (2 * j)) & 3;
    const int       hm = (b[16 * h + l] >> (4 * n + j)) & 1;
    return mul_rn(dl, (float) (q - (hm ? 0 : 4)));
}
template <> B200Q_HD float decode_elem<T_IQ4_NL>(const uint8_t * row, int64_t i) {
    const uint8_t * b = row + (i >> 5) * 18;
    const int       j = (int) (i & 31);
    const float     d = half_bits_to_float(ld_u16(b));
    const int       q = j < 16 ? (b[2 + j] & 0xF) : (b[2 + j - 16] >> 4);
    return mul_rn(d, (float) iq4nl_value(q));
}
template <> B200Q_HD float decode_elem<T_IQ4_XS>(const uint8_t * row, int64_t i) {
    const uint8_t * b  = row + (i >> 8) * 136;
    const int       r  = (int) (i & 255);
    const int       ib = r >> 5, j = r & 31;
    const float     d  = half_bits_to_float(ld_u16(b));
    const int       sh = ld_u16(b + 2);
    const int       ls = ((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xF) | (((sh >> (2 * ib)) & 3) << 4);
    const float     dl = mul_rn(d, (float) (ls - 32));
    const uint8_t   qb = b[8 + 16 * ib + (j & 15)];
    return mul_rn(dl, (float) iq4nl_value(j < 16 ? (qb & 0xF) : (qb >> 4)));
}

}  // namespace b200q
