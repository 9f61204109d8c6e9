/*
 * oracle.c — plain-C restatement of the reference's CPU algorithm for the quantized matmul hot path.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Compiled with -ffp-contract=off so every fp32 operation
 * is rounded separately, like libggml-base (which the reference builds without -march flags).
 *
 * Each function names the reference code whose arithmetic it follows (paths under /root/reference/ggml/src).
 * The structure is ours: blocks are addressed as byte offsets (no struct overlay), K-quant scale unpacking
 * and the per-format integer dots are table/loop driven, and the integer sums are accumulated per
 * activation block and scaled in the same fp32 order as the *_generic kernels.
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------ */
/* scalars                                                                                          */
/* ------------------------------------------------------------------------------------------------ */
static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float    u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* IEEE binary16 -> binary32, exact (GGML_FP16_TO_FP32, ggml-impl.h:357-379) */
float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    const uint32_t e    = (h >> 10) & 31u;
    uint32_t       m    = h & 0x3ffu;
    if (e == 31) return u2f(sign | 0x7f800000u | (m << 13));
    if (e != 0)  return u2f(sign | ((e + 112u) << 23) | (m << 13));
    if (m == 0)  return u2f(sign);
    int sh = 0;
    while (!(m & 0x400u)) { m <<= 1; ++sh; }
    return u2f(sign | ((uint32_t) (113 - sh) << 23) | ((m & 0x3ffu) << 13));
}

/* binary32 -> binary16, round to nearest even, NaN -> 0x7e00 (GGML_FP32_TO_FP16, ggml-impl.h:381-405) */
uint16_t orc_fp32_to_fp16(float f) {
    const uint32_t w    = f2u(f);
    const uint16_t sign = (uint16_t) ((w >> 16) & 0x8000u);
    const uint32_t aw   = w & 0x7fffffffu;
    if (aw > 0x7f800000u) return sign | 0x7e00u;
    if (aw >= 0x477ff000u) return sign | 0x7c00u;          /* >= 65520 rounds to inf */
    if (aw < 0x33000001u) return sign;                     /* <= 2^-25 rounds to zero */
    int32_t  e = (int32_t) (aw >> 23) - 127;
    uint32_t m = (aw & 0x7fffffu) | 0x800000u;
    int      shift;
    uint32_t base;
    if (e < -14) { shift = 13 + (-14 - e); base = 0; }     /* subnormal half */
    else         { shift = 13; base = (uint32_t) (e + 15) << 10; m &= 0x7fffffu; }
    uint32_t q   = m >> shift;
    uint32_t rem = m & ((1u << shift) - 1u);
    uint32_t hlf = 1u << (shift - 1);
    if (rem > hlf || (rem == hlf && (q & 1u))) ++q;        /* carries propagate into the exponent */
    return sign | (uint16_t) (base + q);
}

/* 2^(e-128): ggml_e8m0_to_fp32_half (ggml-impl.h:451-469) */
static inline float e8m0_half(uint8_t e) { return u2f(e < 2 ? (0x00200000u << e) : ((uint32_t) (e - 1) << 23)); }

/* kvalues_mxfp4 (ggml-common.h:1094-1096) */
static const int8_t MXFP4_CODE[16] = { 0, 1, 2, 3, 4, 6, 8, 12, 0, -1, -2, -3, -4, -6, -8, -12 };

static inline uint16_t rd16(const uint8_t * p) { return (uint16_t) (p[0] | (p[1] << 8)); }
static inline void     wr16(uint8_t * p, uint16_t v) { p[0] = (uint8_t) v; p[1] = (uint8_t) (v >> 8); }
static inline float    rdh(const uint8_t * p) { return orc_fp16_to_fp32(rd16(p)); }

/* nearest_int (ggml-quants.c:444-449): round-half-even through the 1.5*2^23 trick */
static inline int nearest_int(float v) { return (int) (f2u(v + 12582912.f) & 0x007fffffu) - 0x00400000; }

/* ------------------------------------------------------------------------------------------------ */
/* type table (ggml.c:586-876, ggml-common.h block sizes)                                           */
/* ------------------------------------------------------------------------------------------------ */
int orc_blck_size(int t) {
    switch (t) {
        case ORC_F32: return 1;
        case ORC_Q4_0: case ORC_Q4_1: case ORC_Q5_0: case ORC_Q5_1: case ORC_Q8_0: case ORC_Q8_1: case ORC_MXFP4: case ORC_IQ4_NL: return 32;
        case ORC_Q4_K: case ORC_Q5_K: case ORC_Q6_K: case ORC_Q8_K: case ORC_Q2_K: case ORC_Q3_K: case ORC_IQ4_XS: return 256;
        default: return 0;
    }
}
size_t orc_type_size(int t) {
    switch (t) {
        case ORC_F32: return 4;   case ORC_Q4_0: return 18;  case ORC_Q4_1: return 20;  case ORC_Q5_0: return 22;
        case ORC_Q5_1: return 24; case ORC_Q8_0: return 34;  case ORC_Q8_1: return 36;  case ORC_MXFP4: return 17;
        case ORC_Q4_K: return 144; case ORC_Q5_K: return 176; case ORC_Q6_K: return 210; case ORC_Q8_K: return 292;
        case ORC_Q2_K: return 84; case ORC_Q3_K: return 110; case ORC_IQ4_NL: return 18; case ORC_IQ4_XS: return 136;
        default: return 0;
    }
}
size_t orc_row_size(int t, int64_t k) { int b = orc_blck_size(t); return b ? (size_t) (k / b) * orc_type_size(t) : 0; }

/* vec_dot_type column of type_traits_cpu (ggml-cpu/ggml-cpu.c:196-376) */
int orc_vec_dot_type(int t) {
    switch (t) {
        case ORC_Q4_0: case ORC_Q5_0: case ORC_Q8_0: case ORC_MXFP4: case ORC_IQ4_NL: return ORC_Q8_0;
        case ORC_Q4_1: case ORC_Q5_1: return ORC_Q8_1;
        case ORC_Q4_K: case ORC_Q5_K: case ORC_Q6_K: case ORC_Q2_K: case ORC_Q3_K: case ORC_IQ4_XS: return ORC_Q8_K;
        default: return -1;
    }
}

/* unpack the 8 (scale, min) 6-bit pairs of a q4_K/q5_K super-block (get_scale_min_k4, ggml-quants.c:703-710) */
static void k_scales(const uint8_t * q, uint8_t sc[8], uint8_t mn[8]) {
    for (int j = 0; j < 4; ++j) {
        sc[j]     = q[j] & 63;
        mn[j]     = q[j + 4] & 63;
        sc[j + 4] = (uint8_t) ((q[j + 8] & 0xF) | ((q[j] >> 6) << 4));
        mn[j + 4] = (uint8_t) ((q[j + 8] >> 4) | ((q[j + 4] >> 6) << 4));
    }
}

/* 5th bit of element j (0..31) of a q5_0/q5_1 block; same bit order as ggml-quants.c:361-362 */
static inline int q5_high(uint32_t qh, int j) { return (int) ((qh >> j) & 1u) << 4; }

/* integer code of element i (0..255) of a q6_K super-block, already re-centred (ggml-quants.c:1776-1779) */
static inline int q6k_code(const uint8_t * b, int i) {
    const int n = i >> 7, quad = (i >> 5) & 3, l = i & 31;
    const uint8_t lo = b[64 * n + l + 32 * (quad & 1)];
    const int nib = (quad & 2) ? (lo >> 4) : (lo & 15);
    const int hi  = (b[128 + 32 * n + l] >> (2 * quad)) & 3;
    return (int) (int8_t) (nib | (hi << 4)) - 32;
}
static inline int q6k_scale(const uint8_t * b, int i) {
    const int n = i >> 7, quad = (i >> 5) & 3, l = i & 31;
    return (int8_t) b[192 + 8 * n + 2 * quad + (l >> 4)];
}

/* kvalues_iq4nl (ggml-common.h:1088-1090) */
static const int8_t IQ4NL_CODE[16] = { -127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113 };

/* the 16 six-bit scales of a q3_K super-block, re-centred: the aux[] shuffle of ggml-quants.c:1147-1152 */
static void q3k_scales(const uint8_t * sc12, int8_t out[16]) {
    uint32_t aux[4];
    memcpy(aux, sc12, 12);
    const uint32_t kmask1 = 0x03030303u, kmask2 = 0x0f0f0f0fu, tmp = aux[2];
    aux[2] = ((aux[0] >> 4) & kmask2) | (((tmp >> 4) & kmask1) << 4);
    aux[3] = ((aux[1] >> 4) & kmask2) | (((tmp >> 6) & kmask1) << 4);
    aux[0] = (aux[0] & kmask2) | (((tmp >> 0) & kmask1) << 4);
    aux[1] = (aux[1] & kmask2) | (((tmp >> 2) & kmask1) << 4);
    memcpy(out, aux, 16);
    for (int j = 0; j < 16; ++j) out[j] = (int8_t) (out[j] - 32);
}
/* integer code of element i (0..255) of a q2_K (qs at +16) / q3_K (qs at +32, hmask at +0) super-block and the index of
 * its 16-element scale: elements come in the order (128-half n, shift j, 16-byte half h, l), ggml-quants.c:797-811 */
static inline int k23_pos(int i, int * is, int * byte, int * shift) {
    const int n = i >> 7, j = (i >> 5) & 3, h = (i >> 4) & 1, l = i & 15;
    *is = 8 * n + 2 * j + h; *byte = 32 * n + 16 * h + l; *shift = 2 * j;
    return 4 * n + j;  /* bit of hmask */
}
static inline int q2k_code(const uint8_t * b, int i, int * is) {
    int byte, shift; k23_pos(i, is, &byte, &shift);
    return (b[16 + byte] >> shift) & 3;
}
static inline int q3k_code(const uint8_t * b, int i, int * is) {
    int byte, shift; const int hb = k23_pos(i, is, &byte, &shift);
    return ((b[32 + byte] >> shift) & 3) - (((b[byte & 31] >> hb) & 1) ? 0 : 4);
}
static inline int iq4xs_scale(const uint8_t * b, int ib) {  /* :2541 */
    return (((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xf) | (((rd16(b + 2) >> (2 * ib)) & 3) << 4)) - 32;
}

/* ------------------------------------------------------------------------------------------------ */
/* dequantize_row_*                                                                                 */
/* ------------------------------------------------------------------------------------------------ */
int orc_dequantize_row(int t, const void * vsrc, float * y, int64_t k) {
    const uint8_t * src = (const uint8_t *) vsrc;
    const int       bs  = orc_blck_size(t);
    if (!bs || k % bs) return -1;
    const size_t  ts = orc_type_size(t);
    const int64_t nb = k / bs;
    for (int64_t ib = 0; ib < nb; ++ib, y += bs) {
        const uint8_t * b = src + (size_t) ib * ts;
        switch (t) {
            case ORC_Q4_0: {  /* ggml-quants.c:307-325 */
                const float d = rdh(b);
                for (int j = 0; j < 16; ++j) {
                    y[j]      = (float) ((b[2 + j] & 15) - 8) * d;
                    y[j + 16] = (float) ((b[2 + j] >> 4) - 8) * d;
                }
            } break;
            case ORC_Q4_1: {  /* :327-346 */
                const float d = rdh(b), m = rdh(b + 2);
                for (int j = 0; j < 16; ++j) {
                    y[j]      = (float) (b[4 + j] & 15) * d + m;
                    y[j + 16] = (float) (b[4 + j] >> 4) * d + m;
                }
            } break;
            case ORC_Q5_0: {  /* :348-372 */
                const float d = rdh(b);
                uint32_t    qh; memcpy(&qh, b + 2, 4);
                for (int j = 0; j < 16; ++j) {
                    y[j]      = (float) (((b[6 + j] & 15) | q5_high(qh, j)) - 16) * d;
                    y[j + 16] = (float) (((b[6 + j] >> 4) | q5_high(qh, j + 16)) - 16) * d;
                }
            } break;
            case ORC_Q5_1: {  /* :374-399 */
                const float d = rdh(b), m = rdh(b + 2);
                uint32_t    qh; memcpy(&qh, b + 4, 4);
                for (int j = 0; j < 16; ++j) {
                    y[j]      = (float) ((b[8 + j] & 15) | q5_high(qh, j)) * d + m;
                    y[j + 16] = (float) ((b[8 + j] >> 4) | q5_high(qh, j + 16)) * d + m;
                }
            } break;
            case ORC_Q8_0: {  /* :401-415 */
                const float d = rdh(b);
                for (int j = 0; j < 32; ++j) y[j] = (float) (int8_t) b[2 + j] * d;
            } break;
            case ORC_MXFP4: {  /* :417-435 */
                const float d = e8m0_half(b[0]);
                for (int j = 0; j < 16; ++j) {
                    y[j]      = (float) MXFP4_CODE[b[1 + j] & 15] * d;
                    y[j + 16] = (float) MXFP4_CODE[b[1 + j] >> 4] * d;
                }
            } break;
            case ORC_Q4_K: case ORC_Q5_K: {  /* :1352-1374, :1554-1579 */
                const float d = rdh(b), dmin = rdh(b + 2);
                uint8_t     sc[8], mn[8];
                k_scales(b + 4, sc, mn);
                const uint8_t * qh = b + 16;
                const uint8_t * qs = b + (t == ORC_Q4_K ? 16 : 48);
                for (int sb = 0; sb < 8; ++sb) {
                    const float dl = d * (float) sc[sb], ml = dmin * (float) mn[sb];
                    for (int l = 0; l < 32; ++l) {
                        const uint8_t byte = qs[32 * (sb >> 1) + l];
                        int           q    = (sb & 1) ? (byte >> 4) : (byte & 15);
                        if (t == ORC_Q5_K && ((qh[l] >> sb) & 1)) q += 16;
                        y[32 * sb + l] = dl * (float) q - ml;
                    }
                }
            } break;
            case ORC_Q2_K: {  /* :784-814 */
                const float d = rdh(b + 80), mn = rdh(b + 82);
                for (int i = 0; i < 256; ++i) {
                    int is; const int q = q2k_code(b, i, &is);
                    const float dl = d * (float) (b[is] & 0xF), ml = mn * (float) (b[is] >> 4);
                    y[i] = dl * (float) q - ml;
                }
            } break;
            case ORC_Q3_K: {  /* :1128-1176 */
                int8_t sc[16]; q3k_scales(b + 96, sc);
                const float d = rdh(b + 108);
                for (int i = 0; i < 256; ++i) {
                    int is; const int q = q3k_code(b, i, &is);
                    const float dl = d * (float) sc[is];
                    y[i] = dl * (float) q;
                }
            } break;
            case ORC_IQ4_NL: {  /* :2512-2528 */
                const float d = rdh(b);
                for (int j = 0; j < 16; ++j) {
                    y[j]      = d * (float) IQ4NL_CODE[b[2 + j] & 0xf];
                    y[j + 16] = d * (float) IQ4NL_CODE[b[2 + j] >> 4];
                }
            } break;
            case ORC_IQ4_XS: {  /* :2530-2550 */
                const float d = rdh(b);
                for (int ib = 0; ib < 8; ++ib) {
                    const float dl = d * (float) iq4xs_scale(b, ib);
                    for (int j = 0; j < 16; ++j) {
                        y[32 * ib + j]      = dl * (float) IQ4NL_CODE[b[8 + 16 * ib + j] & 0xf];
                        y[32 * ib + j + 16] = dl * (float) IQ4NL_CODE[b[8 + 16 * ib + j] >> 4];
                    }
                }
            } break;
            case ORC_Q6_K: {  /* :1762-1792 */
                const float d = rdh(b + 208);
                for (int i = 0; i < 256; ++i) y[i] = d * (float) q6k_scale(b, i) * (float) q6k_code(b, i);
            } break;
            default: return -1;
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------------ */
/* quantize_row_*_ref                                                                               */
/* ------------------------------------------------------------------------------------------------ */
int orc_quantize_row(int t, const float * x, void * vdst, int64_t k) {
    uint8_t * dst = (uint8_t *) vdst;
    const int bs  = orc_blck_size(t);
    if (!bs || k % bs) return -1;
    const size_t  ts = orc_type_size(t);
    const int64_t nb = k / bs;
    for (int64_t ib = 0; ib < nb; ++ib, x += bs) {
        uint8_t * b = dst + (size_t) ib * ts;
        switch (t) {
            case ORC_Q4_0: {  /* quantize_row_q4_0_ref ggml-quants.c:36-70 */
                float amax = 0.f, mx = 0.f;
                for (int j = 0; j < 32; ++j) if (amax < fabsf(x[j])) { amax = fabsf(x[j]); mx = x[j]; }
                const float d = mx / -8, id = d ? 1.0f / d : 0.0f;
                wr16(b, orc_fp32_to_fp16(d));
                for (int j = 0; j < 16; ++j) {
                    int q0 = (int8_t) (x[j] * id + 8.5f), q1 = (int8_t) (x[j + 16] * id + 8.5f);
                    if (q0 > 15) q0 = 15;
                    if (q1 > 15) q1 = 15;
                    b[2 + j] = (uint8_t) ((q0 & 0xff) | (q1 << 4));
                }
            } break;
            case ORC_Q8_0: case ORC_Q8_1: {  /* quantize_row_q8_0_ref :199-222, quantize_row_q8_1_ref :225-258 */
                float amax = 0.f;
                for (int j = 0; j < 32; ++j) amax = amax > fabsf(x[j]) ? amax : fabsf(x[j]);
                const float d = amax / 127, id = d ? 1.0f / d : 0.0f;
                wr16(b, orc_fp32_to_fp16(d));
                int8_t * q   = (int8_t *) (b + (t == ORC_Q8_0 ? 2 : 4));
                int      sum = 0;
                for (int j = 0; j < 32; ++j) { q[j] = (int8_t) roundf(x[j] * id); sum += q[j]; }
                if (t == ORC_Q8_1) wr16(b + 2, orc_fp32_to_fp16((float) sum * d));
            } break;
            case ORC_MXFP4: {  /* quantize_row_mxfp4_ref :273-305 */
                float amax = 0.f;
                for (int j = 0; j < 32; ++j) if (amax < fabsf(x[j])) amax = fabsf(x[j]);
                const uint8_t e = amax > 0.0f ? (uint8_t) (floorf(log2f(amax)) - 2 + 127) : 0;
                const float   d = e8m0_half(e);
                b[0] = e;
                for (int j = 0; j < 16; ++j) {
                    int best[2];
                    for (int h = 0; h < 2; ++h) {
                        const float v = x[j + 16 * h];
                        int   bi = 0;
                        float be = fabsf((float) MXFP4_CODE[0] * d - v);
                        for (int c = 1; c < 16; ++c) {
                            const float err = fabsf((float) MXFP4_CODE[c] * d - v);
                            if (err < be) { be = err; bi = c; }
                        }
                        best[h] = bi;
                    }
                    b[1 + j] = (uint8_t) (best[0] | (best[1] << 4));
                }
            } break;
            case ORC_Q8_K: {  /* quantize_row_q8_K_ref :2555-2592 */
                float mx = 0.f, amax = 0.f;
                for (int j = 0; j < 256; ++j) { const float ax = fabsf(x[j]); if (ax > amax) { amax = ax; mx = x[j]; } }
                int8_t *  q  = (int8_t *) (b + 4);
                int16_t * bs16 = (int16_t *) (b + 4 + 256);
                if (!amax) { memset(b, 0, 292); break; }
                const float iscale = -127.f / mx;
                for (int j = 0; j < 256; ++j) { int v = nearest_int(iscale * x[j]); q[j] = (int8_t) (v > 127 ? 127 : v); }
                for (int g = 0; g < 16; ++g) {
                    int s = 0;
                    for (int j = 0; j < 16; ++j) s += q[16 * g + j];
                    int16_t s16 = (int16_t) s;
                    memcpy(&bs16[g], &s16, 2);
                }
                const float dd = 1 / iscale;
                memcpy(b, &dd, 4);
            } break;
            default: return -1;
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------------ */
/* vec_dot: ggml_vec_dot_<w>_<a>_generic (ggml-cpu/quants.c: q4_0 :115, q4_1 :152, mxfp4 :188, q5_0 :219,  */
/* q5_1 :262, q8_0 :305, q4_K :550, q5_K :626, q6_K :705)                                                 */
/* ------------------------------------------------------------------------------------------------ */
static int code32(int t, const uint8_t * b, int j) {  /* integer weight code of element j of a 32-block */
    switch (t) {
        case ORC_Q4_0: return (j < 16 ? (b[2 + j] & 15) : (b[2 + j - 16] >> 4)) - 8;
        case ORC_Q4_1: return j < 16 ? (b[4 + j] & 15) : (b[4 + j - 16] >> 4);
        case ORC_Q5_0: { uint32_t qh; memcpy(&qh, b + 2, 4);
                         return ((j < 16 ? (b[6 + j] & 15) : (b[6 + j - 16] >> 4)) | q5_high(qh, j)) - 16; }
        case ORC_Q5_1: { uint32_t qh; memcpy(&qh, b + 4, 4);
                         return (j < 16 ? (b[8 + j] & 15) : (b[8 + j - 16] >> 4)) | q5_high(qh, j); }
        case ORC_Q8_0: return (int8_t) b[2 + j];
        case ORC_MXFP4: return MXFP4_CODE[j < 16 ? (b[1 + j] & 15) : (b[1 + j - 16] >> 4)];
        case ORC_IQ4_NL: return IQ4NL_CODE[j < 16 ? (b[2 + j] & 15) : (b[2 + j - 16] >> 4)];
        default: return 0;
    }
}

int orc_vec_dot(int t, int64_t k, const void * vw, const void * va, float * out) {
    const uint8_t * w = (const uint8_t *) vw;
    const uint8_t * a = (const uint8_t *) va;
    const int       at = orc_vec_dot_type(t);
    if (at < 0 || k % orc_blck_size(t)) return -1;
    const size_t wts = orc_type_size(t), ats = orc_type_size(at);
    float        sumf = 0.f;

    if (at == ORC_Q8_0 || at == ORC_Q8_1) {
        for (int64_t ib = 0; ib < k / 32; ++ib) {
            const uint8_t * wb = w + (size_t) ib * wts;
            const uint8_t * ab = a + (size_t) ib * ats;
            const int8_t *  q8 = (const int8_t *) (ab + (at == ORC_Q8_0 ? 2 : 4));
            int s0 = 0, s1 = 0;
            for (int j = 0; j < 16; ++j) {
                s0 += code32(t, wb, j) * q8[j];
                s1 += code32(t, wb, j + 16) * q8[j + 16];
            }
            const int   sumi = s0 + s1;
            const float dy   = rdh(ab);
            switch (t) {
                case ORC_Q4_0:  sumf += (float) sumi * rdh(wb) * dy; break;                    /* :146 */
                case ORC_Q8_0:  sumf += (float) sumi * (rdh(wb) * dy); break;                  /* :329 */
                case ORC_Q5_0:  sumf += (rdh(wb) * dy) * (float) sumi; break;                  /* :257 */
                case ORC_IQ4_NL: sumf += (dy * rdh(wb)) * (float) sumi; break;                 /* :1126-1133 */
                case ORC_MXFP4: sumf += (dy * e8m0_half(wb[0])) * (float) sumi; break;        /* :204-214 */
                case ORC_Q4_1: case ORC_Q5_1:
                    sumf += (rdh(wb) * dy) * (float) sumi + rdh(wb + 2) * rdh(ab + 2); break;  /* :182, :299 */
                default: return -1;
            }
        }
        *out = sumf;
        return 0;
    }

    /* K-quants against q8_K: eight fp32 lanes, lane l sums elements whose index is l mod 8 */
    float lanes[8] = { 0 };
    for (int64_t ib = 0; ib < k / 256; ++ib) {
        const uint8_t * wb = w + (size_t) ib * wts;
        const uint8_t * ab = a + (size_t) ib * 292;
        float           dy; memcpy(&dy, ab, 4);
        const int8_t *  q8 = (const int8_t *) (ab + 4);
        int32_t         acc[8] = { 0 };
        if (t == ORC_Q2_K) {  /* ggml_vec_dot_q2_K_q8_K_generic, ggml-cpu/quants.c:419-469: plain integer sums, one fp32 update */
            int16_t bsum[16]; memcpy(bsum, ab + 4 + 256, 32);
            int     summs = 0, isum = 0;
            for (int j = 0; j < 16; ++j) summs += bsum[j] * (wb[j] >> 4);
            for (int i = 0; i < 256; ++i) { int is; const int q = q2k_code(wb, i, &is); isum += (wb[is] & 0xF) * (q8[i] * q); }
            const float dall = dy * rdh(wb + 80), dmin = dy * rdh(wb + 82);
            sumf += dall * (float) isum - dmin * (float) summs;
        } else if (t == ORC_Q3_K) {  /* _q3_K_q8_K_generic :471-548: eight lanes like q6_K */
            int8_t sc[16]; q3k_scales(wb + 96, sc);
            for (int i = 0; i < 256; ++i) { int is; const int q = q3k_code(wb, i, &is); acc[i & 7] += sc[is] * (int16_t) (q8[i] * q); }
            const float d = rdh(wb + 108) * dy;
            for (int l = 0; l < 8; ++l) lanes[l] += d * (float) acc[l];
        } else if (t == ORC_IQ4_XS) {  /* _iq4_xs_q8_K_generic :1137-1178: one fp32 update per 32-element sub-block */
            const float d4d8 = rdh(wb) * dy;
            for (int ib = 0; ib < 8; ++ib) {
                int s1 = 0, s2 = 0;
                for (int j = 0; j < 16; ++j) {
                    s1 += q8[32 * ib + j] * IQ4NL_CODE[wb[8 + 16 * ib + j] & 0xf];
                    s2 += q8[32 * ib + j + 16] * IQ4NL_CODE[wb[8 + 16 * ib + j] >> 4];
                }
                sumf += (d4d8 * (float) iq4xs_scale(wb, ib)) * (float) (s1 + s2);
            }
        } else if (t == ORC_Q6_K) {
            for (int i = 0; i < 256; ++i) acc[i & 7] += q6k_scale(wb, i) * (int16_t) (q8[i] * q6k_code(wb, i));
            const float d = rdh(wb + 208) * dy;
            for (int l = 0; l < 8; ++l) lanes[l] += d * (float) acc[l];
        } else {
            uint8_t sc[8], mn[8];
            k_scales(wb + 4, sc, mn);
            const uint8_t * qh = wb + 16;
            const uint8_t * qs = wb + (t == ORC_Q4_K ? 16 : 48);
            int16_t         bsum[16]; memcpy(bsum, ab + 4 + 256, 32);
            int             sumi = 0;
            for (int g = 0; g < 16; ++g) sumi += bsum[g] * mn[g >> 1];
            for (int i = 0; i < 256; ++i) {
                const int     sb   = i >> 5, l = i & 31;
                const uint8_t byte = qs[32 * (sb >> 1) + l];
                int           q    = (sb & 1) ? (byte >> 4) : (byte & 15);
                if (t == ORC_Q5_K && ((qh[l] >> sb) & 1)) q += 16;
                acc[i & 7] += (int32_t) sc[sb] * (int16_t) (q8[i] * q);
            }
            const float d = rdh(wb) * dy;
            for (int l = 0; l < 8; ++l) lanes[l] += d * (float) acc[l];
            const float dmin = rdh(wb + 2) * dy;
            sumf -= dmin * (float) sumi;
        }
    }
    for (int l = 0; l < 8; ++l) sumf += lanes[l];
    *out = sumf;
    return 0;
}

/* ------------------------------------------------------------------------------------------------ */
/* MUL_MAT / MUL_MAT_ID / GET_ROWS drivers                                                           */
/* ------------------------------------------------------------------------------------------------ */
int orc_mul_mat(int t, const void * w, size_t w_row_stride, const float * act, float * dst, int64_t m, int64_t k,
                int64_t n) {
    const int at = orc_vec_dot_type(t);
    if (at < 0) return -1;
    const size_t arow = orc_row_size(at, k);
    uint8_t *    aq   = (uint8_t *) malloc(arow ? arow : 1);
    if (!aq) return -2;
    int rc = 0;
    for (int64_t j = 0; j < n && !rc; ++j) {
        rc = orc_quantize_row(at, act + j * k, aq, k);  /* ggml-cpu.c:1264-1299 */
        for (int64_t i = 0; i < m && !rc; ++i) {          /* mul_mat_one_chunk :1112-1200 */
            rc = orc_vec_dot(t, k, (const uint8_t *) w + (size_t) i * w_row_stride, aq, &dst[j * m + i]);
        }
    }
    free(aq);
    return rc;
}

int orc_mul_mat_id(int t, const void * as, const float * b, const int32_t * ids, float * dst, int64_t m, int64_t k,
                   int64_t n_expert, int64_t n_used, int64_t n_tokens, int64_t b_ne1, int64_t ids_stride) {
    const size_t rs = orc_row_size(t, k);
    for (int64_t tk = 0; tk < n_tokens; ++tk) {
        for (int64_t s = 0; s < n_used; ++s) {
            const int32_t e = ids[tk * ids_stride + s];
            if (e < 0 || e >= n_expert) return -3;
            const float * x = b + (tk * b_ne1 + (b_ne1 == 1 ? 0 : s)) * k;   /* i11 = id % ne11, ggml-cpu.c:1608 */
            int rc = orc_mul_mat(t, (const uint8_t *) as + (size_t) e * (size_t) m * rs, rs, x,
                                 dst + (tk * n_used + s) * m, m, k, 1);
            if (rc) return rc;
        }
    }
    return 0;
}

int orc_get_rows(int t, const void * src, const int32_t * rows, float * dst, int64_t k, int64_t n_idx) {
    const size_t rs = orc_row_size(t, k);
    for (int64_t i = 0; i < n_idx; ++i) {
        if (t == ORC_F32) {
            memcpy(dst + i * k, (const uint8_t *) src + (size_t) rows[i] * (size_t) k * 4, (size_t) k * 4);
        } else if (orc_dequantize_row(t, (const uint8_t *) src + (size_t) rows[i] * rs, dst + i * k, k)) {
            return -1;
        }
    }
    return 0;
}
