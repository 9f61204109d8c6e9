// units.cuh — integer-dot "units" for the decode GEMV.
//
// A unit is the smallest piece of a quantized row that one lane decodes on its own:
//   * 32-element block formats (q4_0 q4_1 q5_0 q5_1 q8_0 mxfp4): one block            -> 1 group of 8 words
//   * q4_K / q5_K: one 64-element pair of sub-blocks sharing 32 nibble bytes           -> 2 groups of 8 words
//   * q6_K: 4 x 16 elements that share 4 ql words (lo) + 4 ql words (hi) + 4 qh words  -> 4 groups of 4 words
// decode() turns the unit into int8x4 words `wq` (in activation order), one fp32 scale `ws` per group and
// one fp32 offset `wm` per group such that
//      sum_{i in group} w_i * x_i  ~=  ws * d_x * dot(wq, q_x)  +  wm * s_x ,    s_x = d_x * sum(q_x)
// with (q_x, d_x) the q8 quantization of the activation block the group lies in.  This is the arithmetic of the
// reference's vec_dot_*_q8_1 device functions (ggml-cuda/vecdotq.cuh:51-70 q4_0, :179-191 q8_0, :243-262 mxfp4,
// :409-431 q4_K, :528-548 q6_K), restated; it is NOT bit-comparable with ggml-cpu (which uses q8_0/q8_K
// activations) and is gated by NMSE <= 5e-4 like the reference CUDA backend (tests/test-backend-ops.cpp:3106).
//
// Rows are byte streams: `p` may have ANY byte alignment (17/18/34/210-byte blocks), so all weight
// loads go through ldw(), an aligned 32-bit load pair + funnel shift.  The caller guarantees 8 readable bytes of
// slack behind the row.
#pragma once

#include "blocks.cuh"

namespace b200q {

// ---- intrinsic shims so the same code runs in csrc/selftest on the host ----------------------------------
B200Q_HD int dp4a_(int a, int b, int c) {
#ifdef __CUDA_ARCH__
    return __dp4a(a, b, c);
#else
    for (int i = 0; i < 4; ++i) {
        c += (int) (int8_t) (a >> (8 * i)) * (int) (int8_t) (b >> (8 * i));
    }
    return c;
#endif
}
// a: four UNSIGNED bytes, b: four signed bytes (dp4a.u32.s32)
B200Q_HD int dp4a_us(uint32_t a, int b, int c) {
#ifdef __CUDA_ARCH__
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    for (int i = 0; i < 4; ++i) {
        c += (int) (uint8_t) (a >> (8 * i)) * (int) (int8_t) (b >> (8 * i));
    }
    return c;
#endif
}
B200Q_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, int shift_bits) {
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, shift_bits);
#else
    return shift_bits ? (lo >> shift_bits) | (hi << (32 - shift_bits)) : lo;
#endif
}
B200Q_HD uint32_t vsub4_(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    return __vsub4(a, b);
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        r |= (uint32_t) (uint8_t) ((uint8_t) (a >> (8 * i)) - (uint8_t) (b >> (8 * i))) << (8 * i);
    }
    return r;
#endif
}
B200Q_HD uint32_t byte_perm_(uint32_t a, uint32_t b, uint32_t s) {
#ifdef __CUDA_ARCH__
    return __byte_perm(a, b, s);
#else
    uint64_t v = ((uint64_t) b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        r |= (uint32_t) ((v >> (8 * ((s >> (4 * i)) & 7))) & 0xff) << (8 * i);
    }
    return r;
#endif
}

// ---- address abstraction ----------------------------------------------------------------------------------
// The decoders are templates on the address type A so the SAME code runs
//   * on the device with A = uint32_t: a byte offset into the CTA's dynamic shared memory (b200q_smem) — loads
//     are emitted as LDS because the base is the __shared__ symbol itself (no generic-pointer LD), and all
//     address arithmetic is 32-bit;
//   * on the host (selftest) with A = const uint8_t *.
#ifdef __CUDACC__
extern __shared__ __align__(128) uint8_t b200q_smem[];
__device__ __forceinline__ uint32_t lds32(uint32_t off) { return *(const uint32_t *) (b200q_smem + off); }
__device__ __forceinline__ int4     lds128(uint32_t off) { return *(const int4 *) (b200q_smem + off); }
__device__ __forceinline__ float2   lds64f(uint32_t off) { return *(const float2 *) (b200q_smem + off); }
__device__ __forceinline__ uint32_t addr_align4(uint32_t a) { return a & ~3u; }
__device__ __forceinline__ int      addr_shift(uint32_t a) { return (int) (a & 3u) * 8; }
#endif
B200Q_HD uint32_t        lds32(const uint8_t * p) { return *(const uint32_t *) p; }
struct Words4 { uint32_t x, y, z, w; };
#ifdef __CUDACC__
__device__ __forceinline__ Words4 ldq(uint32_t off) {  // one LDS.128, `off` 16-byte aligned
    const int4 t = *(const int4 *) (b200q_smem + off);
    return Words4{ (uint32_t) t.x, (uint32_t) t.y, (uint32_t) t.z, (uint32_t) t.w };
}
__device__ __forceinline__ bool addr_is16(uint32_t a) { return (a & 15u) == 0; }
#endif
B200Q_HD Words4 ldq(const uint8_t * p) {
    const uint32_t * w = (const uint32_t *) p;
    return Words4{ w[0], w[1], w[2], w[3] };
}
B200Q_HD bool addr_is16(const uint8_t * p) { return ((uintptr_t) p & 15) == 0; }
template <int N, class A> B200Q_HD void ldw_n(A p, uint32_t (&out)[N]);
// N words (N % 4 == 0) from a byte address: LDS.128s when the address is 16-byte aligned (warp-uniform in
// practice: row pitch and block size of the K-quants are multiples of 16), else the funnel-shift path
template <int N, class A> B200Q_HD void ldw_n_fast(A p, uint32_t (&out)[N]) {
    if (addr_is16(p)) {
#pragma unroll
        for (int i = 0; i < N / 4; ++i) {
            const Words4 t = ldq(p + 16 * i);
            out[4 * i + 0] = t.x;
            out[4 * i + 1] = t.y;
            out[4 * i + 2] = t.z;
            out[4 * i + 3] = t.w;
        }
    } else {
        ldw_n<N>(p, out);
    }
}
B200Q_HD const uint8_t * addr_align4(const uint8_t * p) { return (const uint8_t *) ((uintptr_t) p & ~(uintptr_t) 3); }
B200Q_HD int             addr_shift(const uint8_t * p) { return (int) ((uintptr_t) p & 3) * 8; }

// 32-bit little-endian word at an arbitrary byte address (reads the two aligned words that cover it)
template <class A> B200Q_HD uint32_t ldw(A p) {
    const A   w  = addr_align4(p);
    const int sh = addr_shift(p);
    return funnel_r(lds32(w), lds32(w + 4), sh);
}
// N consecutive words starting at an arbitrary byte address: N+1 aligned loads, N funnel shifts
template <int N, class A> B200Q_HD void ldw_n(A p, uint32_t (&out)[N]) {
    const A   w    = addr_align4(p);
    const int sh   = addr_shift(p);
    uint32_t  prev = lds32(w);
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const uint32_t next = lds32(w + 4 * (i + 1));
        out[i]              = funnel_r(prev, next, sh);
        prev                = next;
    }
}

// spread 4 bits (bit i -> bit 8*i) : used for the 5th bit of q5_0 / q5_1 / q5_K
B200Q_HD uint32_t spread4(uint32_t b) { return ((b & 0xFu) * 0x00204081u) & 0x01010101u; }

// 8 nibbles (one word) -> two int8x4 words through the mxfp4 code-point table, lo nibbles and hi nibbles
B200Q_HD void mxfp4_lut(uint32_t q, uint32_t & lo, uint32_t & hi) {
    // table bytes 0..15: 0,1,2,3 | 4,6,8,12 | 0,-1,-2,-3 | -4,-6,-8,-12
    const uint32_t t0 = 0x03020100u, t1 = 0x0C080604u, t2 = 0xFDFEFF00u, t3 = 0xF4F8FAFCu;
    uint32_t       r[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t n   = (q >> (4 * h)) & 0x0F0F0F0Fu;      // 4 nibble indices, one per byte
        const uint32_t sel = n & 0x07070707u;                    // index within the selected 8-entry half
        // prmt selector nibbles: byte i of sel goes to selector nibble i
        const uint32_t s   = (sel & 0xF) | ((sel >> 4) & 0xF0) | ((sel >> 8) & 0xF00) | ((sel >> 12) & 0xF000);
        const uint32_t pos = byte_perm_(t0, t1, s);              // table[idx & 7]        (non-negative half)
        const uint32_t neg = byte_perm_(t2, t3, s);              // table[8 + (idx & 7)]  (negative half)
        // per-byte select on bit 3 of the nibble
        const uint32_t m   = ((n >> 3) & 0x01010101u) * 0xFFu;
        r[h]               = (pos & ~m) | (neg & m);
    }
    lo = r[0];
    hi = r[1];
}

// ---------------------------------------------------------------------------------------------------------
// Unit<T>
//   ELEMS        elements covered by one unit
//   NG, GW       groups per unit, activation words per group
//   act_word(u,g) index of the first int8x4 activation word of group g of unit u
//   act_blk(u,g)  index of the 32-element activation block (for d_x, s_x)
//   decode(row, u, wq, ws, wm)
// ---------------------------------------------------------------------------------------------------------
template <int T> struct Unit;

template <int T, int BYTES> struct UnitBlock32 {
    static constexpr int  ELEMS = 32, NG = 1, GW = 8;
    static constexpr bool UNSIGNED_W = false, MIN16 = false;
    static B200Q_HD int  act_word(int u, int) { return u * 8; }
    static B200Q_HD int  act_blk(int u, int) { return u; }
};

template <> struct Unit<T_Q4_0> : UnitBlock32<T_Q4_0, 18> {
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A b = row + u * 18;
        uint32_t        w[5];
        ldw_n<5>(b, w);  // [d | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        const float d = half_bits_to_float((uint16_t) (w[0] & 0xffff));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = funnel_r(w[i], w[i + 1], 16);
            wq[i]            = (int) (q & 0x0F0F0F0Fu);
            wq[4 + i]        = (int) ((q >> 4) & 0x0F0F0F0Fu);
        }
        ws[0] = d;
        wm[0] = -8.0f * d;
    }
};
template <> struct Unit<T_Q4_1> : UnitBlock32<T_Q4_1, 20> {
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A b = row + u * 20;
        uint32_t        w[5];
        ldw_n<5>(b, w);  // [d m] [qs x4]
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            wq[i]     = (int) (w[1 + i] & 0x0F0F0F0Fu);
            wq[4 + i] = (int) ((w[1 + i] >> 4) & 0x0F0F0F0Fu);
        }
        ws[0] = half_bits_to_float((uint16_t) (w[0] & 0xffff));
        wm[0] = half_bits_to_float((uint16_t) (w[0] >> 16));
    }
};
template <> struct Unit<T_Q5_0> : UnitBlock32<T_Q5_0, 22> {
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A b = row + u * 22;
        uint32_t        w[6];
        ldw_n<6>(b, w);  // [d | qh0 qh1] [qh2 qh3 | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
        const float    d  = half_bits_to_float((uint16_t) (w[0] & 0xffff));
        const uint32_t qh = funnel_r(w[0], w[1], 16);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = funnel_r(w[1 + i], w[2 + i], 16);
            wq[i]            = (int) ((q & 0x0F0F0F0Fu) | (spread4(qh >> (4 * i)) << 4));
            wq[4 + i]        = (int) (((q >> 4) & 0x0F0F0F0Fu) | (spread4(qh >> (16 + 4 * i)) << 4));
        }
        ws[0] = d;
        wm[0] = -16.0f * d;
    }
};
template <> struct Unit<T_Q5_1> : UnitBlock32<T_Q5_1, 24> {
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A b = row + u * 24;
        uint32_t        w[6];
        ldw_n<6>(b, w);  // [d m] [qh] [qs x4]
        const uint32_t qh = w[1];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            wq[i]     = (int) ((w[2 + i] & 0x0F0F0F0Fu) | (spread4(qh >> (4 * i)) << 4));
            wq[4 + i] = (int) (((w[2 + i] >> 4) & 0x0F0F0F0Fu) | (spread4(qh >> (16 + 4 * i)) << 4));
        }
        ws[0] = half_bits_to_float((uint16_t) (w[0] & 0xffff));
        wm[0] = half_bits_to_float((uint16_t) (w[0] >> 16));
    }
};
template <> struct Unit<T_Q8_0> : UnitBlock32<T_Q8_0, 34> {
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A b = row + u * 34;
        uint32_t        w[9];
        ldw_n<9>(b, w);  // [d | q0 q1] ... [q30 q31 | next]
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            wq[i] = (int) funnel_r(w[i], w[i + 1], 16);
        }
        ws[0] = half_bits_to_float((uint16_t) (w[0] & 0xffff));
        wm[0] = 0.0f;
    }
};
template <> struct Unit<T_MXFP4> : UnitBlock32<T_MXFP4, 17> {
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A b = row + u * 17;
        uint32_t        w[5];
        ldw_n<5>(b, w);  // [e | qs0 qs1 qs2] ... [qs15 | next]
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = funnel_r(w[i], w[i + 1], 8);
            uint32_t       lo, hi;
            mxfp4_lut(q, lo, hi);
            wq[i]     = (int) lo;
            wq[4 + i] = (int) hi;
        }
        ws[0] = e8m0_half_to_float((uint8_t) (w[0] & 0xff));
        wm[0] = 0.0f;
    }
};

// scale/min bytes j (0..7) of a K-quant super-block from the three scale words (get_scale_min_k4 in word form)
B200Q_HD void k4_scale_min_words(const uint32_t (&s)[3], int j, int & sc, int & mn) {
    // bytes: s[0] = q[0..3], s[1] = q[4..7], s[2] = q[8..11]
    if (j < 4) {
        sc = (int) ((s[0] >> (8 * j)) & 63);
        mn = (int) ((s[1] >> (8 * j)) & 63);
    } else {
        const int      jj = j - 4;
        const uint32_t hi = (s[2] >> (8 * jj)) & 0xff;  // q[j+4]
        sc                = (int) ((hi & 0xF) | (((s[0] >> (8 * jj + 6)) & 3) << 4));
        mn                = (int) ((hi >> 4) | (((s[1] >> (8 * jj + 6)) & 3) << 4));
    }
}

// (scale, min) pairs 2p and 2p+1 of a K-quant super-block as packed 16-bit pairs (byte 0: pair 2p, byte 1: pair 2p+1)
B200Q_HD void k4_scale_min_pair(uint32_t s0, uint32_t s1, uint32_t s2, int p, uint32_t & scp, uint32_t & mnp) {
    const int      sh = 16 * (p & 1);
    const uint32_t a0 = s0 >> sh, a1 = s1 >> sh, a2 = s2 >> sh;
    if (p < 2) {
        scp = a0 & 0x3F3Fu;
        mnp = a1 & 0x3F3Fu;
    } else {
        scp = (a2 & 0x0F0Fu) | ((a0 >> 2) & 0x3030u);
        mnp = ((a2 >> 4) & 0x0F0Fu) | ((a1 >> 2) & 0x3030u);
    }
}

template <> struct Unit<T_Q4_K> {
    static constexpr int  ELEMS = 64, NG = 2, GW = 8;
    static constexpr bool UNSIGNED_W = true, MIN16 = false;  // wq bytes are unsigned; group 1 holds 16 * nibble (scale pre-divided)
    static B200Q_HD int  act_word(int u, int g) { return u * 16 + g * 8; }
    static B200Q_HD int  act_blk(int u, int g) { return u * 2 + g; }
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[16], float (&ws)[2], float (&wm)[2]) {
        const int p = u & 3;
        const A   b = row + (u >> 2) * 144;
        uint32_t  h[4];
        ldw_n_fast<4>(b, h);  // [d dmin] [scales x3]
        uint32_t q[8];
        ldw_n_fast<8>(b + 16 + 32 * p, q);
        const float d    = half_bits_to_float((uint16_t) (h[0] & 0xffff));
        const float dmin = half_bits_to_float((uint16_t) (h[0] >> 16));
        uint32_t    scp, mnp;
        k4_scale_min_pair(h[1], h[2], h[3], p, scp, mnp);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            wq[i]     = (int) (q[i] & 0x0F0F0F0Fu);
            wq[8 + i] = (int) (q[i] & 0xF0F0F0F0u);  // 16 * high nibble, consumed by the unsigned dp4a
        }
        ws[0] = d * (float) (scp & 0xff);
        ws[1] = d * (float) (scp >> 8) * 0.0625f;
        wm[0] = -dmin * (float) (mnp & 0xff);
        wm[1] = -dmin * (float) (mnp >> 8);
    }
};
template <> struct Unit<T_Q5_K> {
    static constexpr int  ELEMS = 64, NG = 2, GW = 8;
    static constexpr bool UNSIGNED_W = true, MIN16 = false;
    static B200Q_HD int  act_word(int u, int g) { return u * 16 + g * 8; }
    static B200Q_HD int  act_blk(int u, int g) { return u * 2 + g; }
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[16], float (&ws)[2], float (&wm)[2]) {
        const int p = u & 3;
        const A   b = row + (u >> 2) * 176;
        uint32_t  h[4];
        ldw_n_fast<4>(b, h);
        uint32_t qh[8];
        ldw_n_fast<8>(b + 16, qh);
        uint32_t q[8];
        ldw_n_fast<8>(b + 48 + 32 * p, q);
        const float d    = half_bits_to_float((uint16_t) (h[0] & 0xffff));
        const float dmin = half_bits_to_float((uint16_t) (h[0] >> 16));
        uint32_t    scp, mnp;
        k4_scale_min_pair(h[1], h[2], h[3], p, scp, mnp);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t hb = qh[i] >> (2 * p);
            wq[i]             = (int) ((q[i] & 0x0F0F0F0Fu) | ((hb & 0x01010101u) << 4));
            wq[8 + i]         = (int) (((q[i] >> 4) & 0x0F0F0F0Fu) | ((hb & 0x02020202u) << 3));
        }
        ws[0] = d * (float) (scp & 0xff);
        ws[1] = d * (float) (scp >> 8);
        wm[0] = -dmin * (float) (mnp & 0xff);
        wm[1] = -dmin * (float) (mnp >> 8);
    }
};
template <> struct Unit<T_Q6_K> {
    // unit u = (super-block sb, half n, 16-lane half h): elements 256 sb + 128 n + 32 quad + 16 h + [0,16), quad = 0..3
    static constexpr int  ELEMS = 64, NG = 4, GW = 4;
    static constexpr bool UNSIGNED_W = false, MIN16 = false;
    static B200Q_HD int  elem0(int u, int g) { return (u >> 2) * 256 + ((u >> 1) & 1) * 128 + g * 32 + (u & 1) * 16; }
    static B200Q_HD int  act_word(int u, int g) { return elem0(u, g) >> 2; }
    static B200Q_HD int  act_blk(int u, int g) { return elem0(u, g) >> 5; }
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[16], float (&ws)[4], float (&wm)[4]) {
        const int       n = (u >> 1) & 1, h = u & 1;
        const A b = row + (u >> 2) * 210;
        uint32_t        l0[4], l1[4], qh[4];
        ldw_n<4>(b + 64 * n + 16 * h, l0);        // ql[l],      l = 16h .. 16h+15
        ldw_n<4>(b + 64 * n + 32 + 16 * h, l1);   // ql[l + 32]
        ldw_n<4>(b + 128 + 32 * n + 16 * h, qh);  // qh[l]
        // scales: sc[8n + 2*quad + h], d at +208
        const uint32_t sw0 = ldw(b + 192 + 8 * n);      // sc[8n+0..3]
        const uint32_t sw1 = ldw(b + 192 + 8 * n + 4);  // sc[8n+4..7]
        const float    d   = half_bits_to_float((uint16_t) (ldw(b + 208) & 0xffff));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q0 = (l0[i] & 0x0F0F0F0Fu) | (((qh[i] >> 0) & 0x03030303u) << 4);
            const uint32_t q1 = (l1[i] & 0x0F0F0F0Fu) | (((qh[i] >> 2) & 0x03030303u) << 4);
            const uint32_t q2 = ((l0[i] >> 4) & 0x0F0F0F0Fu) | (((qh[i] >> 4) & 0x03030303u) << 4);
            const uint32_t q3 = ((l1[i] >> 4) & 0x0F0F0F0Fu) | (((qh[i] >> 6) & 0x03030303u) << 4);
            wq[0 + i]         = (int) vsub4_(q0, 0x20202020u);
            wq[4 + i]         = (int) vsub4_(q1, 0x20202020u);
            wq[8 + i]         = (int) vsub4_(q2, 0x20202020u);
            wq[12 + i]        = (int) vsub4_(q3, 0x20202020u);
        }
        ws[0] = d * (float) (int8_t) (sw0 >> (8 * h));
        ws[1] = d * (float) (int8_t) (sw0 >> (16 + 8 * h));
        ws[2] = d * (float) (int8_t) (sw1 >> (8 * h));
        ws[3] = d * (float) (int8_t) (sw1 >> (16 + 8 * h));
        wm[0] = wm[1] = wm[2] = wm[3] = 0.0f;
    }
};

// 16-entry int8 table look-up on the 8 nibbles of one word (PRMT pairs, as mxfp4_lut): kvalues_iq4nl
B200Q_HD void iq4nl_lut(uint32_t q, uint32_t & lo, uint32_t & hi) {
    const uint32_t t0 = 0xBFAD9881u, t1 = 0xF6EADDCFu, t2 = 0x26190D01u, t3 = 0x71594535u;
    uint32_t       r[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t n   = (q >> (4 * h)) & 0x0F0F0F0Fu;
        const uint32_t sel = n & 0x07070707u;
        const uint32_t s   = (sel & 0xF) | ((sel >> 4) & 0xF0) | ((sel >> 8) & 0xF00) | ((sel >> 12) & 0xF000);
        const uint32_t a   = byte_perm_(t0, t1, s);  // entries 0..7
        const uint32_t b   = byte_perm_(t2, t3, s);  // entries 8..15
        const uint32_t m   = ((n >> 3) & 0x01010101u) * 0xFFu;
        r[h]               = (a & ~m) | (b & m);
    }
    lo = r[0];
    hi = r[1];
}
template <> struct Unit<T_IQ4_NL> : UnitBlock32<T_IQ4_NL, 18> {  // vec_dot_iq4_nl_q8_1, vecdotq.cuh:1124-1146
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const A  b = row + u * 18;
        uint32_t w[5];
        ldw_n<5>(b, w);  // [d | qs0 qs1] [qs2..5] [qs6..9] [qs10..13] [qs14 qs15 | next]
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            uint32_t lo, hi;
            iq4nl_lut(funnel_r(w[i], w[i + 1], 16), lo, hi);
            wq[i]     = (int) lo;
            wq[4 + i] = (int) hi;
        }
        ws[0] = half_bits_to_float((uint16_t) (w[0] & 0xffff));
        wm[0] = 0.0f;
    }
};
template <> struct Unit<T_IQ4_XS> : UnitBlock32<T_IQ4_XS, 0> {  // vec_dot_iq4_xs_q8_1, vecdotq.cuh:1148-1171
    // unit u = sub-block ib = u & 7 of super-block u >> 3 (136 bytes: d, scales_h, scales_l[4], qs[128])
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[8], float (&ws)[1], float (&wm)[1]) {
        const int ib = u & 7;
        const A   b  = row + (u >> 3) * 136;
        uint32_t  h[2];
        ldw_n<2>(b, h);  // [d | scales_h] [scales_l x4]
        uint32_t q[4];
        ldw_n<4>(b + 8 + 16 * ib, q);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            uint32_t lo, hi;
            iq4nl_lut(q[i], lo, hi);
            wq[i]     = (int) lo;
            wq[4 + i] = (int) hi;
        }
        const int ls = (int) (((h[1] >> (4 * ib)) & 0xF) | ((((h[0] >> 16) >> (2 * ib)) & 3) << 4));
        ws[0] = half_bits_to_float((uint16_t) (h[0] & 0xffff)) * (float) (ls - 32);
        wm[0] = 0.0f;
    }
};
// q2_K / q3_K: unit u = (super-block, 128-element half n, 16-byte half h) like q6_K: 16 bytes of 2-bit fields are four groups
// (shift 2 g) of 16 elements 256 sb + 128 n + 32 g + 16 h + [0, 16), each with its own scale (vec_dot_q2_K_q8_1
// vecdotq.cuh:268-311, vec_dot_q3_K_q8_1 :313-404)
template <> struct Unit<T_Q2_K> {
    static constexpr int  ELEMS = 64, NG = 4, GW = 4;
    static constexpr bool UNSIGNED_W = true, MIN16 = true;  // the min of a group multiplies the sum of ITS 16 activations
    static B200Q_HD int  elem0(int u, int g) { return (u >> 2) * 256 + ((u >> 1) & 1) * 128 + g * 32 + (u & 1) * 16; }
    static B200Q_HD int  act_word(int u, int g) { return elem0(u, g) >> 2; }
    static B200Q_HD int  act_blk(int u, int g) { return elem0(u, g) >> 5; }
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[16], float (&ws)[4], float (&wm)[4]) {
        const int n = (u >> 1) & 1, h = u & 1;
        const A   b = row + (u >> 2) * 84;
        uint32_t  q[4];
        ldw_n<4>(b + 16 + 32 * n + 16 * h, q);
        const uint32_t s0 = ldw(b + 8 * n), s1 = ldw(b + 8 * n + 4);  // scales[8n + 0..3], [8n + 4..7]
        const uint32_t dm = ldw(b + 80);
        const float    d = half_bits_to_float((uint16_t) (dm & 0xffff)), dmin = half_bits_to_float((uint16_t) (dm >> 16));
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
            for (int i = 0; i < 4; ++i) wq[4 * g + i] = (int) ((q[i] >> (2 * g)) & 0x03030303u);
            const uint32_t sc = ((g < 2 ? s0 : s1) >> (16 * (g & 1) + 8 * h)) & 0xff;  // scales[8n + 2g + h]
            ws[g] = d * (float) (sc & 0xF);
            wm[g] = -dmin * (float) (sc >> 4);
        }
    }
};
template <> struct Unit<T_Q3_K> {
    static constexpr int  ELEMS = 64, NG = 4, GW = 4;
    static constexpr bool UNSIGNED_W = false, MIN16 = false;
    static B200Q_HD int  elem0(int u, int g) { return (u >> 2) * 256 + ((u >> 1) & 1) * 128 + g * 32 + (u & 1) * 16; }
    static B200Q_HD int  act_word(int u, int g) { return elem0(u, g) >> 2; }
    static B200Q_HD int  act_blk(int u, int g) { return elem0(u, g) >> 5; }
    template <class A> static B200Q_HD void decode(A row, int u, int (&wq)[16], float (&ws)[4], float (&wm)[4]) {
        const int n = (u >> 1) & 1, h = u & 1;
        const A   b = row + (u >> 2) * 110;
        uint32_t  q[4], hm[4], sc[3];
        ldw_n<4>(b + 32 + 32 * n + 16 * h, q);
        ldw_n<4>(b + 16 * h, hm);
        ldw_n<3>(b + 96, sc);  // 12 packed scale bytes
        const float d = half_bits_to_float((uint16_t) (ldw(b + 108) & 0xffff));
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t v = ((q[i] >> (2 * g)) & 0x03030303u) | (((hm[i] >> (4 * n + g)) & 0x01010101u) << 2);
                wq[4 * g + i]    = (int) vsub4_(v, 0x04040404u);  // (q & 3) - (high bit set ? 0 : 4)
            }
            const int      j  = 8 * n + 2 * g + h;                // scale index
            const uint32_t lw = sc[(j & 7) >> 2] >> (8 * (j & 3));
            const int      lo = (int) (j < 8 ? (lw & 0xF) : ((lw >> 4) & 0xF));
            const int      hi = (int) ((sc[2] >> (8 * (j & 3) + 2 * (j >> 2))) & 3);
            ws[g] = d * (float) ((lo | (hi << 4)) - 32);
            wm[g] = 0.0f;
        }
    }
};

// acc += sum over groups of  ws*d_x*dot + wm*s_x   against one activation column held as
// int8x4 words `qa` and (d, s) pairs `ds`
template <int T> B200Q_HD float unit_dot(const int (&wq)[Unit<T>::NG * Unit<T>::GW], const float (&ws)[Unit<T>::NG],
                                         const float (&wm)[Unit<T>::NG], int u, const int * qa, const float2 * ds) {
    using U   = Unit<T>;
    float acc = 0.0f;
#pragma unroll
    for (int g = 0; g < U::NG; ++g) {
        const int * a    = qa + U::act_word(u, g);
        int         sumi = 0;
#pragma unroll
        for (int i = 0; i < U::GW; ++i) {
            sumi = U::UNSIGNED_W ? dp4a_us((uint32_t) wq[g * U::GW + i], a[i], sumi) : dp4a_(wq[g * U::GW + i], a[i], sumi);
        }
        float2 x = ds[U::act_blk(u, g)];
        if (U::MIN16) {
            int asum = 0;
#pragma unroll
            for (int i = 0; i < U::GW; ++i) asum = dp4a_(0x01010101, a[i], asum);
            x.y = x.x * (float) asum;
        }
        acc += ws[g] * (x.x * (float) sumi) + wm[g] * x.y;
    }
    return acc;
}

}  // namespace b200q
