// ops.cu — the operators that bracket the quantized matmuls in a decoder layer, so that a layer's graph does not bounce to
// the CPU backend between them: RMS_NORM, ADD / SUB / MUL / DIV with ggml broadcasting, SCALE, the unary
// activations, GLU (SWIGLU / GEGLU / REGLU / SWIGLU_OAI ...), SOFT_MAX (mask, ALiBi slopes, sinks), ROPE (normal / NeoX,
// frequency factors, YaRN), CPY / CONT / DUP between f32 / f16 / bf16, SET_ROWS (the KV-cache write).
//
// Each kernel restates the arithmetic of the reference's CPU operator (ggml/src/ggml-cpu/ops.cpp: rms_norm :4138-4187,
// soft_max :5218-5330, rope :6049-6330, glu :3285-3700, set_rows :5342-5400; binary-ops.cpp; unary-ops.cpp; vec.h
// :601-700 for the activation formulas) and replaces the reference's norm.cu / binbcast.cu / scale.cu / unary.cu /
// softmax.cu / rope.cu / cpy.cu / set-rows.cu.  All of it is HBM-bound elementwise or row work: coalesced accesses,
// one CTA per row for the reductions, grid-stride loops sized in multiples of the SM count elsewhere; nothing here is
// shaped into a GEMM.  Gate: NMSE <= 1e-7 against ggml-cpu (test-backend-ops' default, tests/test-backend-ops.cpp:948-950).
#include "../../include/b200q.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "pdl.cuh"
#include <cuda_runtime.h>

#include <math.h>
#include <stdint.h>

namespace b200q {
int ops_fail(int code, const char * msg);  // api.cu: records the message for b200q_last_error
}
using b200q::ops_fail;
using b200q::launch_pdl;
using b200q::pdl_grid_sync;

namespace {

struct T4 {  // a tensor as a kernel sees it
    char *    data;
    long long ne[4];
    long long nb[4];
};
T4 t4(const b200q_tensor * t) {
    T4 r;
    r.data = (char *) t->data;
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = (long long) t->nb[i]; }
    return r;
}
long long nelem(const b200q_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
long long nrows(const b200q_tensor * t) { return t->ne[1] * t->ne[2] * t->ne[3]; }
size_t tsize(int type) { return type == B200Q_TYPE_F32 ? 4 : (type == B200Q_TYPE_F16 || type == B200Q_TYPE_BF16) ? 2 : type == B200Q_TYPE_I32 ? 4 : type == B200Q_TYPE_I64 ? 8 : 0; }
bool contiguous(const b200q_tensor * t) {
    size_t s = tsize(t->type);
    if (!s) return false;
    for (int i = 0; i < 4; ++i) {
        if (t->ne[i] != 1 && t->nb[i] != s) return false;
        s *= (size_t) t->ne[i];
    }
    return true;
}
bool contiguous_rows(const b200q_tensor * t) {  // ggml_is_contiguous_1: dims >= 1 dense, rows at any pitch
    size_t s = t->nb[1];
    if (t->nb[0] != tsize(t->type)) return false;
    for (int i = 1; i < 4; ++i) {
        if (t->ne[i] != 1 && t->nb[i] != s) return false;
        s *= (size_t) t->ne[i];
    }
    return true;
}
bool same_shape(const b200q_tensor * a, const b200q_tensor * b) {
    for (int i = 0; i < 4; ++i) if (a->ne[i] != b->ne[i]) return false;
    return true;
}
int grid_for(long long n, int threads) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long want = (n + threads - 1) / threads;
    const long long cap  = (long long) sms * 16;
    return (int) (want < 1 ? 1 : want < cap ? want : cap);
}
int launched() {
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : ops_fail(B200Q_ECUDA, cudaGetErrorString(e));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <bool MAX> __device__ __forceinline__ float block_reduce(float v, float * sh) {
    v = MAX ? warp_max(v) : warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
    __syncthreads();  // sh may still be read from a previous reduction
    if (l == 0) sh[w] = v;
    __syncthreads();
    v = l < nw ? sh[l] : (MAX ? -INFINITY : 0.0f);
    return MAX ? warp_max(v) : warp_sum(v);
}

// ---- element load / store by type -----------------------------------------------------------------------------------------
template <int TYPE> struct Elem;
template <> struct Elem<B200Q_TYPE_F32> {
    static __device__ __forceinline__ float ld(const char * p) { return *(const float *) p; }
    static __device__ __forceinline__ void  st(char * p, float v) { *(float *) p = v; }
};
template <> struct Elem<B200Q_TYPE_F16> {
    static __device__ __forceinline__ float ld(const char * p) { return __half2float(*(const __half *) p); }
    static __device__ __forceinline__ void  st(char * p, float v) { *(__half *) p = __float2half_rn(v); }
};
template <> struct Elem<B200Q_TYPE_BF16> {
    static __device__ __forceinline__ float ld(const char * p) { return __bfloat162float(*(const __nv_bfloat16 *) p); }
    static __device__ __forceinline__ void  st(char * p, float v) { *(__nv_bfloat16 *) p = __float2bfloat16_rn(v); }
};

// ---- RMS_NORM: y = x / sqrt(mean(x^2) + eps), one CTA per row -------------------------------------------------------------
__global__ void __launch_bounds__(256) rms_norm_kernel(T4 x, T4 y, float eps) {
    pdl_grid_sync();
    __shared__ float sh[32];
    const long long row = blockIdx.x;
    const long long i3 = row / (x.ne[1] * x.ne[2]), i2 = (row / x.ne[1]) % x.ne[2], i1 = row % x.ne[1];
    const float * xp = (const float *) (x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *       yp = (float *) (y.data + i1 * y.nb[1] + i2 * y.nb[2] + i3 * y.nb[3]);
    const int     n  = (int) x.ne[0];
    float s = 0.0f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = xp[i];
        s += v * v;
    }
    s = block_reduce<false>(s, sh);
    const float scale = 1.0f / sqrtf(s / (float) n + eps);
    for (int i = threadIdx.x; i < n; i += blockDim.x) yp[i] = xp[i] * scale;
}

// ---- [ADD +] RMS_NORM [+ MUL] in one launch: x = a (+ b, optionally written out), y = (x / sqrt(mean(x^2) + eps)) (* w) ----------
// The three operators are consecutive nodes twice per decoder layer (residual add -> norm -> norm weight); a decode token pays
// ~3.5 us per launch, and the stand-alone norm re-read its row through possibly-aliasing pointers.  One CTA per row, the row
// lives in registers (n <= 8 * blockDim.x), same arithmetic as the separate kernels (sum, then x * scale, then * w: no FMA
// across the operator boundaries).  w is broadcast over rows as ggml's MUL does (index modulo its extents).
template <bool ADD, bool MUL> __global__ void __launch_bounds__(1024) norm_fused_kernel(T4 a, T4 b, T4 so, T4 w, T4 y, float eps, int write_sum, int vec4) {
    pdl_grid_sync();
    __shared__ float sh[32];
    const long long row = blockIdx.x;
    const long long i3 = row / (a.ne[1] * a.ne[2]), i2 = (row / a.ne[1]) % a.ne[2], i1 = row % a.ne[1];
    const float * __restrict__ ap = (const float *) (a.data + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    const float * __restrict__ bp = ADD ? (const float *) (b.data + i1 * b.nb[1] + i2 * b.nb[2] + i3 * b.nb[3]) : nullptr;
    float *       sp = ADD && write_sum ? (float *) (so.data + i1 * so.nb[1] + i2 * so.nb[2] + i3 * so.nb[3]) : nullptr;
    const float * __restrict__ wp = MUL ? (const float *) (w.data + (i1 % w.ne[1]) * w.nb[1] + (i2 % w.ne[2]) * w.nb[2] + (i3 % w.ne[3]) * w.nb[3]) : nullptr;
    float *       yp = (float *) (y.data + i1 * y.nb[1] + i2 * y.nb[2] + i3 * y.nb[3]);
    const int     n  = (int) a.ne[0];
    float v[8], wv[8];
    float s = 0.0f;
    if (vec4) {
        // 16-byte accesses, two per thread and operand (n <= 8 * blockDim.x, n % 4 == 0, rows 16-byte aligned); the norm weight
        // is requested BEFORE the reduction so that its latency hides behind it (the decode step runs ONE row per launch)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int i = 4 * (threadIdx.x + j * blockDim.x);
            float4 x = make_float4(0.0f, 0.0f, 0.0f, 0.0f), ww = x;
            if (i < n) {
                x = *(const float4 *) (ap + i);
                if (ADD) {
                    const float4 bb = *(const float4 *) (bp + i);
                    x.x = x.x + bb.x; x.y = x.y + bb.y; x.z = x.z + bb.z; x.w = x.w + bb.w;
                    if (sp) *(float4 *) (sp + i) = x;
                }
                if (MUL) ww = *(const float4 *) (wp + i);
            }
            v[4 * j] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
            wv[4 * j] = ww.x; wv[4 * j + 1] = ww.y; wv[4 * j + 2] = ww.z; wv[4 * j + 3] = ww.w;
            s += x.x * x.x; s += x.y * x.y; s += x.z * x.z; s += x.w * x.w;
        }
        s = block_reduce<false>(s, sh);
        const float scale = 1.0f / sqrtf(s / (float) n + eps);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int i = 4 * (threadIdx.x + j * blockDim.x);
            if (i < n) {
                float4 o;
                o.x = v[4 * j] * scale; o.y = v[4 * j + 1] * scale; o.z = v[4 * j + 2] * scale; o.w = v[4 * j + 3] * scale;
                if (MUL) { o.x = o.x * wv[4 * j]; o.y = o.y * wv[4 * j + 1]; o.z = o.z * wv[4 * j + 2]; o.w = o.w * wv[4 * j + 3]; }
                *(float4 *) (yp + i) = o;
            }
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int i = threadIdx.x + j * blockDim.x;
        float x = 0.0f;
        wv[j] = 0.0f;
        if (i < n) {
            x = ap[i];
            if (ADD) {
                x = x + bp[i];
                if (sp) sp[i] = x;
            }
            if (MUL) wv[j] = wp[i];
        }
        v[j] = x;
        s += x * x;
    }
    s = block_reduce<false>(s, sh);
    const float scale = 1.0f / sqrtf(s / (float) n + eps);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int i = threadIdx.x + j * blockDim.x;
        if (i < n) {
            const float t = v[j] * scale;
            yp[i] = MUL ? t * wv[j] : t;
        }
    }
}

// ---- binary ops with ggml broadcasting: dst = src0 (op) src1, src1 repeated along every dim ---------------------------------
template <int OP> __device__ __forceinline__ float bin_apply(float a, float b) {
    return OP == B200Q_BIN_ADD ? a + b : OP == B200Q_BIN_SUB ? a - b : OP == B200Q_BIN_MUL ? a * b : a / b;
}
template <int OP> __global__ void __launch_bounds__(256) bin_kernel(T4 a, T4 b, T4 d, long long n) {
    pdl_grid_sync();
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const long long i0 = i % d.ne[0], r = i / d.ne[0];
        const long long i1 = r % d.ne[1], r2 = r / d.ne[1];
        const long long i2 = r2 % d.ne[2], i3 = r2 / d.ne[2];
        const float av = *(const float *) (a.data + i0 * a.nb[0] + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
        const float bv = *(const float *) (b.data + (i0 % b.ne[0]) * b.nb[0] + (i1 % b.ne[1]) * b.nb[1] + (i2 % b.ne[2]) * b.nb[2] + (i3 % b.ne[3]) * b.nb[3]);
        *(float *) (d.data + i0 * d.nb[0] + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = bin_apply<OP>(av, bv);
    }
}

// ---- SCALE and the unary activations (contiguous f32) ---------------------------------------------------------------------
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x))); }
__device__ __forceinline__ float gelu_quick_f(float x) { return x * (1.0f / (1.0f + expf(-1.702f * x))); }
__device__ __forceinline__ float gelu_erf_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float unary_apply(int op, float x) {
    switch (op) {
        case B200Q_UNARY_SILU:       return silu_f(x);
        case B200Q_UNARY_GELU:       return gelu_f(x);
        case B200Q_UNARY_GELU_QUICK: return gelu_quick_f(x);
        case B200Q_UNARY_GELU_ERF:   return gelu_erf_f(x);
        case B200Q_UNARY_RELU:       return fmaxf(x, 0.0f);
        case B200Q_UNARY_SIGMOID:    return 1.0f / (1.0f + expf(-x));
        case B200Q_UNARY_TANH:       return tanhf(x);
        case B200Q_UNARY_NEG:        return -x;
        case B200Q_UNARY_ABS:        return fabsf(x);
        case B200Q_UNARY_EXP:        return expf(x);
        default:                     return x;
    }
}
__global__ void __launch_bounds__(256) unary_kernel(const float * x, float * y, long long n, int op) {
    pdl_grid_sync();
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) y[i] = unary_apply(op, x[i]);
}
__global__ void __launch_bounds__(256) scale_kernel(const float * x, float * y, long long n, float s, float b) {
    pdl_grid_sync();
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) y[i] = x[i] * s + b;
}

// ---- GLU: dst[r][c] = act(x[r][c]) * g[r][c]; x and g are the two halves of one row or two tensors ------------------------------
__global__ void __launch_bounds__(256) glu_kernel(const char * x, const char * g, char * d, long long nr, int nc, long long x_nb1,
                                                  long long g_nb1, long long d_nb1, int op, float alpha, float limit) {
    pdl_grid_sync();
    auto apply = [&](float xv, float gv) {
        switch (op) {
            case B200Q_GLU_REGLU:       return fmaxf(xv, 0.0f) * gv;
            case B200Q_GLU_GEGLU:       return gelu_f(xv) * gv;
            case B200Q_GLU_GEGLU_ERF:   return gelu_erf_f(xv) * gv;
            case B200Q_GLU_GEGLU_QUICK: return gelu_quick_f(xv) * gv;
            case B200Q_GLU_SWIGLU_OAI: {
                const float xc = fminf(xv, limit), yc = fminf(fmaxf(gv, -limit), limit);
                return (xc / (1.0f + expf(alpha * (-xc)))) * (yc + 1.0f);
            }
            default:                    return silu_f(xv) * gv;  // SWIGLU
        }
    };
    const bool vec = (nc & 3) == 0 && ((((uintptr_t) x | (uintptr_t) g | (uintptr_t) d) | (unsigned long long) (x_nb1 | g_nb1 | d_nb1)) & 15) == 0;
    if (vec) {  // four values per thread, 16-byte accesses (prompt batches: 512 x 14336 values per layer)
        const int       nc4 = nc >> 2;
        const long long n4  = nr * nc4;
        for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long) gridDim.x * blockDim.x) {
            const long long r = i / nc4;
            const int       c = (int) (i - r * nc4);
            const float4 xv = ((const float4 *) (x + r * x_nb1))[c], gv = ((const float4 *) (g + r * g_nb1))[c];
            float4 o;
            o.x = apply(xv.x, gv.x); o.y = apply(xv.y, gv.y); o.z = apply(xv.z, gv.z); o.w = apply(xv.w, gv.w);
            ((float4 *) (d + r * d_nb1))[c] = o;
        }
        return;
    }
    const long long n = nr * nc;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const long long r = i / nc;
        const int       c = (int) (i - r * nc);
        const float xv = ((const float *) (x + r * x_nb1))[c], gv = ((const float *) (g + r * g_nb1))[c];
        ((float *) (d + r * d_nb1))[c] = apply(xv, gv);
    }
}

// ---- SOFT_MAX over rows: softmax(x * scale + slope(head) * mask [, sink]), one CTA per row ----------------------------------
template <bool MASK_F16> __global__ void __launch_bounds__(256) soft_max_kernel(T4 x, T4 m, const float * sinks, T4 d, float scale, float max_bias,
                                                                                float m0, float m1, unsigned n_head_log2) {
    pdl_grid_sync();
    __shared__ float sh[32];
    const long long row = blockIdx.x;
    const long long i3 = row / (x.ne[1] * x.ne[2]), i2 = (row / x.ne[1]) % x.ne[2], i1 = row % x.ne[1];
    const float * xp = (const float *) (x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *       dp = (float *) (d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const char *  mp = m.data ? m.data + i1 * m.nb[1] + (i2 % m.ne[2]) * m.nb[2] + (i3 % m.ne[3]) * m.nb[3] : nullptr;
    const int     n  = (int) x.ne[0];
    const unsigned h = (unsigned) i2;
    const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float) (h + 1)) : powf(m1, (float) (2 * (h - n_head_log2) + 1))) : 1.0f;
    auto val = [&](int i) {
        float v = xp[i] * scale;
        if (mp) v += slope * (MASK_F16 ? __half2float(((const __half *) mp)[i]) : ((const float *) mp)[i]);
        return v;
    };
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < n; i += blockDim.x) mx = fmaxf(mx, val(i));
    mx = block_reduce<true>(mx, sh);
    if (sinks) mx = fmaxf(mx, sinks[i2]);
    float s = 0.0f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float e = expf(val(i) - mx);
        dp[i] = e;
        s += e;
    }
    s = block_reduce<false>(s, sh);
    if (sinks) s += expf(sinks[i2] - mx);
    const float inv = 1.0f / s;
    for (int i = threadIdx.x; i < n; i += blockDim.x) dp[i] *= inv;
}

// rows of at most 1024 values (attention over a short context, MoE routers): one WARP per row, the row in registers, no block
// barriers — at pp512 the 16384 rows of a layer's K^T Q took 61 us with a CTA each (1.1 TB/s)
template <bool MASK_F16> __global__ void __launch_bounds__(256) soft_max_warp_kernel(T4 x, T4 m, const float * sinks, T4 d, float scale, float max_bias,
                                                                                     float m0, float m1, unsigned n_head_log2, long long nrows) {
    pdl_grid_sync();
    const int       lane = threadIdx.x & 31;
    const long long row  = (long long) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= nrows) return;
    const long long i3 = row / (x.ne[1] * x.ne[2]), i2 = (row / x.ne[1]) % x.ne[2], i1 = row % x.ne[1];
    const float * xp = (const float *) (x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *       dp = (float *) (d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const char *  mp = m.data ? m.data + i1 * m.nb[1] + (i2 % m.ne[2]) * m.nb[2] + (i3 % m.ne[3]) * m.nb[3] : nullptr;
    const int     n  = (int) x.ne[0];
    const unsigned h = (unsigned) i2;
    const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float) (h + 1)) : powf(m1, (float) (2 * (h - n_head_log2) + 1))) : 1.0f;
    float v[32];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int i = lane + 32 * j;
        float t = -INFINITY;
        if (i < n) {
            t = xp[i] * scale;
            if (mp) t += slope * (MASK_F16 ? __half2float(((const __half *) mp)[i]) : ((const float *) mp)[i]);
        }
        v[j] = t;
        mx   = fmaxf(mx, t);
    }
    mx = warp_max(mx);
    if (sinks) mx = fmaxf(mx, sinks[i2]);
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        if (lane + 32 * j < n) {
            v[j] = expf(v[j] - mx);
            s += v[j];
        }
    }
    s = warp_sum(s);
    if (sinks) s += expf(sinks[i2] - mx);
    const float inv = 1.0f / s;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int i = lane + 32 * j;
        if (i < n) dp[i] = v[j] * inv;
    }
}

// ---- decode attention core in one launch: softmax(scale * K^T q + mask [, sink]) applied to V -------------------------------------
// llama.cpp without flash attention emits MUL_MAT(k, q) -> SOFT_MAX -> MUL_MAT(v, kq) -> PERMUTE -> CONT; at decode these are
// four ~3.5 us launches per layer around a few hundred KB of KV cache.  One CTA per (head, token): scores of all cached
// positions into shared memory (a warp per position, K rows are read 64 bytes per warp-load), the softmax in place, then a
// warp per output element over the transposed V rows.  f32 arithmetic throughout, same operations as the separate kernels
// (different summation order).  q [hd, n_tok, n_head] f32; k [hd, n_kv, n_head_kv] f16; v [n_kv, hd, n_head_kv] f16 (the
// transposed V cache); mask [n_kv, >= n_tok, ...] f32 | f16 or none; dst [hd, n_head, n_tok] f32.
struct AttnP {
    T4            q, k, v, m, d;
    const float * sinks;
    float         scale;
    int           mask_f16;
};
__device__ __forceinline__ float dot8_h(const uint4 raw, const float * x) {
    const __half2 * h = (const __half2 *) &raw;
    float a = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = __half22float2(h[i]);
        a = fmaf(x[2 * i], f.x, a);
        a = fmaf(x[2 * i + 1], f.y, a);
    }
    return a;
}
__global__ void __launch_bounds__(256) attn_decode_kernel(const AttnP p) {
    pdl_grid_sync();
    extern __shared__ float at_smem[];  // [n_kv] scores | [hd] q | [256] partial outputs
    __shared__ float sh[32];
    const int h = blockIdx.x, t = blockIdx.y, tid = threadIdx.x;
    const int hd = (int) p.q.ne[0], n_kv = (int) p.k.ne[1];
    const int hk = h / (int) (p.q.ne[2] / p.k.ne[2]);  // ggml's batch broadcast: kv head = head / (n_head / n_head_kv)
    float * sc = at_smem;
    float * qs = at_smem + ((n_kv + 3) & ~3);
    float * pt = qs + ((hd + 3) & ~3);
    const float * qp = (const float *) (p.q.data + (long long) t * p.q.nb[1] + (long long) h * p.q.nb[2]);
    for (int e = tid; e < hd; e += 256) qs[e] = qp[e];
    __syncthreads();
    // scores: one THREAD per cached position, its K row (hd f16, contiguous) in 16-byte loads — up to hd / 8 independent loads
    // in flight per thread (a warp per position with a shuffle reduction was latency-bound: 22 us per layer)
    const char * kb = p.k.data + (long long) hk * p.k.nb[2];
    const char * mp = p.m.data ? p.m.data + (long long) t * p.m.nb[1] + (long long) (h % p.m.ne[2]) * p.m.nb[2] : nullptr;
    const bool kvec = (hd & 7) == 0 && ((((uintptr_t) p.k.data) | (unsigned long long) p.k.nb[1] | (unsigned long long) p.k.nb[2]) & 15) == 0;
    for (int j = tid; j < n_kv; j += 256) {
        const char * kr = kb + (long long) j * p.k.nb[1];
        float acc = 0.0f;
        if (kvec) {
#pragma unroll 4
            for (int e = 0; e < hd; e += 8) acc += dot8_h(__ldg((const uint4 *) (kr + 2 * e)), qs + e);
        } else {
            for (int e = 0; e < hd; ++e) acc = fmaf(qs[e], __half2float(((const __half *) kr)[e]), acc);
        }
        float x = acc * p.scale;
        if (mp) x += p.mask_f16 ? __half2float(((const __half *) mp)[j]) : ((const float *) mp)[j];
        sc[j] = x;
    }
    __syncthreads();
    float mx = -INFINITY;
    for (int j = tid; j < n_kv; j += 256) mx = fmaxf(mx, sc[j]);
    mx = block_reduce<true>(mx, sh);
    if (p.sinks) mx = fmaxf(mx, p.sinks[h]);
    float s = 0.0f;
    for (int j = tid; j < n_kv; j += 256) {
        const float e = expf(sc[j] - mx);
        sc[j] = e;
        s += e;
    }
    s = block_reduce<false>(s, sh);  // (its barriers also publish the exponentials)
    if (p.sinks) s += expf(p.sinks[h] - mx);
    const float inv = 1.0f / s;
    // output: thread (element e, part) walks its share of the transposed V row e (contiguous over positions, 16-byte loads);
    // the parts are added in a fixed order
    const char * vb = p.v.data + (long long) hk * p.v.nb[2];
    const int  parts = (hd <= 256 && 256 % hd == 0) ? 256 / hd : 1;
    const bool vvec  = ((((uintptr_t) p.v.data) | (unsigned long long) p.v.nb[1] | (unsigned long long) p.v.nb[2]) & 15) == 0;
    const int  chunk = (((n_kv + parts - 1) / parts) + 7) & ~7;
    float *    dp    = (float *) (p.d.data + (long long) h * p.d.nb[1] + (long long) t * p.d.nb[2]);
    for (int e0 = 0; e0 < hd; e0 += 256) {
        const int e = e0 + (parts > 1 ? tid % hd : tid), part = parts > 1 ? tid / hd : 0;
        float acc = 0.0f;
        if (e < hd) {
            const char * vr = vb + (long long) e * p.v.nb[1];
            const int j0 = part * chunk, j1 = min(n_kv, j0 + chunk);
            int j = j0;
            if (vvec) {
#pragma unroll 4
                for (; j + 8 <= j1; j += 8) acc += dot8_h(__ldg((const uint4 *) (vr + 2 * j)), sc + j);
            }
            for (; j < j1; ++j) acc = fmaf(sc[j], __half2float(((const __half *) vr)[j]), acc);
        }
        if (parts > 1) {
            pt[tid] = acc;
            __syncthreads();
            if (tid < hd) {
                float o = pt[tid];
                for (int q = 1; q < parts; ++q) o += pt[tid + q * hd];
                dp[tid] = o * inv;
            }
        } else if (e < hd) {
            dp[e] = acc * inv;
        }
    }
}

// ---- ROPE: one thread per rotated pair ----------------------------------------------------------------------------------------
struct RopeP {
    int   n_dims, neox;
    float theta_scale, freq_scale, ext_factor, attn_factor, corr0, corr1;
};
// rope_yarn (ops.cpp:6056-6071) on theta = pos * theta_scale^(i0/2) / freq_factor: cosine and sine of pair ip at position pos
__device__ __forceinline__ void rope_cos_sin(const RopeP & p, int pos, long long ip, const float * ff, float & c, float & s) {
    const float theta_extrap = (float) pos * powf(p.theta_scale, (float) ip) / (ff ? ff[ip] : 1.0f);
    const float theta_interp = p.freq_scale * theta_extrap;
    float theta = theta_interp, mscale = p.attn_factor;
    if (p.ext_factor != 0.0f) {
        const float y    = ((float) ip - p.corr0) / fmaxf(0.001f, p.corr1 - p.corr0);
        const float ramp = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * p.ext_factor;
        theta            = theta_interp * (1.0f - ramp) + theta_extrap * ramp;
        mscale *= 1.0f + 0.1f * logf(1.0f / p.freq_scale);
    }
    c = cosf(theta) * mscale;
    s = sinf(theta) * mscale;
}
template <int TYPE> __global__ void __launch_bounds__(256) rope_kernel(T4 x, const int * pos, const float * ff, T4 d, RopeP p) {
    pdl_grid_sync();
    const long long half = x.ne[0] / 2, n = half * x.ne[1] * x.ne[2] * x.ne[3];
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const long long ip = i % half, r = i / half;
        const long long i1 = r % x.ne[1], r2 = r / x.ne[1], i2 = r2 % x.ne[2], i3 = r2 / x.ne[2];
        const char * xs = x.data + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3];
        char *       ds = d.data + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3];
        const long long i0 = 2 * ip;  // pair index in the rotation order
        constexpr long long es = TYPE == B200Q_TYPE_F32 ? 4 : 2;
        if (i0 >= p.n_dims) {  // channels past n_dims are copied
            Elem<TYPE>::st(ds + i0 * es, Elem<TYPE>::ld(xs + i0 * es));
            Elem<TYPE>::st(ds + (i0 + 1) * es, Elem<TYPE>::ld(xs + (i0 + 1) * es));
            continue;
        }
        float c, s;
        rope_cos_sin(p, pos[i2], ip, ff, c, s);
        const long long a0 = p.neox ? ip : i0, a1 = p.neox ? ip + p.n_dims / 2 : i0 + 1;
        const float x0 = Elem<TYPE>::ld(xs + a0 * es), x1 = Elem<TYPE>::ld(xs + a1 * es);
        Elem<TYPE>::st(ds + a0 * es, x0 * c - x1 * s);
        Elem<TYPE>::st(ds + a1 * es, x0 * s + x1 * c);
    }
}

// ---- CPY / CONT / DUP: element i of src (logical order) -> element i of dst (logical order), with type conversion ----------------
template <int TS, int TD> __global__ void __launch_bounds__(256) cpy_kernel(T4 s, T4 d, long long n) {
    pdl_grid_sync();
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        long long r = i;
        const long long s0 = r % s.ne[0]; r /= s.ne[0];
        const long long s1 = r % s.ne[1]; r /= s.ne[1];
        const long long s2 = r % s.ne[2], s3 = r / s.ne[2];
        r = i;
        const long long d0 = r % d.ne[0]; r /= d.ne[0];
        const long long d1 = r % d.ne[1]; r /= d.ne[1];
        const long long d2 = r % d.ne[2], d3 = r / d.ne[2];
        Elem<TD>::st(d.data + d0 * d.nb[0] + d1 * d.nb[1] + d2 * d.nb[2] + d3 * d.nb[3],
                     Elem<TS>::ld(s.data + s0 * s.nb[0] + s1 * s.nb[1] + s2 * s.nb[2] + s3 * s.nb[3]));
    }
}

// f32 -> f32 with unit-stride rows of a multiple of four values on both sides, everything 16-byte aligned (the CONT behind the
// attention's permute at prompt sizes: 8 MB per layer): four values per thread, one index decomposition per four
__global__ void __launch_bounds__(256) cpy_f32_vec4_kernel(T4 s, T4 d, long long n4) {
    pdl_grid_sync();
    for (long long q = (long long) blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (long long) gridDim.x * blockDim.x) {
        const long long i = q * 4;
        long long r = i;
        const long long s0 = r % s.ne[0]; r /= s.ne[0];
        const long long s1 = r % s.ne[1]; r /= s.ne[1];
        const long long s2 = r % s.ne[2], s3 = r / s.ne[2];
        r = i;
        const long long d0 = r % d.ne[0]; r /= d.ne[0];
        const long long d1 = r % d.ne[1]; r /= d.ne[1];
        const long long d2 = r % d.ne[2], d3 = r / d.ne[2];
        *(float4 *) (d.data + d0 * 4 + d1 * d.nb[1] + d2 * d.nb[2] + d3 * d.nb[3]) =
            *(const float4 *) (s.data + s0 * 4 + s1 * s.nb[1] + s2 * s.nb[2] + s3 * s.nb[3]);
    }
}

// ---- SET_ROWS: dst[idx[i1, i2 % ne11, i3 % ne12]][i2][i3] = src[i1][i2][i3] (rows), idx i64 ------------------------------------
template <int TD> __global__ void __launch_bounds__(256) set_rows_kernel(T4 s, T4 idx, T4 d, long long n) {
    pdl_grid_sync();
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        long long r = i;
        const long long i0 = r % s.ne[0]; r /= s.ne[0];
        const long long i1 = r % s.ne[1]; r /= s.ne[1];
        const long long i2 = r % s.ne[2], i3 = r / s.ne[2];
        const long long row = *(const long long *) (idx.data + i1 * idx.nb[0] + (i2 % idx.ne[1]) * idx.nb[1] + (i3 % idx.ne[2]) * idx.nb[2]);
        const float v = *(const float *) (s.data + i0 * s.nb[0] + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]);
        Elem<TD>::st(d.data + i0 * d.nb[0] + row * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3], v);
    }
}

// two SET_ROWS in one launch (the K and the V write of a layer are consecutive nodes; blockIdx.y picks the tensor)
template <int TD> __global__ void __launch_bounds__(256) set_rows_pair_kernel(T4 s0, T4 idx0, T4 d0, long long n0, T4 s1, T4 idx1, T4 d1, long long n1) {
    pdl_grid_sync();
    const T4 &      s   = blockIdx.y ? s1 : s0;
    const T4 &      idx = blockIdx.y ? idx1 : idx0;
    const T4 &      d   = blockIdx.y ? d1 : d0;
    const long long n   = blockIdx.y ? n1 : n0;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        long long r = i;
        const long long i0 = r % s.ne[0]; r /= s.ne[0];
        const long long i1 = r % s.ne[1]; r /= s.ne[1];
        const long long i2 = r % s.ne[2], i3 = r / s.ne[2];
        const long long row = *(const long long *) (idx.data + i1 * idx.nb[0] + (i2 % idx.ne[1]) * idx.nb[1] + (i3 % idx.ne[2]) * idx.nb[2]);
        const float v = *(const float *) (s.data + i0 * s.nb[0] + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]);
        Elem<TD>::st(d.data + i0 * d.nb[0] + row * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3], v);
    }
}

// what follows a decode step's q / k / v matmuls, in one launch (blockIdx.y picks the part): ROPE of q into its own result,
// ROPE of k stored straight into the K cache rows (the rotated k has no other reader), v stored into the V cache rows.
// Arithmetic and rounding are those of rope_kernel followed by set_rows_kernel (rotate in f32, one conversion at the store).
template <int TD> __global__ void __launch_bounds__(256) rope_store_kernel(T4 q, T4 qd, T4 k, const int * pos, const float * ff, RopeP p,
                                                                            T4 kidx, T4 kd, T4 v, T4 vidx, T4 vd, long long nv,
                                                                            const float * qb, const float * kb, const float * vb, long long vb_len) {
    pdl_grid_sync();
    if (blockIdx.y == 2) {  // v: a plain SET_ROWS (+ the bias row: element e of a token is element (flat index % width) of the bias)
        for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long) gridDim.x * blockDim.x) {
            const long long i0 = i % v.ne[0], i1 = i / v.ne[0];
            const long long row = *(const long long *) (vidx.data + i1 * vidx.nb[0]);
            float x = *(const float *) (v.data + i0 * v.nb[0] + i1 * v.nb[1]);
            if (vb) x += vb[i % vb_len];
            Elem<TD>::st(vd.data + i0 * vd.nb[0] + row * vd.nb[1], x);
        }
        return;
    }
    const bool      store = blockIdx.y == 1;
    const T4 &      x     = store ? k : q;
    const float *   bias  = store ? kb : qb;
    const long long half = x.ne[0] / 2, n = half * x.ne[1] * x.ne[2];
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const long long ip = i % half, r = i / half, i1 = r % x.ne[1], i2 = r / x.ne[1], i0 = 2 * ip;
        const char * xs = x.data + i1 * x.nb[1] + i2 * x.nb[2];
        long long a0 = i0, a1 = i0 + 1;
        if (i0 < p.n_dims && p.neox) { a0 = ip; a1 = ip + p.n_dims / 2; }
        float x0 = *(const float *) (xs + a0 * 4), x1 = *(const float *) (xs + a1 * 4);
        if (bias) {  // the ADD in front of the ROPE (q / k / v biases): one rounding, as the separate operator
            x0 += bias[i1 * x.ne[0] + a0];
            x1 += bias[i1 * x.ne[0] + a1];
        }
        float y0 = x0, y1 = x1;
        if (i0 < p.n_dims) {
            float c, s;
            rope_cos_sin(p, pos[i2], ip, ff, c, s);
            y0 = x0 * c - x1 * s;
            y1 = x0 * s + x1 * c;
        }
        if (store) {  // element (a, head i1) of token i2 is element i1 * ne0 + a of the cache row idx[i2]
            char * ds = kd.data + *(const long long *) (kidx.data + i2 * kidx.nb[0]) * kd.nb[1] + i1 * x.ne[0] * kd.nb[0];
            Elem<TD>::st(ds + a0 * kd.nb[0], y0);
            Elem<TD>::st(ds + a1 * kd.nb[0], y1);
        } else {
            char * ds = qd.data + i1 * qd.nb[1] + i2 * qd.nb[2];
            *(float *) (ds + a0 * 4) = y0;
            *(float *) (ds + a1 * 4) = y1;
        }
    }
}

// ---- expert-parallel combine: row r of src goes to row row_index[r] of dst on THIS GPU and on every peer (NVLink stores) --------
// (no reference counterpart: experts are sharded by id, activations and routing are replicated, so "dispatch" is
// free and "combine" is each rank storing the rows it computed into every rank's result — no all-to-all)
struct ScatterP {
    const float * src;
    const int *   row_index;
    float *       dst;
    long long     m, n_rows, dst_row_stride;  // floats
    int           n_peers;
    long long     peer_delta[7];              // bytes, as b200q_mul_mat_group_peers
};
__global__ void __launch_bounds__(256) scatter_rows_peers_kernel(const ScatterP p) {
    pdl_grid_sync();
    const long long n = p.n_rows * p.m;
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const long long r = i / p.m, c = i - r * p.m;
        const float     v = p.src[i];
        float *         d = p.dst + (long long) p.row_index[r] * p.dst_row_stride + c;
        *d = v;
        for (int q = 0; q < p.n_peers; ++q) *(float *) ((char *) d + p.peer_delta[q]) = v;
    }
}

// ---- MoE combine: dst[i0, t] = sum_s e[i0, s, t] * w[s, t] — llama.cpp emits MUL(experts, weights), one VIEW per used expert and a
// chain of ADDs (src/llama-graph.cpp build_moe_ffn: 4 launches per layer for top-4); same arithmetic and order: every product
// rounded, then ((p0 + p1) + p2) + p3
__global__ void __launch_bounds__(256) moe_weighted_sum_kernel(T4 e, T4 w, T4 d, long long n) {
    pdl_grid_sync();
    const int n_used = (int) e.ne[1];
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const long long i0 = i % e.ne[0], t = i / e.ne[0];
        const char * ep = e.data + i0 * 4 + t * e.nb[2];
        const char * wp = w.data + t * w.nb[2];
        float acc = __fmul_rn(*(const float *) ep, *(const float *) wp);
        for (int s = 1; s < n_used; ++s) {
            acc = __fadd_rn(acc, __fmul_rn(*(const float *) (ep + (long long) s * e.nb[1]), *(const float *) (wp + (long long) s * w.nb[1])));
        }
        *(float *) (d.data + i0 * 4 + t * d.nb[1]) = acc;
    }
}

// ---- ADD_ID: dst[i0, i1, i2] = a[i0, i1, i2] + b[i0, ids[i1, i2]] (the per-expert bias of a MoE layer; add-id.cu) ----------------
__global__ void __launch_bounds__(256) add_id_kernel(T4 a, T4 b, T4 ids, T4 d, long long n) {
    pdl_grid_sync();
    for (long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        long long r = i;
        const long long i0 = r % a.ne[0]; r /= a.ne[0];
        const long long i1 = r % a.ne[1], i2 = r / a.ne[1];
        const int e = *(const int *) (ids.data + i1 * ids.nb[0] + i2 * ids.nb[1]);
        const float x = *(const float *) (a.data + i0 * 4 + i1 * a.nb[1] + i2 * a.nb[2]);
        const float y = *(const float *) (b.data + i0 * 4 + (long long) e * b.nb[1]);
        *(float *) (d.data + i0 * 4 + i1 * d.nb[1] + i2 * d.nb[2]) = x + y;
    }
}

// ---- ARGSORT: per row, the indices that sort it (MoE routers: top-k = a view of the descending argsort; argsort.cu) ----------
// one CTA per row, bitonic network over the next power of two (padding sorts last), ne0 <= 1024
__global__ void __launch_bounds__(1024) argsort_kernel(T4 s, T4 d, int npad, int desc) {
    pdl_grid_sync();
    extern __shared__ int as_idx[];
    const long long r  = blockIdx.x;
    const long long i1 = r % s.ne[1], i2 = (r / s.ne[1]) % s.ne[2], i3 = r / (s.ne[1] * s.ne[2]);
    const char * row = s.data + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3];
    const int n = (int) s.ne[0], t = threadIdx.x;
    auto val = [&](int i) { return *(const float *) (row + (long long) i * 4); };
    if (t < npad) as_idx[t] = t;
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int o = t ^ j;
            if (t < npad && o > t) {
                const int a = as_idx[t], b = as_idx[o];
                // "a sorts after b": padding (index >= n) after everything, then by value
                const bool after = a >= n ? b < n : (b < n && (desc ? val(a) < val(b) : val(a) > val(b)));
                const bool before = b >= n ? a < n : (a < n && (desc ? val(b) < val(a) : val(b) > val(a)));
                const bool up = (t & k) == 0;
                if (up ? after : before) {
                    as_idx[t] = b;
                    as_idx[o] = a;
                }
            }
            __syncthreads();
        }
    }
    if (t < n) *(int *) (d.data + (long long) t * 4 + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = as_idx[t];
}

// ---- MoE router behind the logits matmul (gpt-oss gating: src/llama-graph.cpp build_moe_ffn with a router bias and
// LLAMA_EXPERT_GATING_FUNC_TYPE_SOFTMAX_WEIGHT): ADD(bias) -> ARGSORT(desc) -> top-k view -> GET_ROWS -> SOFT_MAX as one launch.
// One CTA per token; the comparisons are argsort_kernel's (same order, ties included), the soft_max is max / expf / sum / scale
// by the reciprocal over the n_used selected logits in slot order.
__global__ void __launch_bounds__(1024) moe_route_kernel(T4 lg, const float * bias, T4 biased, T4 sorted, T4 wout, int npad, int n_used) {
    pdl_grid_sync();
    extern __shared__ int rt_smem[];
    int *   idx = rt_smem;
    float * v   = (float *) (rt_smem + npad);
    const long long tok = blockIdx.x;
    const int n = (int) lg.ne[0], t = threadIdx.x;
    if (t < n) {
        float x = *(const float *) (lg.data + (long long) t * 4 + tok * lg.nb[1]);
        if (bias) x += bias[t];
        v[t] = x;
        *(float *) (biased.data + (long long) t * 4 + tok * biased.nb[1]) = x;
    }
    if (t < npad) idx[t] = t;
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int o = t ^ j;
            if (t < npad && o > t) {
                const int a = idx[t], b = idx[o];
                const bool after  = a >= n ? b < n : (b < n && v[a] < v[b]);
                const bool before = b >= n ? a < n : (a < n && v[b] < v[a]);
                const bool up = (t & k) == 0;
                if (up ? after : before) {
                    idx[t] = b;
                    idx[o] = a;
                }
            }
            __syncthreads();
        }
    }
    if (t < n) *(int *) (sorted.data + (long long) t * 4 + tok * sorted.nb[1]) = idx[t];
    if (t == 0 && n_used > 0) {
        float mx = -INFINITY, sum = 0.0f;
        for (int q = 0; q < n_used; ++q) mx = fmaxf(mx, v[idx[q]]);
        float * wp = (float *) (wout.data + tok * wout.nb[1]);
        for (int q = 0; q < n_used; ++q) {
            const float e = expf(v[idx[q]] - mx);
            wp[q] = e;
            sum += e;
        }
        const float inv = 1.0f / sum;
        for (int q = 0; q < n_used; ++q) wp[q] *= inv;
    }
}

// GET_ROWS of single logits by the selected expert ids + SOFT_MAX over them (the second half of the router: llama.cpp's graph
// order puts the expert matmuls between the top-k and this): weights[q, t] = soft_max_q(probs[ids[q, t], t]); a thread per token
__global__ void __launch_bounds__(128) moe_gate_weights_kernel(T4 probs, T4 ids, T4 wout, long long n_tok, int n_used) {
    pdl_grid_sync();
    const long long tok = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (tok >= n_tok) return;
    const char * ip = ids.data + tok * ids.nb[1];
    const char * pp = probs.data + tok * probs.nb[2];
    float *      wp = (float *) (wout.data + tok * wout.nb[1]);
    float mx = -INFINITY, sum = 0.0f;
    for (int q = 0; q < n_used; ++q) mx = fmaxf(mx, *(const float *) (pp + (long long) *(const int *) (ip + q * ids.nb[0]) * probs.nb[1]));
    for (int q = 0; q < n_used; ++q) {
        const float e = expf(*(const float *) (pp + (long long) *(const int *) (ip + q * ids.nb[0]) * probs.nb[1]) - mx);
        wp[q] = e;
        sum += e;
    }
    const float inv = 1.0f / sum;
    for (int q = 0; q < n_used; ++q) wp[q] *= inv;
}

template <int TS> int cpy_to(const b200q_tensor * s, const b200q_tensor * d, long long n, cudaStream_t st) {
    if (TS == B200Q_TYPE_F32 && d->type == B200Q_TYPE_F32 && s->nb[0] == 4 && d->nb[0] == 4 && s->ne[0] % 4 == 0 && d->ne[0] % 4 == 0) {
        bool al = true;
        for (const b200q_tensor * t : { s, d }) {
            al = al && (((uintptr_t) t->data | (unsigned long long) t->nb[1] | (unsigned long long) t->nb[2] | (unsigned long long) t->nb[3]) & 15) == 0;
        }
        if (al) {
            launch_pdl(cpy_f32_vec4_kernel, dim3(grid_for(n / 4, 256)), dim3(256), 0, st, t4(s), t4(d), n / 4);
            return launched();
        }
    }
    const int g = grid_for(n, 256);
    switch (d->type) {
        case B200Q_TYPE_F32:  launch_pdl(cpy_kernel<TS, B200Q_TYPE_F32>, dim3(g), dim3(256), 0, st, t4(s), t4(d), n); break;
        case B200Q_TYPE_F16:  launch_pdl(cpy_kernel<TS, B200Q_TYPE_F16>, dim3(g), dim3(256), 0, st, t4(s), t4(d), n); break;
        case B200Q_TYPE_BF16: launch_pdl(cpy_kernel<TS, B200Q_TYPE_BF16>, dim3(g), dim3(256), 0, st, t4(s), t4(d), n); break;
        default: return ops_fail(B200Q_ENOTSUP, "cpy: destination type");
    }
    return launched();
}

}  // namespace

extern "C" {

// every row of every operand starts on a 16-byte boundary and holds a multiple of four values
static int norm_vec4(const b200q_tensor * a, const b200q_tensor * b, const b200q_tensor * s, const b200q_tensor * w, const b200q_tensor * d) {
    if (a->ne[0] % 4) return 0;
    for (const b200q_tensor * t : { a, b, s, w, d }) {
        if (!t) continue;
        if (((uintptr_t) t->data | (unsigned long long) t->nb[1] | (unsigned long long) t->nb[2] | (unsigned long long) t->nb[3]) & 15) return 0;
    }
    return 1;
}

int b200q_op_rms_norm(const b200q_tensor * src, const b200q_tensor * dst, float eps, void * stream) {
    if (!src || !dst || src->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 || !same_shape(src, dst) || src->nb[0] != 4 || dst->nb[0] != 4) {
        return ops_fail(B200Q_ENOTSUP, "rms_norm: f32 rows with unit stride, same shape");
    }
    const long long nr = nrows(src);
    if (nr == 0 || src->ne[0] == 0) return 0;
    if (nr > 0x7fffffffLL) return ops_fail(B200Q_ENOTSUP, "rms_norm: too many rows");
    if (src->ne[0] <= 8192) {  // the row fits the registers of one CTA
        const int threads = src->ne[0] <= 2048 ? 256 : 1024;
        launch_pdl(norm_fused_kernel<false, false>, dim3((unsigned) nr), dim3(threads), 0, (cudaStream_t) stream, t4(src), t4(src), t4(src), t4(src), t4(dst), eps, 0,
                   norm_vec4(src, nullptr, nullptr, nullptr, dst));
    } else {
        launch_pdl(rms_norm_kernel, dim3((unsigned) nr), dim3(256), 0, (cudaStream_t) stream, t4(src), t4(dst), eps);
    }
    return launched();
}

// fused [ADD +] RMS_NORM + MUL (see norm_fused_kernel): a (+ b) -> sum_out (optional) ; dst = rms_norm(a (+ b)) * w
int b200q_op_norm_fused(const b200q_tensor * a, const b200q_tensor * b, const b200q_tensor * sum_out, const b200q_tensor * w,
                        const b200q_tensor * dst, float eps, void * stream) {
    auto f32row = [](const b200q_tensor * t) { return t && t->type == B200Q_TYPE_F32 && t->nb[0] == 4; };
    if (!f32row(a) || !f32row(w) || !f32row(dst) || !same_shape(a, dst) || w->ne[0] != a->ne[0] || a->ne[0] > 8192 || a->ne[0] < 1 ||
        (b && (!f32row(b) || !same_shape(a, b))) || (sum_out && (!b || !f32row(sum_out) || !same_shape(a, sum_out))) ||
        a->ne[1] % w->ne[1] || a->ne[2] % w->ne[2] || a->ne[3] % w->ne[3]) {
        return ops_fail(B200Q_ENOTSUP, "norm_fused: f32 rows of at most 8192 values, b / sum_out shaped like a, w broadcast over rows");
    }
    const long long nr = nrows(a);
    if (nr == 0) return 0;
    if (nr > 0x7fffffffLL) return ops_fail(B200Q_ENOTSUP, "norm_fused: too many rows");
    const int threads = a->ne[0] <= 2048 ? 256 : 1024;
    cudaStream_t st = (cudaStream_t) stream;
    const int vec4 = norm_vec4(a, b, sum_out, w, dst);
    if (b) launch_pdl(norm_fused_kernel<true, true>, dim3((unsigned) nr), dim3(threads), 0, st, t4(a), t4(b), t4(sum_out ? sum_out : a), t4(w), t4(dst), eps, sum_out ? 1 : 0, vec4);
    else launch_pdl(norm_fused_kernel<false, true>, dim3((unsigned) nr), dim3(threads), 0, st, t4(a), t4(a), t4(a), t4(w), t4(dst), eps, 0, vec4);
    return launched();
}

int b200q_op_bin(int op, const b200q_tensor * a, const b200q_tensor * b, const b200q_tensor * dst, void * stream) {
    if (!a || !b || !dst || a->type != B200Q_TYPE_F32 || b->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 || !same_shape(a, dst)) {
        return ops_fail(B200Q_ENOTSUP, "bin: f32 operands, dst shaped like src0");
    }
    for (int i = 0; i < 4; ++i) if (b->ne[i] < 1 || dst->ne[i] % b->ne[i]) return ops_fail(B200Q_EINVAL, "bin: src1 does not broadcast");
    const long long n = nelem(dst);
    if (n == 0) return 0;
    const int g = grid_for(n, 256);
    cudaStream_t st = (cudaStream_t) stream;
    switch (op) {
        case B200Q_BIN_ADD: launch_pdl(bin_kernel<B200Q_BIN_ADD>, dim3(g), dim3(256), 0, st, t4(a), t4(b), t4(dst), n); break;
        case B200Q_BIN_SUB: launch_pdl(bin_kernel<B200Q_BIN_SUB>, dim3(g), dim3(256), 0, st, t4(a), t4(b), t4(dst), n); break;
        case B200Q_BIN_MUL: launch_pdl(bin_kernel<B200Q_BIN_MUL>, dim3(g), dim3(256), 0, st, t4(a), t4(b), t4(dst), n); break;
        case B200Q_BIN_DIV: launch_pdl(bin_kernel<B200Q_BIN_DIV>, dim3(g), dim3(256), 0, st, t4(a), t4(b), t4(dst), n); break;
        default: return ops_fail(B200Q_EINVAL, "bin: unknown op");
    }
    return launched();
}

int b200q_op_scale(const b200q_tensor * src, const b200q_tensor * dst, float scale, float bias, void * stream) {
    if (!src || !dst || src->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 || !contiguous(src) || !contiguous(dst) || nelem(src) != nelem(dst)) {
        return ops_fail(B200Q_ENOTSUP, "scale: contiguous f32");
    }
    const long long n = nelem(src);
    if (n == 0) return 0;
    launch_pdl(scale_kernel, dim3(grid_for(n, 256)), dim3(256), 0, (cudaStream_t) stream, (const float *) src->data, (float *) dst->data, n, scale, bias);
    return launched();
}

int b200q_op_unary(int op, const b200q_tensor * src, const b200q_tensor * dst, void * stream) {
    if (!src || !dst || src->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 || !contiguous(src) || !contiguous(dst) || nelem(src) != nelem(dst) ||
        op < B200Q_UNARY_SILU || op > B200Q_UNARY_EXP) {
        return ops_fail(B200Q_ENOTSUP, "unary: contiguous f32, known op");
    }
    const long long n = nelem(src);
    if (n == 0) return 0;
    launch_pdl(unary_kernel, dim3(grid_for(n, 256)), dim3(256), 0, (cudaStream_t) stream, (const float *) src->data, (float *) dst->data, n, op);
    return launched();
}

int b200q_op_glu(int op, const b200q_tensor * src0, const b200q_tensor * src1, const b200q_tensor * dst, int swapped, float alpha, float limit,
                 void * stream) {
    if (!src0 || !dst || src0->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 || !contiguous_rows(src0) || !contiguous_rows(dst) ||
        (src1 && (src1->type != B200Q_TYPE_F32 || !contiguous_rows(src1) || !same_shape(src0, src1))) || op < B200Q_GLU_REGLU || op > B200Q_GLU_SWIGLU_OAI) {
        return ops_fail(B200Q_ENOTSUP, "glu: f32 tensors with dense rows");
    }
    const long long nc = src1 ? src0->ne[0] : src0->ne[0] / 2, nr = nrows(src0);
    if (dst->ne[0] != nc || nrows(dst) != nr) return ops_fail(B200Q_EINVAL, "glu: dst shape");
    if (nc == 0 || nr == 0) return 0;
    const char * x = (const char *) src0->data;
    const char * g = src1 ? (const char *) src1->data : (const char *) src0->data;
    if (!src1) {  // the two halves of a row; `swapped` says which half is the gate
        x += swapped ? nc * 4 : 0;
        g += swapped ? 0 : nc * 4;
    }
    launch_pdl(glu_kernel, dim3(grid_for(nr * nc, 256)), dim3(256), 0, (cudaStream_t) stream, x, g, (char *) dst->data, nr, (int) nc, (long long) src0->nb[1],
                                                                         (long long) (src1 ? src1->nb[1] : src0->nb[1]), (long long) dst->nb[1], op, alpha, limit);
    return launched();
}

int b200q_op_soft_max(const b200q_tensor * src, const b200q_tensor * mask, const b200q_tensor * sinks, const b200q_tensor * dst, float scale,
                      float max_bias, void * stream) {
    if (!src || !dst || src->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 || !same_shape(src, dst) || src->nb[0] != 4 || !contiguous(dst)) {
        return ops_fail(B200Q_ENOTSUP, "soft_max: f32 rows, contiguous dst");
    }
    if (mask && ((mask->type != B200Q_TYPE_F32 && mask->type != B200Q_TYPE_F16) || mask->nb[0] != tsize(mask->type) || mask->ne[0] != src->ne[0] ||
                 mask->ne[1] < src->ne[1] || mask->ne[2] < 1 || mask->ne[3] < 1)) {
        return ops_fail(B200Q_ENOTSUP, "soft_max: mask layout");
    }
    if (sinks && (sinks->type != B200Q_TYPE_F32 || sinks->ne[0] != src->ne[2] || !contiguous(sinks))) return ops_fail(B200Q_ENOTSUP, "soft_max: sinks");
    const long long nr = nrows(src);
    if (nr == 0 || src->ne[0] == 0) return 0;
    if (nr > 0x7fffffffLL) return ops_fail(B200Q_ENOTSUP, "soft_max: too many rows");
    const unsigned n_head = (unsigned) src->ne[2], n_head_log2 = 1u << (unsigned) floor(log2((double) n_head));
    const float m0 = powf(2.0f, -(max_bias) / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    T4 m{};
    if (mask) m = t4(mask);
    cudaStream_t st = (cudaStream_t) stream;
    const float * sk = sinks ? (const float *) sinks->data : nullptr;
    if (src->ne[0] <= 1024 && nr >= 64) {  // many short rows: a warp each
        const unsigned g = (unsigned) ((nr + 7) / 8);
        if (mask && mask->type == B200Q_TYPE_F16) launch_pdl(soft_max_warp_kernel<true>, dim3(g), dim3(256), 0, st, t4(src), m, sk, t4(dst), scale, max_bias, m0, m1, n_head_log2, nr);
        else launch_pdl(soft_max_warp_kernel<false>, dim3(g), dim3(256), 0, st, t4(src), m, sk, t4(dst), scale, max_bias, m0, m1, n_head_log2, nr);
        return launched();
    }
    if (mask && mask->type == B200Q_TYPE_F16) {
        launch_pdl(soft_max_kernel<true>, dim3((unsigned) nr), dim3(256), 0, st, t4(src), m, sinks ? (const float *) sinks->data : nullptr, t4(dst), scale, max_bias, m0, m1, n_head_log2);
    } else {
        launch_pdl(soft_max_kernel<false>, dim3((unsigned) nr), dim3(256), 0, st, t4(src), m, sinks ? (const float *) sinks->data : nullptr, t4(dst), scale, max_bias, m0, m1, n_head_log2);
    }
    return launched();
}

int b200q_op_attn_decode(const b200q_tensor * q, const b200q_tensor * k, const b200q_tensor * v, const b200q_tensor * mask,
                         const b200q_tensor * sinks, const b200q_tensor * dst, float scale, void * stream) {
    if (!q || !k || !v || !dst || q->type != B200Q_TYPE_F32 || k->type != B200Q_TYPE_F16 || v->type != B200Q_TYPE_F16 || dst->type != B200Q_TYPE_F32 ||
        q->nb[0] != 4 || k->nb[0] != 2 || v->nb[0] != 2 || dst->nb[0] != 4 || q->ne[3] != 1 || k->ne[3] != 1 || v->ne[3] != 1 ||
        k->ne[0] != q->ne[0] || v->ne[0] != k->ne[1] || v->ne[1] != q->ne[0] || v->ne[2] != k->ne[2] || k->ne[2] < 1 || q->ne[2] % k->ne[2] ||
        dst->ne[0] != q->ne[0] || dst->ne[1] != q->ne[2] || dst->ne[2] != q->ne[1] || q->ne[0] > 512 || k->ne[1] < 1 || k->ne[1] > 8192 ||
        q->ne[1] > 65535 || (mask && ((mask->type != B200Q_TYPE_F32 && mask->type != B200Q_TYPE_F16) || mask->ne[0] < k->ne[1] || mask->ne[1] < q->ne[1] ||
                                      mask->nb[0] != tsize(mask->type) || mask->ne[2] < 1 || mask->ne[3] != 1)) ||
        (sinks && (sinks->type != B200Q_TYPE_F32 || sinks->ne[0] < q->ne[2] || sinks->nb[0] != 4))) {
        return ops_fail(B200Q_ENOTSUP, "attn_decode: q f32 [hd, n_tok, n_head], k f16 [hd, n_kv, n_head_kv], v f16 [n_kv, hd, n_head_kv], n_kv <= 8192");
    }
    if (q->ne[1] == 0 || q->ne[2] == 0) return 0;
    AttnP p{};
    p.q = t4(q); p.k = t4(k); p.v = t4(v); p.d = t4(dst);
    if (mask) p.m = t4(mask);
    p.sinks = sinks ? (const float *) sinks->data : nullptr;
    p.scale = scale; p.mask_f16 = mask && mask->type == B200Q_TYPE_F16;
    const size_t smem = (size_t) (((k->ne[1] + 3) & ~3) + ((q->ne[0] + 3) & ~3) + 256) * sizeof(float);
    launch_pdl(attn_decode_kernel, dim3((unsigned) q->ne[2], (unsigned) q->ne[1]), dim3(256), smem, (cudaStream_t) stream, p);
    return launched();
}

int b200q_op_rope(const b200q_tensor * src, const b200q_tensor * pos, const b200q_tensor * freq_factors, const b200q_tensor * dst,
                  const b200q_rope_params * rp, void * stream) {
    if (!src || !pos || !dst || !rp || (src->type != B200Q_TYPE_F32 && src->type != B200Q_TYPE_F16) || dst->type != src->type || !same_shape(src, dst) ||
        src->nb[0] != tsize(src->type) || dst->nb[0] != tsize(dst->type) || pos->type != B200Q_TYPE_I32 || pos->ne[0] < src->ne[2] || pos->nb[0] != 4) {
        return ops_fail(B200Q_ENOTSUP, "rope: f32 / f16 rows with unit stride, i32 positions");
    }
    if (rp->mode != B200Q_ROPE_NORMAL && rp->mode != B200Q_ROPE_NEOX) return ops_fail(B200Q_ENOTSUP, "rope: only the normal and NeoX modes");
    if (rp->n_dims <= 0 || rp->n_dims % 2 || rp->n_dims > src->ne[0] || src->ne[0] % 2) return ops_fail(B200Q_EINVAL, "rope: n_dims");
    if (freq_factors && (freq_factors->type != B200Q_TYPE_F32 || freq_factors->ne[0] < rp->n_dims / 2 || freq_factors->nb[0] != 4)) {
        return ops_fail(B200Q_ENOTSUP, "rope: frequency factors");
    }
    const long long n = nelem(src) / 2;
    if (n == 0) return 0;
    RopeP p;
    p.n_dims = rp->n_dims; p.neox = rp->mode == B200Q_ROPE_NEOX;
    p.theta_scale = powf(rp->freq_base, -2.0f / rp->n_dims);
    p.freq_scale = rp->freq_scale; p.ext_factor = rp->ext_factor; p.attn_factor = rp->attn_factor;
    p.corr0 = rp->corr_dims[0]; p.corr1 = rp->corr_dims[1];
    const float * ff = freq_factors ? (const float *) freq_factors->data : nullptr;
    cudaStream_t  st = (cudaStream_t) stream;
    if (src->type == B200Q_TYPE_F32) launch_pdl(rope_kernel<B200Q_TYPE_F32>, dim3(grid_for(n, 256)), dim3(256), 0, st, t4(src), (const int *) pos->data, ff, t4(dst), p);
    else launch_pdl(rope_kernel<B200Q_TYPE_F16>, dim3(grid_for(n, 256)), dim3(256), 0, st, t4(src), (const int *) pos->data, ff, t4(dst), p);
    return launched();
}

int b200q_op_cpy(const b200q_tensor * src, const b200q_tensor * dst, void * stream) {
    if (!src || !dst || nelem(src) != nelem(dst)) return ops_fail(B200Q_EINVAL, "cpy: element counts differ");
    const long long n = nelem(src);
    if (n == 0) return 0;
    switch (src->type) {
        case B200Q_TYPE_F32:  return cpy_to<B200Q_TYPE_F32>(src, dst, n, (cudaStream_t) stream);
        case B200Q_TYPE_F16:  return cpy_to<B200Q_TYPE_F16>(src, dst, n, (cudaStream_t) stream);
        case B200Q_TYPE_BF16: return cpy_to<B200Q_TYPE_BF16>(src, dst, n, (cudaStream_t) stream);
        default: return ops_fail(B200Q_ENOTSUP, "cpy: source type");
    }
}

int b200q_op_add_id(const b200q_tensor * a, const b200q_tensor * b, const b200q_tensor * ids, const b200q_tensor * dst, void * stream) {
    if (!a || !b || !ids || !dst || a->type != B200Q_TYPE_F32 || b->type != B200Q_TYPE_F32 || ids->type != B200Q_TYPE_I32 ||
        dst->type != B200Q_TYPE_F32 || a->nb[0] != 4 || b->nb[0] != 4 || dst->nb[0] != 4 || a->ne[3] != 1 || b->ne[0] != a->ne[0] ||
        ids->ne[0] != a->ne[1] || ids->ne[1] != a->ne[2] || nelem(a) != nelem(dst)) {
        return ops_fail(B200Q_ENOTSUP, "add_id: a f32 [n, used, tokens], b f32 [n, experts], ids i32 [used, tokens]");
    }
    const long long n = nelem(a);
    if (n == 0) return 0;
    launch_pdl(add_id_kernel, dim3(grid_for(n, 256)), dim3(256), 0, (cudaStream_t) stream, t4(a), t4(b), t4(ids), t4(dst), n);
    return launched();
}

int b200q_op_argsort(const b200q_tensor * src, const b200q_tensor * dst, int descending, void * stream) {
    if (!src || !dst || src->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_I32 || src->nb[0] != 4 || dst->nb[0] != 4 ||
        src->ne[0] < 1 || src->ne[0] > 1024 || nelem(src) != nelem(dst)) {
        return ops_fail(B200Q_ENOTSUP, "argsort: f32 rows of at most 1024 values -> i32");
    }
    const long long rows = nelem(src) / src->ne[0];
    if (rows == 0) return 0;
    if (rows > 0x7fffffffll) return ops_fail(B200Q_ENOTSUP, "argsort: too many rows");
    int npad = 1;
    while (npad < src->ne[0]) npad <<= 1;
    launch_pdl(argsort_kernel, dim3((unsigned) rows), dim3(npad < 32 ? 32 : npad), npad * sizeof(int), (cudaStream_t) stream, t4(src), t4(dst), npad, descending ? 1 : 0);
    return launched();
}

int b200q_op_moe_route(const b200q_tensor * logits, const b200q_tensor * bias, const b200q_tensor * biased, const b200q_tensor * sorted,
                       const b200q_tensor * weights, void * stream) {
    if (!logits || !biased || !sorted || logits->type != B200Q_TYPE_F32 || biased->type != B200Q_TYPE_F32 || sorted->type != B200Q_TYPE_I32 ||
        logits->nb[0] != 4 || biased->nb[0] != 4 || sorted->nb[0] != 4 || logits->ne[0] < 1 ||
        logits->ne[0] > 1024 || logits->ne[2] != 1 || logits->ne[3] != 1 || !same_shape(logits, biased) || !same_shape(logits, sorted) ||
        (weights && (weights->type != B200Q_TYPE_F32 || weights->nb[0] != 4 || weights->ne[0] < 1 || weights->ne[0] > logits->ne[0] ||
                     weights->ne[1] != logits->ne[1] || weights->ne[2] != 1 || weights->ne[3] != 1)) ||
        (bias && (bias->type != B200Q_TYPE_F32 || bias->nb[0] != 4 || bias->ne[0] != logits->ne[0] || bias->ne[1] != 1 || bias->ne[2] != 1 || bias->ne[3] != 1))) {
        return ops_fail(B200Q_ENOTSUP, "moe_route: f32 logits [experts <= 1024, tokens], f32 bias [experts], i32 order, f32 weights [used, tokens]");
    }
    if (logits->ne[1] == 0) return 0;
    if (logits->ne[1] > 0x7fffffffll) return ops_fail(B200Q_ENOTSUP, "moe_route: too many tokens");
    int npad = 1;
    while (npad < logits->ne[0]) npad <<= 1;
    launch_pdl(moe_route_kernel, dim3((unsigned) logits->ne[1]), dim3(npad < 32 ? 32 : npad), npad * 8, (cudaStream_t) stream, t4(logits),
               bias ? (const float *) bias->data : nullptr, t4(biased), t4(sorted), weights ? t4(weights) : t4(biased), npad, weights ? (int) weights->ne[0] : 0);
    return launched();
}

int b200q_op_moe_gate_weights(const b200q_tensor * probs, const b200q_tensor * ids, const b200q_tensor * weights, void * stream) {
    if (!probs || !ids || !weights || probs->type != B200Q_TYPE_F32 || ids->type != B200Q_TYPE_I32 || weights->type != B200Q_TYPE_F32 ||
        probs->ne[0] != 1 || probs->ne[3] != 1 || ids->ne[2] != 1 || ids->ne[3] != 1 || ids->ne[1] != probs->ne[2] || ids->ne[0] < 1 || ids->ne[0] > 1024 ||
        weights->nb[0] != 4 || weights->ne[0] != ids->ne[0] || weights->ne[1] != ids->ne[1] || weights->ne[2] != 1 || weights->ne[3] != 1) {
        return ops_fail(B200Q_ENOTSUP, "moe_gate_weights: f32 probs [1, experts, tokens], i32 ids [used, tokens], f32 weights [used, tokens]");
    }
    const long long n_tok = ids->ne[1];
    if (n_tok == 0) return 0;
    launch_pdl(moe_gate_weights_kernel, dim3(grid_for(n_tok, 128)), dim3(128), 0, (cudaStream_t) stream, t4(probs), t4(ids), t4(weights), n_tok, (int) ids->ne[0]);
    return launched();
}

static bool set_rows_ok(const b200q_tensor * src, const b200q_tensor * idx, const b200q_tensor * dst) {
    return src && idx && dst && src->type == B200Q_TYPE_F32 && idx->type == B200Q_TYPE_I64 && src->ne[0] == dst->ne[0] && src->ne[2] == dst->ne[2] &&
           src->ne[3] == dst->ne[3] && idx->ne[0] == src->ne[1] && idx->ne[1] >= 1 && idx->ne[2] >= 1 && src->ne[2] % idx->ne[1] == 0 &&
           src->ne[3] % idx->ne[2] == 0;
}
// two independent SET_ROWS with the same destination type in one launch (the K and V cache writes of a layer)
int b200q_op_set_rows_pair(const b200q_tensor * src0, const b200q_tensor * idx0, const b200q_tensor * dst0, const b200q_tensor * src1,
                           const b200q_tensor * idx1, const b200q_tensor * dst1, void * stream) {
    if (!set_rows_ok(src0, idx0, dst0) || !set_rows_ok(src1, idx1, dst1) || dst0->type != dst1->type) {
        return ops_fail(B200Q_ENOTSUP, "set_rows_pair: two set_rows (f32 rows, i64 indices) with one destination type");
    }
    const long long n0 = nelem(src0), n1 = nelem(src1);
    if (n0 == 0 && n1 == 0) return 0;
    const dim3 g((unsigned) grid_for(n0 > n1 ? n0 : n1, 256), 2);
    cudaStream_t st = (cudaStream_t) stream;
    switch (dst0->type) {
        case B200Q_TYPE_F32:  launch_pdl(set_rows_pair_kernel<B200Q_TYPE_F32>, g, dim3(256), 0, st, t4(src0), t4(idx0), t4(dst0), n0, t4(src1), t4(idx1), t4(dst1), n1); break;
        case B200Q_TYPE_F16:  launch_pdl(set_rows_pair_kernel<B200Q_TYPE_F16>, g, dim3(256), 0, st, t4(src0), t4(idx0), t4(dst0), n0, t4(src1), t4(idx1), t4(dst1), n1); break;
        case B200Q_TYPE_BF16: launch_pdl(set_rows_pair_kernel<B200Q_TYPE_BF16>, g, dim3(256), 0, st, t4(src0), t4(idx0), t4(dst0), n0, t4(src1), t4(idx1), t4(dst1), n1); break;
        default: return ops_fail(B200Q_ENOTSUP, "set_rows_pair: destination type");
    }
    return launched();
}

int b200q_op_rope_store(const b200q_tensor * q_src, const b200q_tensor * q_dst, const b200q_tensor * k_src, const b200q_tensor * pos,
                        const b200q_tensor * freq_factors, const b200q_rope_params * rp, const b200q_tensor * k_idx, const b200q_tensor * k_dst,
                        const b200q_tensor * v_src, const b200q_tensor * v_idx, const b200q_tensor * v_dst, const b200q_tensor * q_bias,
                        const b200q_tensor * k_bias, const b200q_tensor * v_bias, void * stream) {
    auto bias_ok = [](const b200q_tensor * b, long long width) {
        return !b || (b->type == B200Q_TYPE_F32 && b->nb[0] == 4 && b->ne[0] == width && b->ne[1] == 1 && b->ne[2] == 1 && b->ne[3] == 1);
    };
    auto rope_ok = [&](const b200q_tensor * x) {
        return x && x->type == B200Q_TYPE_F32 && x->nb[0] == 4 && x->ne[0] % 2 == 0 && x->ne[3] == 1 && pos && pos->ne[0] >= x->ne[2];
    };
    auto store_ok = [&](long long width, long long rows, const b200q_tensor * idx, const b200q_tensor * d) {
        return idx && d && idx->type == B200Q_TYPE_I64 && idx->ne[0] == rows && idx->ne[1] == 1 && idx->ne[2] == 1 && d->ne[0] == width &&
               d->ne[2] == 1 && d->ne[3] == 1;
    };
    if (!rp || !rope_ok(q_src) || !rope_ok(k_src) || !q_dst || q_dst->type != B200Q_TYPE_F32 || q_dst->nb[0] != 4 || !same_shape(q_src, q_dst) ||
        pos->type != B200Q_TYPE_I32 || pos->nb[0] != 4 || q_src->ne[2] != k_src->ne[2] || q_src->ne[0] != k_src->ne[0] ||
        !store_ok(k_src->ne[0] * k_src->ne[1], k_src->ne[2], k_idx, k_dst) || !v_src || v_src->type != B200Q_TYPE_F32 || v_src->ne[2] != 1 ||
        v_src->ne[3] != 1 || !store_ok(v_src->ne[0], v_src->ne[1], v_idx, v_dst) || k_dst->type != v_dst->type) {
        return ops_fail(B200Q_ENOTSUP, "rope_store: f32 q / k [head dim, heads, tokens], f32 v [width, tokens], i64 row indices, caches of one type");
    }
    if (!bias_ok(q_bias, q_src->ne[0] * q_src->ne[1]) || !bias_ok(k_bias, k_src->ne[0] * k_src->ne[1]) ||
        (v_bias && (v_bias->type != B200Q_TYPE_F32 || v_bias->nb[0] != 4 || v_bias->ne[0] < 1 || v_bias->ne[1] != 1 || v_bias->ne[2] != 1 || v_bias->ne[3] != 1 ||
                    (v_src->ne[0] * v_src->ne[1]) % v_bias->ne[0] || v_src->nb[1] != v_src->ne[0] * 4))) {
        return ops_fail(B200Q_ENOTSUP, "rope_store: biases are f32 vectors of a token's width (v: contiguous rows)");
    }
    if (rp->mode != B200Q_ROPE_NORMAL && rp->mode != B200Q_ROPE_NEOX) return ops_fail(B200Q_ENOTSUP, "rope_store: only the normal and NeoX modes");
    if (rp->n_dims <= 0 || rp->n_dims % 2 || rp->n_dims > q_src->ne[0]) return ops_fail(B200Q_EINVAL, "rope_store: n_dims");
    if (freq_factors && (freq_factors->type != B200Q_TYPE_F32 || freq_factors->ne[0] < rp->n_dims / 2 || freq_factors->nb[0] != 4)) {
        return ops_fail(B200Q_ENOTSUP, "rope_store: frequency factors");
    }
    const long long nq = nelem(q_src) / 2, nk = nelem(k_src) / 2, nv = nelem(v_src);
    long long most = nq > nk ? nq : nk;
    if (nv > most) most = nv;
    if (most == 0) return 0;
    RopeP p;
    p.n_dims = rp->n_dims; p.neox = rp->mode == B200Q_ROPE_NEOX;
    p.theta_scale = powf(rp->freq_base, -2.0f / rp->n_dims);
    p.freq_scale = rp->freq_scale; p.ext_factor = rp->ext_factor; p.attn_factor = rp->attn_factor;
    p.corr0 = rp->corr_dims[0]; p.corr1 = rp->corr_dims[1];
    const float * ff = freq_factors ? (const float *) freq_factors->data : nullptr;
    const dim3    g((unsigned) grid_for(most, 256), 3);
    cudaStream_t  st = (cudaStream_t) stream;
#define B200Q_ROPE_STORE(TD)                                                                                                               \
    launch_pdl(rope_store_kernel<TD>, g, dim3(256), 0, st, t4(q_src), t4(q_dst), t4(k_src), (const int *) pos->data, ff, p, t4(k_idx), t4(k_dst), \
               t4(v_src), t4(v_idx), t4(v_dst), nv, q_bias ? (const float *) q_bias->data : nullptr, k_bias ? (const float *) k_bias->data : nullptr, \
               v_bias ? (const float *) v_bias->data : nullptr, v_bias ? (long long) v_bias->ne[0] : 1ll)
    switch (k_dst->type) {
        case B200Q_TYPE_F32:  B200Q_ROPE_STORE(B200Q_TYPE_F32); break;
        case B200Q_TYPE_F16:  B200Q_ROPE_STORE(B200Q_TYPE_F16); break;
        case B200Q_TYPE_BF16: B200Q_ROPE_STORE(B200Q_TYPE_BF16); break;
        default: return ops_fail(B200Q_ENOTSUP, "rope_store: cache type");
    }
#undef B200Q_ROPE_STORE
    return launched();
}

int b200q_scatter_rows_peers(const float * src, const int32_t * row_index, float * dst, int64_t n_rows, int64_t m, int64_t dst_row_stride,
                             int n_peers, const int64_t * peer_delta, void * stream) {
    if (n_rows < 0 || m < 0 || n_peers < 0 || n_peers > 7 || (n_peers && !peer_delta) || (n_rows && m && (!src || !row_index || !dst))) {
        return ops_fail(B200Q_EINVAL, "scatter_rows_peers: arguments (0..7 peers)");
    }
    if (n_rows == 0 || m == 0) return 0;
    ScatterP p{};
    p.src = src; p.row_index = row_index; p.dst = dst; p.m = m; p.n_rows = n_rows; p.dst_row_stride = dst_row_stride; p.n_peers = n_peers;
    for (int q = 0; q < n_peers; ++q) p.peer_delta[q] = peer_delta[q];
    launch_pdl(scatter_rows_peers_kernel, dim3(grid_for(n_rows * m, 256)), dim3(256), 0, (cudaStream_t) stream, p);
    return launched();
}

// experts f32 [n, n_used, n_tok] * weights f32 [1, n_used, n_tok], summed over the used experts -> dst f32 [n, n_tok]
int b200q_op_moe_weighted_sum(const b200q_tensor * experts, const b200q_tensor * weights, const b200q_tensor * dst, void * stream) {
    if (!experts || !weights || !dst || experts->type != B200Q_TYPE_F32 || weights->type != B200Q_TYPE_F32 || dst->type != B200Q_TYPE_F32 ||
        experts->nb[0] != 4 || dst->nb[0] != 4 || experts->ne[3] != 1 || weights->ne[0] != 1 || weights->ne[1] != experts->ne[1] ||
        weights->ne[2] != experts->ne[2] || weights->ne[3] != 1 || dst->ne[0] != experts->ne[0] || dst->ne[1] != experts->ne[2] || experts->ne[1] < 1) {
        return ops_fail(B200Q_ENOTSUP, "moe_weighted_sum: experts f32 [n, n_used, n_tok], weights f32 [1, n_used, n_tok], dst f32 [n, n_tok]");
    }
    const long long n = experts->ne[0] * experts->ne[2];
    if (n == 0) return 0;
    launch_pdl(moe_weighted_sum_kernel, dim3(grid_for(n, 256)), dim3(256), 0, (cudaStream_t) stream, t4(experts), t4(weights), t4(dst), n);
    return launched();
}

int b200q_op_set_rows(const b200q_tensor * src, const b200q_tensor * idx, const b200q_tensor * dst, void * stream) {
    if (!src || !idx || !dst || src->type != B200Q_TYPE_F32 || idx->type != B200Q_TYPE_I64 || src->ne[0] != dst->ne[0] || src->ne[2] != dst->ne[2] ||
        src->ne[3] != dst->ne[3] || idx->ne[0] != src->ne[1] || idx->ne[1] < 1 || idx->ne[2] < 1 || src->ne[2] % idx->ne[1] || src->ne[3] % idx->ne[2]) {
        return ops_fail(B200Q_ENOTSUP, "set_rows: f32 rows, i64 indices [ne01, ne11 | ne02, ne12 | ne03]");
    }
    const long long n = nelem(src);
    if (n == 0) return 0;
    const int g = grid_for(n, 256);
    cudaStream_t st = (cudaStream_t) stream;
    switch (dst->type) {
        case B200Q_TYPE_F32:  launch_pdl(set_rows_kernel<B200Q_TYPE_F32>, dim3(g), dim3(256), 0, st, t4(src), t4(idx), t4(dst), n); break;
        case B200Q_TYPE_F16:  launch_pdl(set_rows_kernel<B200Q_TYPE_F16>, dim3(g), dim3(256), 0, st, t4(src), t4(idx), t4(dst), n); break;
        case B200Q_TYPE_BF16: launch_pdl(set_rows_kernel<B200Q_TYPE_BF16>, dim3(g), dim3(256), 0, st, t4(src), t4(idx), t4(dst), n); break;
        default: return ops_fail(B200Q_ENOTSUP, "set_rows: destination type");
    }
    return launched();
}

}  // extern "C"
