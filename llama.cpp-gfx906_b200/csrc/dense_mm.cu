// dense_mm.cu — MUL_MAT with unquantized weights (f16 / bf16 / f32) where the tensor-core kernel does not apply: any k
// (the attention products K^T q and V softmax(.) have k = head size and k = number of cached positions), strided rows
// (views into the KV cache), and few columns.  Replaces the reference's mul_mat_vec_f / mul_mat_f / cuBLAS branch for
// these types (ggml-cuda.cu:2035-2075 dispatch; mmvf.cu, mmf.cu).  Weight values are converted to f32 exactly, products are
// accumulated in f32 in a fixed order (bit-reproducible for a given shape).
//
//   * n <= 8 columns: one warp per weight row, the row is read ONCE for all columns (HBM-bound at decode: the KV cache is
//     streamed once per step); 16-byte loads when the rows are 16-byte aligned;
//   * n  > 8 columns: 64 x 64 x 16 shared-memory tiles, 4 x 4 outputs per thread, bounds-checked (odd k, m, n).
// Products with k % 64 == 0 and more than 8 columns keep going to the tcgen05 kernel (mmq.cu).
#include "dense_mm.cuh"

#include "blocks.cuh"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "pdl.cuh"
#include <cuda_runtime.h>

namespace b200q {

namespace {

template <int T> struct Elem;
template <> struct Elem<T_F16> {
    using type = __half;
    static __device__ __forceinline__ float cvt(type v) { return __half2float(v); }
};
template <> struct Elem<T_BF16> {
    using type = __nv_bfloat16;
    static __device__ __forceinline__ float cvt(type v) { return __bfloat162float(v); }
};
template <> struct Elem<T_F32W> {
    using type = float;
    static __device__ __forceinline__ float cvt(type v) { return v; }
};

struct DenseParams {
    const uint8_t * w;
    const uint8_t * act;
    uint8_t *       dst;
    int             k, m, n;
    int             ne02, ne03, ne12, ne13;
    long long       w_nb1, w_nb2, w_nb3, a_nb1, a_nb2, a_nb3, d_nb1, d_nb2, d_nb3;
    int             vec;  // rows and columns 16-byte aligned
};

// batch `b` of dst reads weight batch (i3 / r3, i2 / r2) — ggml's broadcast rule (ggml.c ggml_mul_mat: ne12 % ne02 == 0)
__device__ __forceinline__ void batch_ptrs(const DenseParams & p, int b, const uint8_t *& w, const uint8_t *& act, uint8_t *& dst) {
    const int i2 = b % p.ne12, i3 = b / p.ne12;
    const int r2 = p.ne12 / p.ne02, r3 = p.ne13 / p.ne03;
    w   = p.w + (long long) (i2 / r2) * p.w_nb2 + (long long) (i3 / r3) * p.w_nb3;
    act = p.act + (long long) i2 * p.a_nb2 + (long long) i3 * p.a_nb3;
    dst = p.dst + (long long) i2 * p.d_nb2 + (long long) i3 * p.d_nb3;
}

constexpr int GV_WARPS = 4;

// WIDE: few rows (a MoE router: 32 rows of 2880 values took 10 us with one warp per row) — one CTA per row, its four warps
// split k and add their partial sums in warp order
template <int T, int NC, bool WIDE> __global__ void __launch_bounds__(GV_WARPS * 32) dense_gemv_kernel(const DenseParams p) {
    pdl_grid_sync();
    using E = typename Elem<T>::type;
    constexpr int VEC = 16 / (int) sizeof(E);
    constexpr int LANES = WIDE ? GV_WARPS * 32 : 32;  // threads that share one row
    __shared__ float part[GV_WARPS][NC];
    const int warp = threadIdx.x >> 5, lane = WIDE ? (int) threadIdx.x : (int) (threadIdx.x & 31);
    const int row  = WIDE ? (int) blockIdx.x : (int) (blockIdx.x * GV_WARPS + warp);
    if (row >= p.m) return;
    for (int b = blockIdx.y; b < p.ne12 * p.ne13; b += gridDim.y) {
        const uint8_t *w, *act;
        uint8_t *      dst;
        batch_ptrs(p, b, w, act, dst);
        const E * wr = (const E *) (w + (long long) row * p.w_nb1);
        float     acc[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[c] = 0.0f;
        int k0 = 0;
        if (p.vec) {
            const int nv = p.k / VEC;
            for (int v = lane; v < nv; v += LANES) {
                const uint4 raw = __ldg((const uint4 *) wr + v);
                const E *   e   = (const E *) &raw;
                float       wv[VEC];
#pragma unroll
                for (int i = 0; i < VEC; ++i) wv[i] = Elem<T>::cvt(e[i]);
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    if (c < p.n) {
                        const float4 * ar = (const float4 *) (act + (long long) c * p.a_nb1) + (long long) v * (VEC / 4);
#pragma unroll
                        for (int q = 0; q < VEC / 4; ++q) {
                            const float4 x = __ldg(ar + q);
                            acc[c] = fmaf(wv[4 * q + 0], x.x, acc[c]);
                            acc[c] = fmaf(wv[4 * q + 1], x.y, acc[c]);
                            acc[c] = fmaf(wv[4 * q + 2], x.z, acc[c]);
                            acc[c] = fmaf(wv[4 * q + 3], x.w, acc[c]);
                        }
                    }
                }
            }
            k0 = nv * VEC;
        }
        for (int i = k0 + lane; i < p.k; i += LANES) {
            const float wv = Elem<T>::cvt(wr[i]);
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                if (c < p.n) acc[c] = fmaf(wv, *((const float *) (act + (long long) c * p.a_nb1) + i), acc[c]);
            }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            float s = acc[c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (WIDE) {
                if ((threadIdx.x & 31) == 0) part[warp][c] = s;
            } else if (lane == 0 && c < p.n) {
                *((float *) (dst + (long long) c * p.d_nb1) + row) = s;
            }
        }
        if (WIDE) {
            __syncthreads();
            if (threadIdx.x < NC && (int) threadIdx.x < p.n) {
                float s = part[0][threadIdx.x];
#pragma unroll
                for (int q = 1; q < GV_WARPS; ++q) s += part[q][threadIdx.x];
                *((float *) (dst + (long long) threadIdx.x * p.d_nb1) + row) = s;
            }
            __syncthreads();
        }
    }
}

constexpr int TM = 64, TN = 64, TK = 16;

template <int T> __global__ void __launch_bounds__(256) dense_tile_kernel(const DenseParams p) {
    pdl_grid_sync();
    using E = typename Elem<T>::type;
    __shared__ float As[TK][TM + 4];
    __shared__ float Bs[TK][TN + 4];
    const int t  = threadIdx.x;
    const int m0 = blockIdx.x * TM, n0 = blockIdx.y * TN;
    const int ty = t >> 4, tx = t & 15;
    const int lr = t >> 2, lk = (t & 3) * 4;  // loader: row / column lr, 4 consecutive k
    for (int b = blockIdx.z; b < p.ne12 * p.ne13; b += gridDim.z) {
        const uint8_t *w, *act;
        uint8_t *      dst;
        batch_ptrs(p, b, w, act, dst);
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
        const E *     wr = (const E *) (w + (long long) min(m0 + lr, p.m - 1) * p.w_nb1);
        const float * ar = (const float *) (act + (long long) min(n0 + lr, p.n - 1) * p.a_nb1);
        const bool    wv = m0 + lr < p.m, av = n0 + lr < p.n;
        for (int kb = 0; kb < p.k; kb += TK) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int kk = kb + lk + i;
                As[lk + i][lr] = (wv && kk < p.k) ? Elem<T>::cvt(wr[kk]) : 0.0f;
                Bs[lk + i][lr] = (av && kk < p.k) ? ar[kk] : 0.0f;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < TK; ++kk) {
                const float4 a4 = *(const float4 *) &As[kk][ty * 4];
                const float4 b4 = *(const float4 *) &Bs[kk][tx * 4];
                const float  a[4] = { a4.x, a4.y, a4.z, a4.w }, bb[4] = { b4.x, b4.y, b4.z, b4.w };
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = n0 + tx * 4 + j;
            if (col >= p.n) continue;
            float * dr = (float *) (dst + (long long) col * p.d_nb1);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int row = m0 + ty * 4 + i;
                if (row < p.m) dr[row] = acc[i][j];
            }
        }
    }
}

template <int T> int launch_t(const DenseParams & p, cudaStream_t st) {
    const int nbatch = p.ne12 * p.ne13;
    if (p.n <= 8) {
        const bool wide = (long long) p.m * nbatch <= 592 && p.k >= 1024;  // fewer rows than 4 warps per SM can take one each
        const dim3 grid((unsigned) (wide ? p.m : (p.m + GV_WARPS - 1) / GV_WARPS), (unsigned) (nbatch < 65535 ? nbatch : 65535));
        const dim3 block(GV_WARPS * 32);
        if (wide) {
            if (p.n == 1) launch_pdl(dense_gemv_kernel<T, 1, true>, grid, block, 0, st, p);
            else if (p.n == 2) launch_pdl(dense_gemv_kernel<T, 2, true>, grid, block, 0, st, p);
            else if (p.n <= 4) launch_pdl(dense_gemv_kernel<T, 4, true>, grid, block, 0, st, p);
            else launch_pdl(dense_gemv_kernel<T, 8, true>, grid, block, 0, st, p);
        } else if (p.n == 1) launch_pdl(dense_gemv_kernel<T, 1, false>, grid, block, 0, st, p);
        else if (p.n == 2) launch_pdl(dense_gemv_kernel<T, 2, false>, grid, block, 0, st, p);
        else if (p.n <= 4) launch_pdl(dense_gemv_kernel<T, 4, false>, grid, block, 0, st, p);
        else launch_pdl(dense_gemv_kernel<T, 8, false>, grid, block, 0, st, p);
    } else {
        const long long gy = (p.n + TN - 1) / TN;
        if (gy > 65535) return B200Q_ENOTSUP;
        const dim3 grid((unsigned) ((p.m + TM - 1) / TM), (unsigned) gy, (unsigned) (nbatch < 65535 ? nbatch : 65535));
        launch_pdl(dense_tile_kernel<T>, dim3(grid), dim3(256), 0, st, p);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : B200Q_ECUDA;
}

}  // namespace

bool dense_mm_supported(int type, int64_t k, int64_t m, int64_t n, int64_t nbatch) {
    return dense_type(type) && k >= 1 && k < (1ll << 31) && m < (1ll << 31) && n < (1ll << 31) && nbatch < (1ll << 31);
}

int dense_mm_launch(const MmqArgs & a, cudaStream_t st) {
    DenseParams p{};
    p.w = a.w; p.act = (const uint8_t *) a.act; p.dst = (uint8_t *) a.dst;
    p.k = (int) a.k; p.m = (int) a.m; p.n = (int) a.n;
    p.ne02 = (int) a.ne02; p.ne03 = (int) a.ne03; p.ne12 = (int) a.ne12; p.ne13 = (int) a.ne13;
    p.w_nb1 = a.w_nb1; p.w_nb2 = a.w_nb2; p.w_nb3 = a.w_nb3;
    p.a_nb1 = a.a_nb1; p.a_nb2 = a.a_nb2; p.a_nb3 = a.a_nb3;
    p.d_nb1 = a.d_nb1; p.d_nb2 = a.d_nb2; p.d_nb3 = a.d_nb3;
    auto al16 = [](long long v) { return (v & 15) == 0; };
    p.vec = al16((long long) (uintptr_t) a.w) && al16(a.w_nb1) && al16(a.w_nb2) && al16(a.w_nb3) && al16((long long) (uintptr_t) a.act) &&
            al16(a.a_nb1) && al16(a.a_nb2) && al16(a.a_nb3);
    switch (a.type) {
        case T_F16:  return launch_t<T_F16>(p, st);
        case T_BF16: return launch_t<T_BF16>(p, st);
        case T_F32W: return launch_t<T_F32W>(p, st);
        default: return B200Q_ENOTSUP;
    }
}

}  // namespace b200q
