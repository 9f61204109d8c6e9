// oracle/ref_driver.cpp — TEST INFRASTRUCTURE ONLY.
//
// A thin extern "C" driver over the UNMODIFIED reference libraries built by oracle/Makefile
// (libggml-base / libggml-cpu from /root/reference/ggml).  It lets tests/ and bench.py's
// cpu_baseline leg run the reference's own CPU implementation of the hot path on arbitrary
// host buffers without writing C++:
//
//   ref_quantize      -> ggml_quantize_chunk            (ggml/src/ggml.c:6946)
//   ref_dequantize    -> type_traits[type].to_float     (ggml/src/ggml.c:586-876, ggml-quants.c)
//   ref_act_quantize  -> CPU type-traits from_float of the vec_dot_type (ggml-cpu/ggml-cpu.c:196-376)
//   ref_vec_dot       -> CPU type-traits vec_dot         (arch/x86/quants.c or ggml-cpu/quants.c)
//   ref_mul_mat       -> ggml_mul_mat graph on the CPU backend  (ggml-cpu/ggml-cpu.c:1202-1394)
//   ref_mul_mat_id    -> ggml_mul_mat_id graph on the CPU backend (ggml-cpu/ggml-cpu.c:1476-1654)
//   ref_get_rows      -> ggml_get_rows graph on the CPU backend (ggml-cpu/ops.cpp:5119-5340)
//
// Nothing in the product links or loads this file.

#include <ggml.h>
#include <ggml-backend.h>
#include <ggml-cpu.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct cpu_session {
    ggml_backend_t backend = nullptr;
    int            threads = 0;
    ggml_backend_t get(int n_threads) {
        if (!backend) {
            backend = ggml_backend_cpu_init();
        }
        if (n_threads <= 0) {
            n_threads = (int) std::thread::hardware_concurrency();
        }
        if (n_threads != threads) {
            ggml_backend_cpu_set_n_threads(backend, n_threads);
            threads = n_threads;
        }
        return backend;
    }
};

cpu_session g_cpu;

struct graph_scope {
    ggml_context * ctx;
    explicit graph_scope(size_t n_tensors = 64) {
        ggml_init_params ip{};
        ip.mem_size   = ggml_tensor_overhead() * n_tensors + ggml_graph_overhead() + 4096;
        ip.mem_buffer = nullptr;
        ip.no_alloc   = true;  // tensors point straight at the caller's host buffers
        ctx           = ggml_init(ip);
    }
    ~graph_scope() { ggml_free(ctx); }
};

double run_graph(ggml_context * ctx, ggml_tensor * out, int n_threads, int repeat) {
    ggml_cgraph * gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, out);
    ggml_backend_t be = g_cpu.get(n_threads);
    double best = 1e30;
    for (int r = 0; r < (repeat < 1 ? 1 : repeat); ++r) {
        auto t0 = std::chrono::steady_clock::now();
        if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) {
            return -1.0;
        }
        auto t1 = std::chrono::steady_clock::now();
        double s = std::chrono::duration<double>(t1 - t0).count();
        if (s < best) best = s;
    }
    return best;
}

}  // namespace

extern "C" {

int ref_n_threads_default(void) { return (int) std::thread::hardware_concurrency(); }

// feature string so a report can say which CPU kernels were compiled in
const char * ref_cpu_features(void) {
    static char buf[256];
    snprintf(buf, sizeof(buf), "avx2=%d fma=%d f16c=%d avx512=%d avx512_vnni=%d avx512_vbmi=%d amx_int8=%d llamafile=1 openmp=0",
             ggml_cpu_has_avx2(), ggml_cpu_has_fma(), ggml_cpu_has_f16c(), ggml_cpu_has_avx512(),
             ggml_cpu_has_avx512_vnni(), ggml_cpu_has_avx512_vbmi(), ggml_cpu_has_amx_int8());
    return buf;
}

int64_t ref_blck_size(int type) { return ggml_blck_size((ggml_type) type); }
size_t  ref_type_size(int type) { return ggml_type_size((ggml_type) type); }
size_t  ref_row_size(int type, int64_t ne) { return ggml_row_size((ggml_type) type, ne); }
const char * ref_type_name(int type) { return ggml_type_name((ggml_type) type); }
int ref_vec_dot_type(int type) { return (int) ggml_get_type_traits_cpu((ggml_type) type)->vec_dot_type; }

// weights: f32 [nrows, n_per_row] -> quantized rows (imatrix = NULL)
size_t ref_quantize(int type, const float * src, void * dst, int64_t nrows, int64_t n_per_row) {
    return ggml_quantize_chunk((ggml_type) type, src, dst, 0, nrows, n_per_row, nullptr);
}

// quantized -> f32 through the type-traits table (the dequantize oracle)
int ref_dequantize(int type, const void * src, float * dst, int64_t n) {
    const ggml_type_traits * tt = ggml_get_type_traits((ggml_type) type);
    if (!tt || !tt->to_float) return -1;
    tt->to_float(src, dst, n);
    return 0;
}

// activation row -> the CPU backend's vec_dot_type for `wtype` (q8_0 / q8_1 / q8_K)
int ref_act_quantize(int wtype, const float * src, void * dst, int64_t n) {
    ggml_cpu_init();  // fp16 lookup tables used by the CPU kernels
    const ggml_type_traits_cpu * tc = ggml_get_type_traits_cpu((ggml_type) wtype);
    if (!tc) return -1;
    const ggml_type_traits_cpu * tv = ggml_get_type_traits_cpu(tc->vec_dot_type);
    if (!tv || !tv->from_float) return -1;
    tv->from_float(src, dst, n);
    return (int) tc->vec_dot_type;
}

// one dot product, exactly as mul_mat_one_chunk calls it (nrc = 1)
int ref_vec_dot(int wtype, int64_t n, const void * w_row, const void * act_q, float * out) {
    ggml_cpu_init();
    const ggml_type_traits_cpu * tc = ggml_get_type_traits_cpu((ggml_type) wtype);
    if (!tc || !tc->vec_dot) return -1;
    tc->vec_dot((int) n, out, 0, w_row, 0, act_q, 0, 1);
    return 0;
}

// dst[n][m] = W[m][k] (quantized) x act[n][k] (f32).  Returns best-of-`repeat` seconds (<0 on error).
double ref_mul_mat(int type, const void * w, const float * act, float * dst, int64_t m, int64_t k, int64_t n,
                   int n_threads, int repeat) {
    graph_scope gs;
    ggml_tensor * a = ggml_new_tensor_2d(gs.ctx, (ggml_type) type, k, m);
    ggml_tensor * b = ggml_new_tensor_2d(gs.ctx, GGML_TYPE_F32, k, n);
    a->data = const_cast<void *>(w);
    b->data = const_cast<float *>(act);
    ggml_tensor * d = ggml_mul_mat(gs.ctx, a, b);
    d->data = dst;
    return run_graph(gs.ctx, d, n_threads, repeat);
}

// batched / broadcast form: W[ne03][ne02][m][k], act[ne13][ne12][n][k] with ne12 % ne02 == 0, ne13 % ne03 == 0
double ref_mul_mat_batched(int type, const void * w, const float * act, float * dst, int64_t m, int64_t k, int64_t n,
                           int64_t ne02, int64_t ne03, int64_t ne12, int64_t ne13, int n_threads, int repeat) {
    graph_scope gs;
    ggml_tensor * a = ggml_new_tensor_4d(gs.ctx, (ggml_type) type, k, m, ne02, ne03);
    ggml_tensor * b = ggml_new_tensor_4d(gs.ctx, GGML_TYPE_F32, k, n, ne12, ne13);
    a->data = const_cast<void *>(w);
    b->data = const_cast<float *>(act);
    ggml_tensor * d = ggml_mul_mat(gs.ctx, a, b);
    d->data = dst;
    return run_graph(gs.ctx, d, n_threads, repeat);
}

// MoE: as[n_expert][m][k], b[n_tokens][nb1][k] (nb1 = n_used or 1), ids[n_tokens][ids_stride] i32 (first n_used used)
// dst[n_tokens][n_used][m]
double ref_mul_mat_id(int type, const void * as, const float * b, const int32_t * ids, float * dst, int64_t m,
                      int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens, int64_t b_ne1,
                      int64_t ids_stride, int n_threads, int repeat) {
    graph_scope gs;
    ggml_tensor * ta = ggml_new_tensor_3d(gs.ctx, (ggml_type) type, k, m, n_expert);
    ggml_tensor * tb = ggml_new_tensor_3d(gs.ctx, GGML_TYPE_F32, k, b_ne1, n_tokens);
    ggml_tensor * ti = ggml_new_tensor_2d(gs.ctx, GGML_TYPE_I32, ids_stride, n_tokens);
    ta->data = const_cast<void *>(as);
    tb->data = const_cast<float *>(b);
    ti->data = const_cast<int32_t *>(ids);
    ggml_tensor * tv = ti;
    if (ids_stride != n_used) {
        tv       = ggml_view_2d(gs.ctx, ti, n_used, n_tokens, ti->nb[1], 0);
        tv->data = ti->data;
    }
    ggml_tensor * d = ggml_mul_mat_id(gs.ctx, ta, tb, tv);
    d->data = dst;
    return run_graph(gs.ctx, d, n_threads, repeat);
}

// GET_ROWS: src[n_rows_total][k] quantized (or f32), rows i32[n_idx] -> dst f32 [n_idx][k]
double ref_get_rows(int type, const void * src, const int32_t * rows, float * dst, int64_t k, int64_t n_rows_total,
                    int64_t n_idx, int n_threads) {
    graph_scope gs;
    ggml_tensor * a = ggml_new_tensor_2d(gs.ctx, (ggml_type) type, k, n_rows_total);
    ggml_tensor * r = ggml_new_tensor_1d(gs.ctx, GGML_TYPE_I32, n_idx);
    a->data = const_cast<void *>(src);
    r->data = const_cast<int32_t *>(rows);
    ggml_tensor * d = ggml_get_rows(gs.ctx, a, r);
    d->data = dst;
    return run_graph(gs.ctx, d, n_threads, 1);
}

}  // extern "C"
