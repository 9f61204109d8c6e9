// tests/plugin_driver.cpp — TEST INFRASTRUCTURE ONLY (built by oracle/Makefile into oracle/_ref/plugin_driver.so).
//
// Drives the ggml backend plugin (libggml-b200.so) IN-PROCESS through the reference's own public API, the way
// llama.cpp does: ggml_backend_load() -> device -> backend -> buffers -> graph_compute / ggml_backend_sched.
// tests/test_gpu_plugin.py loads this with ctypes so that the plugin is exercised (and recorded as loaded) inside
// the pytest process; the reference libraries it links are the unmodified ones of oracle/_ref.
//
//   pd_load             ggml_backend_load                         ggml/src/ggml-backend-reg.cpp:232-276, 362-365
//   pd_mul_mat_seq      direct backend use (test-backend-ops' pattern, tests/test-backend-ops.cpp:1108-1230):
//                       alloc_ctx_tensors + tensor_set + graph_compute + tensor_get
//   pd_sched_mlp        ggml_backend_sched with [B200, CPU] (libllama's pattern, ggml-backend.cpp:852-1443): the ops the
//                       plugin declines stay on the CPU backend, tensors cross backends through set/get/cpy
#include <ggml.h>
#include <ggml-alloc.h>
#include <ggml-backend.h>
#include <ggml-cpu.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace {
ggml_backend_reg_t g_reg = nullptr;

struct ctx_scope {
    ggml_context * ctx;
    explicit ctx_scope(size_t n_tensors) {
        ggml_init_params ip{};
        ip.mem_size   = ggml_tensor_overhead() * n_tensors + ggml_graph_overhead() + 4096;
        ip.mem_buffer = nullptr;
        ip.no_alloc   = true;
        ctx           = ggml_init(ip);
    }
    ~ctx_scope() { ggml_free(ctx); }
};
}  // namespace

extern "C" {

// returns the number of devices the plugin registered (<0: load failed)
int pd_load(const char * path) {
    if (!g_reg) {
        g_reg = ggml_backend_load(path);
        if (!g_reg) return -1;
    }
    return (int) ggml_backend_reg_dev_count(g_reg);
}

const char * pd_device_name(int dev) {
    if (!g_reg || dev < 0 || dev >= (int) ggml_backend_reg_dev_count(g_reg)) return "";
    return ggml_backend_dev_name(ggml_backend_reg_dev_get(g_reg, dev));
}

// address of a registry extension (ggml_backend_split_buffer_type, ...) or NULL
void * pd_proc_address(const char * name) { return g_reg ? ggml_backend_reg_get_proc_address(g_reg, name) : nullptr; }

// One graph with n_count MUL_MAT nodes on the SAME weights and growing column counts n_list[i] (the kernels' workspace
// grows inside a single graph_compute).  act: [max n][k] f32, out: concatenation of [n_i][m] f32.  Returns 0 on success.
int pd_mul_mat_seq(int dev, int type, int64_t m, int64_t k, const void * w, const int64_t * n_list, int n_count,
                   const float * act, float * out) {
    if (!g_reg) return -1;
    ggml_backend_dev_t d  = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_t     be = ggml_backend_dev_init(d, nullptr);
    if (!be) return -2;
    int rc = 0;
    {
        ctx_scope wc(4), cc(4 * n_count + 8);
        ggml_tensor * a = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, m);
        ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, be);
        if (!wbuf) { ggml_backend_free(be); return -3; }
        ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        ggml_backend_tensor_set(a, w, 0, ggml_nbytes(a));
        std::vector<ggml_tensor *> bs, ds;
        ggml_cgraph * gf = ggml_new_graph(cc.ctx);
        for (int i = 0; i < n_count; ++i) {
            ggml_tensor * b = ggml_new_tensor_2d(cc.ctx, GGML_TYPE_F32, k, n_list[i]);
            ggml_tensor * o = ggml_mul_mat(cc.ctx, a, b);
            if (!ggml_backend_supports_op(be, o)) rc = -4;
            ggml_build_forward_expand(gf, o);
            bs.push_back(b);
            ds.push_back(o);
        }
        ggml_backend_buffer_t cbuf = rc == 0 ? ggml_backend_alloc_ctx_tensors(cc.ctx, be) : nullptr;
        if (rc == 0 && !cbuf) rc = -5;
        if (rc == 0) {
            for (int i = 0; i < n_count; ++i) ggml_backend_tensor_set(bs[i], act, 0, ggml_nbytes(bs[i]));
            if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -6;
            size_t off = 0;
            for (int i = 0; i < n_count && rc == 0; ++i) {
                ggml_backend_tensor_get(ds[i], (char *) out + off, 0, ggml_nbytes(ds[i]));
                off += ggml_nbytes(ds[i]);
            }
        }
        if (cbuf) ggml_backend_buffer_free(cbuf);
        ggml_backend_buffer_free(wbuf);
    }
    ggml_backend_free(be);
    return rc;
}

// y[n][k] = W2[k][m] . silu( W1[m][k] . x[n][k] )   through ggml_backend_sched over [B200 dev, CPU]; the weights live
// in the plugin's buffer type, SILU is declined by the plugin (or taken, once it supports it) — either way the scheduler
// must produce the CPU backend's result.  y_cpu = the same graph on the CPU backend alone.  n_splits_out: scheduler
// splits of the mixed run.  Returns 0 on success.
int pd_sched_mlp(int dev, int type, int64_t k, int64_t m, const void * w1, const void * w2, const float * x, int64_t n,
                 float * y, float * y_cpu, int * n_splits_out) {
    if (!g_reg) return -1;
    ggml_backend_dev_t d   = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_t     be  = ggml_backend_dev_init(d, nullptr);
    ggml_backend_t     cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    if (!be || !cpu) return -2;
    int rc = 0;
    for (int pass = 0; pass < 2 && rc == 0; ++pass) {  // 0: [B200, CPU] through the scheduler, 1: CPU alone
        ctx_scope wc(4), cc(16);
        ggml_tensor * a1 = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, m);
        ggml_tensor * a2 = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, m, k);
        ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, pass == 0 ? be : cpu);
        if (!wbuf) { rc = -3; break; }
        ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        ggml_backend_tensor_set(a1, w1, 0, ggml_nbytes(a1));
        ggml_backend_tensor_set(a2, w2, 0, ggml_nbytes(a2));
        ggml_tensor * xin = ggml_new_tensor_2d(cc.ctx, GGML_TYPE_F32, k, n);
        ggml_set_input(xin);
        ggml_tensor * h   = ggml_silu(cc.ctx, ggml_mul_mat(cc.ctx, a1, xin));
        ggml_tensor * out = ggml_mul_mat(cc.ctx, a2, h);
        ggml_set_output(out);
        ggml_cgraph * gf = ggml_new_graph(cc.ctx);
        ggml_build_forward_expand(gf, out);
        ggml_backend_t backends[2] = { be, cpu };
        ggml_backend_sched_t sched = pass == 0 ? ggml_backend_sched_new(backends, nullptr, 2, 64, false, true)
                                               : ggml_backend_sched_new(backends + 1, nullptr, 1, 64, false, true);
        if (!ggml_backend_sched_alloc_graph(sched, gf)) rc = -4;
        if (rc == 0) {
            ggml_backend_tensor_set(xin, x, 0, ggml_nbytes(xin));
            if (ggml_backend_sched_graph_compute(sched, gf) != GGML_STATUS_SUCCESS) rc = -5;
        }
        if (rc == 0) {
            ggml_backend_tensor_get(out, pass == 0 ? y : y_cpu, 0, ggml_nbytes(out));
            if (pass == 0 && n_splits_out) *n_splits_out = ggml_backend_sched_get_n_splits(sched);
        }
        ggml_backend_sched_free(sched);
        ggml_backend_buffer_free(wbuf);
    }
    ggml_backend_free(cpu);
    ggml_backend_free(be);
    return rc;
}

}  // extern "C"
