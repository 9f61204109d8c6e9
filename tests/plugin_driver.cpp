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
//   pd_mul_mat_split    llama's `-sm row` (src/llama-model.cpp:366-393): the split buffer type from get_proc_address, a weight
//                       allocated in it, MUL_MAT on the main device's backend next to the same MUL_MAT on a plain buffer
//   pd_async_roundtrip  events, cpy_tensor_async between two backends (two streams, or two devices), the pinned host
//                       buffer type (ggml-backend.cpp:327-376, 432-470; llama's input staging)
#include <ggml.h>
#include <ggml-alloc.h>
#include <ggml-backend.h>
#include <ggml-cpu.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <cmath>
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

// The same MUL_MAT twice in one graph: weights in the row-split buffer type (tensor_split: MAX 16 proportions, NULL = equal)
// and in the device's ordinary buffer.  out_split / out_plain: [n][m] f32 each.  Returns 0 on success, 100 + i when the
// bytes read back from the split tensor differ from w.
int pd_mul_mat_split(int dev, const float * tensor_split, int type, int64_t m, int64_t k, const void * w, int64_t n,
                     const float * act, float * out_split, float * out_plain) {
    if (!g_reg) return -1;
    typedef ggml_backend_buffer_type_t (*split_fn)(int, const float *);
    split_fn get_split = (split_fn) ggml_backend_reg_get_proc_address(g_reg, "ggml_backend_split_buffer_type");
    if (!get_split) return -2;
    ggml_backend_dev_t         d     = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_buffer_type_t sbuft = get_split(dev, tensor_split);
    if (!sbuft) return -3;
    if (!ggml_backend_dev_supports_buft(d, sbuft)) return -4;
    ggml_backend_t be = ggml_backend_dev_init(d, nullptr);
    if (!be) return -5;
    int rc = 0;
    {
        ctx_scope sc(4), wc(4), cc(16);
        ggml_tensor * as = ggml_new_tensor_2d(sc.ctx, (ggml_type) type, k, m);
        ggml_tensor * ap = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, m);
        ggml_backend_buffer_t sbuf = ggml_backend_alloc_ctx_tensors_from_buft(sc.ctx, sbuft);
        ggml_backend_buffer_t pbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, be);
        if (!sbuf || !pbuf) rc = -6;
        ggml_backend_buffer_t cbuf = nullptr;
        if (rc == 0) {
            ggml_backend_buffer_set_usage(sbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
            ggml_backend_buffer_set_usage(pbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
            ggml_backend_tensor_set(as, w, 0, ggml_nbytes(as));
            ggml_backend_tensor_set(ap, w, 0, ggml_nbytes(ap));
            std::vector<uint8_t> back(ggml_nbytes(as));
            ggml_backend_tensor_get(as, back.data(), 0, back.size());
            if (memcmp(back.data(), w, back.size()) != 0) rc = 100;
        }
        if (rc == 0) {
            ggml_tensor * b  = ggml_new_tensor_2d(cc.ctx, GGML_TYPE_F32, k, n);
            ggml_tensor * os = ggml_mul_mat(cc.ctx, as, b);
            ggml_tensor * op = ggml_mul_mat(cc.ctx, ap, b);
            if (!ggml_backend_supports_op(be, os) || !ggml_backend_supports_op(be, op)) rc = -7;
            ggml_cgraph * gf = ggml_new_graph(cc.ctx);
            ggml_build_forward_expand(gf, os);
            ggml_build_forward_expand(gf, op);
            cbuf = rc == 0 ? ggml_backend_alloc_ctx_tensors(cc.ctx, be) : nullptr;
            if (rc == 0 && !cbuf) rc = -8;
            if (rc == 0) {
                ggml_backend_tensor_set(b, act, 0, ggml_nbytes(b));
                for (int rep = 0; rep < 2 && rc == 0; ++rep) {  // twice: the second pass reuses the peers' streams and buffers
                    if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -9;
                }
                if (rc == 0) {
                    ggml_backend_tensor_get(os, out_split, 0, ggml_nbytes(os));
                    ggml_backend_tensor_get(op, out_plain, 0, ggml_nbytes(op));
                }
            }
        }
        if (cbuf) ggml_backend_buffer_free(cbuf);
        if (pbuf) ggml_backend_buffer_free(pbuf);
        if (sbuf) ggml_backend_buffer_free(sbuf);
    }
    ggml_backend_free(be);
    return rc;
}

// x (host, n floats) -> pinned host buffer -> backend A (set_tensor_async) -> backend B on dev_b (cpy_tensor_async, ordered by
// the plugin's events) -> y (get_tensor_async + an event the host waits on).  Returns 0 on success.
int pd_async_roundtrip(int dev_a, int dev_b, const float * x, int64_t n, float * y) {
    if (!g_reg) return -1;
    ggml_backend_dev_t da = ggml_backend_reg_dev_get(g_reg, dev_a), db = ggml_backend_reg_dev_get(g_reg, dev_b);
    ggml_backend_t     ba = ggml_backend_dev_init(da, nullptr), bb = ggml_backend_dev_init(db, nullptr);
    if (!ba || !bb) return -2;
    int rc = 0;
    ggml_backend_buffer_type_t hbuft = ggml_backend_dev_host_buffer_type(da);
    if (!hbuft || !ggml_backend_buft_is_host(hbuft)) rc = -3;
    ggml_backend_dev_props props;
    ggml_backend_dev_get_props(da, &props);
    if (rc == 0 && (!props.caps.async || !props.caps.events || !props.caps.host_buffer)) rc = -4;
    if (rc == 0) {
        ctx_scope ca(2), cb(2), ch(2);
        ggml_tensor * ta = ggml_new_tensor_1d(ca.ctx, GGML_TYPE_F32, n);
        ggml_tensor * tb = ggml_new_tensor_1d(cb.ctx, GGML_TYPE_F32, n);
        ggml_tensor * th = ggml_new_tensor_1d(ch.ctx, GGML_TYPE_F32, n);
        ggml_backend_buffer_t bufa = ggml_backend_alloc_ctx_tensors(ca.ctx, ba);
        ggml_backend_buffer_t bufb = ggml_backend_alloc_ctx_tensors(cb.ctx, bb);
        ggml_backend_buffer_t bufh = ggml_backend_alloc_ctx_tensors_from_buft(ch.ctx, hbuft);
        if (!bufa || !bufb || !bufh) rc = -5;
        if (rc == 0) {
            memcpy(th->data, x, (size_t) n * sizeof(float));  // a host buffer: tensor->data is host memory
            ggml_backend_event_t ev = ggml_backend_event_new(db);
            if (!ev) rc = -6;
            if (rc == 0) {
                ggml_backend_tensor_set_async(ba, ta, th->data, 0, ggml_nbytes(ta));
                ggml_backend_tensor_copy_async(ba, bb, ta, tb);      // ordered after the upload on A, before the read on B
                ggml_backend_tensor_get_async(bb, tb, y, 0, ggml_nbytes(tb));
                ggml_backend_event_record(ev, bb);
                ggml_backend_event_synchronize(ev);
                // a second stream waits for the event, too (event_wait)
                ggml_backend_event_wait(ba, ev);
                ggml_backend_synchronize(ba);
                ggml_backend_event_free(ev);
            }
        }
        if (bufh) ggml_backend_buffer_free(bufh);
        if (bufb) ggml_backend_buffer_free(bufb);
        if (bufa) ggml_backend_buffer_free(bufa);
    }
    ggml_backend_free(bb);
    ggml_backend_free(ba);
    return rc;
}

// a large pageable upload (the load path's pinned double buffering) read back: returns 0 when the bytes match
int pd_big_upload(int dev, const uint8_t * src, int64_t nbytes, uint8_t * back) {
    if (!g_reg) return -1;
    ggml_backend_dev_t d  = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_t     be = ggml_backend_dev_init(d, nullptr);
    if (!be) return -2;
    int rc = 0;
    {
        ctx_scope c(2);
        ggml_tensor * t = ggml_new_tensor_1d(c.ctx, GGML_TYPE_I8, nbytes);
        ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(c.ctx, be);
        if (!buf) rc = -3;
        if (rc == 0) {
            ggml_backend_tensor_set(t, src, 0, (size_t) nbytes);
            ggml_backend_tensor_get(t, back, 0, (size_t) nbytes);
            if (memcmp(src, back, (size_t) nbytes) != 0) rc = 1;
        }
        if (buf) ggml_backend_buffer_free(buf);
    }
    ggml_backend_free(be);
    return rc;
}

// A decode-shaped graph on immutable weights: q = W1 x, k = W2 x, v = W3 x (one shared vector, W3 of type_v), o = W4 q (reads
// what the first node wrote).  The plugin runs the four nodes as ONE persistent list (or one by one with GGML_B200_NO_LIST).
// All matrices are [k][k]; x: k floats; out: 4 x k floats (q, k, v, o).  Returns 0 on success.
int pd_decode_run(int dev, int type, int type_v, int64_t k, const void * w1, const void * w2, const void * w3, const void * w4,
                  const float * x, float * out) {
    if (!g_reg) return -1;
    ggml_backend_dev_t d  = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_t     be = ggml_backend_dev_init(d, nullptr);
    if (!be) return -2;
    int rc = 0;
    {
        ctx_scope wc(8), cc(16);
        ggml_tensor * a1 = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, k);
        ggml_tensor * a2 = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, k);
        ggml_tensor * a3 = ggml_new_tensor_2d(wc.ctx, (ggml_type) type_v, k, k);
        ggml_tensor * a4 = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, k);
        ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, be);
        if (!wbuf) { ggml_backend_free(be); return -3; }
        ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        ggml_backend_tensor_set(a1, w1, 0, ggml_nbytes(a1));
        ggml_backend_tensor_set(a2, w2, 0, ggml_nbytes(a2));
        ggml_backend_tensor_set(a3, w3, 0, ggml_nbytes(a3));
        ggml_backend_tensor_set(a4, w4, 0, ggml_nbytes(a4));
        ggml_tensor * xin = ggml_new_tensor_1d(cc.ctx, GGML_TYPE_F32, k);
        ggml_tensor * q = ggml_mul_mat(cc.ctx, a1, xin);
        ggml_tensor * kk = ggml_mul_mat(cc.ctx, a2, xin);
        ggml_tensor * v = ggml_mul_mat(cc.ctx, a3, xin);
        ggml_tensor * o = ggml_mul_mat(cc.ctx, a4, q);
        ggml_cgraph * gf = ggml_new_graph(cc.ctx);
        ggml_build_forward_expand(gf, q);
        ggml_build_forward_expand(gf, kk);
        ggml_build_forward_expand(gf, v);
        ggml_build_forward_expand(gf, o);
        ggml_backend_buffer_t cbuf = ggml_backend_alloc_ctx_tensors(cc.ctx, be);
        if (!cbuf) rc = -4;
        if (rc == 0) {
            ggml_backend_tensor_set(xin, x, 0, ggml_nbytes(xin));
            for (int rep = 0; rep < 3 && rc == 0; ++rep) {  // the second and third pass hit the plugin's list cache
                if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -5;
            }
            ggml_tensor * outs[4] = { q, kk, v, o };
            for (int i = 0; i < 4 && rc == 0; ++i) ggml_backend_tensor_get(outs[i], out + i * k, 0, (size_t) k * sizeof(float));
        }
        if (cbuf) ggml_backend_buffer_free(cbuf);
        ggml_backend_buffer_free(wbuf);
    }
    ggml_backend_free(be);
    return rc;
}

// llama.cpp picks a weight's buffer type by asking supports_op for the operator the weight is used in (src/llama-model.cpp:152-300):
// with `-sm row` a norm weight (MUL) or an embedding (GET_ROWS) placed in the split buffer type must be declined — only MUL_MAT
// can read row slices (ggml-cuda.cu:3265-3283).  Returns a bit mask of wrongly accepted operators (0 = all declined), < 0 on error.
int pd_split_supports(int dev) {
    if (!g_reg) return -1;
    typedef ggml_backend_buffer_type_t (*split_fn)(int, const float *);
    split_fn get_split = (split_fn) ggml_backend_reg_get_proc_address(g_reg, "ggml_backend_split_buffer_type");
    if (!get_split) return -2;
    ggml_backend_dev_t         d     = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_buffer_type_t sbuft = get_split(dev, nullptr);
    ggml_backend_t             be    = ggml_backend_dev_init(d, nullptr);
    if (!sbuft || !be) return -3;
    int bad = 0;
    {
        ctx_scope sc(4), cc(16);
        ggml_tensor * w1 = ggml_new_tensor_1d(sc.ctx, GGML_TYPE_F32, 256);
        ggml_tensor * w2 = ggml_new_tensor_2d(sc.ctx, GGML_TYPE_Q8_0, 256, 512);
        ggml_backend_buffer_t sbuf = ggml_backend_alloc_ctx_tensors_from_buft(sc.ctx, sbuft);
        if (!sbuf) { ggml_backend_free(be); return -4; }
        ggml_tensor * x   = ggml_new_tensor_2d(cc.ctx, GGML_TYPE_F32, 256, 4);
        ggml_tensor * idx = ggml_new_tensor_1d(cc.ctx, GGML_TYPE_I32, 4);
        ggml_backend_buffer_t cbuf = ggml_backend_alloc_ctx_tensors(cc.ctx, be);
        if (ggml_backend_supports_op(be, ggml_mul(cc.ctx, x, w1))) bad |= 1;
        if (ggml_backend_supports_op(be, ggml_add(cc.ctx, x, w1))) bad |= 2;
        if (ggml_backend_supports_op(be, ggml_get_rows(cc.ctx, w2, idx))) bad |= 4;
        if (!ggml_backend_supports_op(be, ggml_mul_mat(cc.ctx, w2, x))) bad |= 8;  // the one operator that must be accepted
        if (cbuf) ggml_backend_buffer_free(cbuf);
        ggml_backend_buffer_free(sbuf);
    }
    ggml_backend_free(be);
    return bad;
}

// A layer-like graph (rms_norm -> mul_mat -> silu -> mul_mat -> add, `layers` times: >= 16 nodes, so the plugin records it as a
// CUDA graph on the second pass and replays it from the third) computed `reps` times on inputs x * (rep + 1); out [reps][k].
// Results must not depend on whether a pass ran eagerly, was being recorded, or was a replay.
// One Llama decoder layer exactly as llama.cpp builds it without flash attention (src/llama-graph.cpp build_attn_mha,
// src/llama-kv-cache.cpp cpy_k / cpy_v with a non-transposed V cache, src/models llama: norm -> q/k/v -> rope -> KV store ->
// K^T q -> soft_max(mask, scale) -> V kq -> wo -> residual -> norm -> swiglu ffn -> residual), run once on the plugin and once
// on the reference's CPU backend with the same synthetic weights, inputs and pre-filled cache.  Returns 0 and fills
// *nmse (plugin vs CPU on the layer output) and *unsupported (nodes the plugin's supports_op declines: must be 0 for the
// layer to stay on the device).  type: weight type of the seven matrices (k-quants need n_embd % 256 == 0).
namespace {
struct lcg { uint32_t s; float next() { s = s * 1664525u + 1013904223u; return ((s >> 8) * (1.0f / 8388608.0f)) - 1.0f; } };
struct layer_dims { int64_t n_embd = 512, n_head = 8, n_head_kv = 2, hd = 64, n_ff = 1024, n_ctx = 256; };

bool g_layer_vtrans = false;  // the V cache transposed as llama.cpp keeps it without flash attention (llama-kv-cache.cpp cpy_v / get_v,
                              // v_trans): element e of position p at (e * n_ctx + p); written by a SET_ROWS of one-element rows
bool g_layer_bias = false;  // q / k / v / o biases as gpt-oss and Qwen2 have them (an ADD behind each of the four matmuls)
int run_layer(ggml_backend_t be, int type, int n_tokens, int n_past, std::vector<float> & out, int * unsupported) {
    const layer_dims D;
    const int64_t n_kv   = ((n_past + n_tokens + 31) / 32) * 32;
    const int64_t kv_dim = D.hd * D.n_head_kv;
    ctx_scope wc(24), cc(112);
    const ggml_type wt = (ggml_type) type;
    ggml_tensor * wq = ggml_new_tensor_2d(wc.ctx, wt, D.n_embd, D.n_embd);
    ggml_tensor * wk = ggml_new_tensor_2d(wc.ctx, wt, D.n_embd, kv_dim);
    ggml_tensor * wv = ggml_new_tensor_2d(wc.ctx, wt, D.n_embd, kv_dim);
    ggml_tensor * wo = ggml_new_tensor_2d(wc.ctx, wt, D.n_embd, D.n_embd);
    ggml_tensor * wg = ggml_new_tensor_2d(wc.ctx, wt, D.n_embd, D.n_ff);
    ggml_tensor * wu = ggml_new_tensor_2d(wc.ctx, wt, D.n_embd, D.n_ff);
    ggml_tensor * wd = ggml_new_tensor_2d(wc.ctx, wt, D.n_ff, D.n_embd);
    ggml_tensor * an = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, D.n_embd);
    ggml_tensor * fn = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, D.n_embd);
    ggml_tensor * bq = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, D.n_embd);
    ggml_tensor * bk = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, kv_dim);
    ggml_tensor * bv = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, kv_dim);
    ggml_tensor * bo = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, D.n_embd);
    ctx_scope kvc(4);  // the KV cache lives in its own buffer: the weights' buffer is marked immutable, as llama.cpp's is
    ggml_tensor * kc = ggml_new_tensor_2d(kvc.ctx, GGML_TYPE_F16, kv_dim, D.n_ctx);
    ggml_tensor * vc = ggml_new_tensor_2d(kvc.ctx, GGML_TYPE_F16, kv_dim, D.n_ctx);
    ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, be);
    if (!wbuf) return -3;
    ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
    ggml_backend_buffer_t kvbuf = ggml_backend_alloc_ctx_tensors(kvc.ctx, be);
    if (!kvbuf) { ggml_backend_buffer_free(wbuf); return -3; }
    struct kv_guard { ggml_backend_buffer_t b; ~kv_guard() { ggml_backend_buffer_free(b); } } kv_free{ kvbuf };
    lcg rng{ 12345u };
    auto fill_q = [&](ggml_tensor * t, float amp) {
        const int64_t n = ggml_nelements(t);
        std::vector<float> f((size_t) n);
        for (auto & v : f) v = rng.next() * amp;
        std::vector<uint8_t> q(ggml_nbytes(t));
        if (t->type == GGML_TYPE_F32) memcpy(q.data(), f.data(), q.size());
        else ggml_quantize_chunk(t->type, f.data(), q.data(), 0, n / t->ne[0], t->ne[0], nullptr);
        ggml_backend_tensor_set(t, q.data(), 0, q.size());
    };
    for (ggml_tensor * t : { wq, wk, wv, wo, wg, wu, wd }) fill_q(t, 0.06f);
    for (ggml_tensor * t : { an, fn }) fill_q(t, 1.0f);
    for (ggml_tensor * t : { bq, bk, bv, bo }) fill_q(t, 0.3f);
    for (ggml_tensor * t : { kc, vc }) fill_q(t, 0.5f);  // positions < n_past hold earlier tokens; the rest is masked out

    ggml_tensor * x    = ggml_new_tensor_2d(cc.ctx, GGML_TYPE_F32, D.n_embd, n_tokens);
    ggml_tensor * pos  = ggml_new_tensor_1d(cc.ctx, GGML_TYPE_I32, n_tokens);
    ggml_tensor * idx  = ggml_new_tensor_1d(cc.ctx, GGML_TYPE_I64, n_tokens);
    ggml_tensor * mask = ggml_new_tensor_2d(cc.ctx, GGML_TYPE_F32, n_kv, n_tokens);
    ggml_tensor * idxv = ggml_new_tensor_1d(cc.ctx, GGML_TYPE_I64, g_layer_vtrans ? kv_dim * n_tokens : 1);
    ggml_context * c = cc.ctx;
    ggml_tensor * cur = ggml_mul(c, ggml_rms_norm(c, x, 1e-5f), an);
    auto biased = [&](ggml_tensor * t, ggml_tensor * b) { return g_layer_bias ? ggml_add(c, t, b) : t; };
    ggml_tensor * Q = ggml_reshape_3d(c, biased(ggml_mul_mat(c, wq, cur), bq), D.hd, D.n_head, n_tokens);
    ggml_tensor * K = ggml_reshape_3d(c, biased(ggml_mul_mat(c, wk, cur), bk), D.hd, D.n_head_kv, n_tokens);
    ggml_tensor * V = biased(ggml_mul_mat(c, wv, cur), bv);
    Q = ggml_rope_ext(c, Q, pos, nullptr, (int) D.hd, 0, 8192, 500000.0f, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
    K = ggml_rope_ext(c, K, pos, nullptr, (int) D.hd, 0, 8192, 500000.0f, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
    ggml_tensor * k_store = ggml_set_rows(c, kc, ggml_reshape_2d(c, K, kv_dim, n_tokens), idx);
    ggml_tensor * v_store = g_layer_vtrans ? ggml_set_rows(c, ggml_reshape_2d(c, vc, 1, ggml_nelements(vc)), ggml_reshape_2d(c, V, 1, ggml_nelements(V)), idxv)
                                           : ggml_set_rows(c, vc, V, idx);
    ggml_tensor * q = ggml_permute(c, Q, 0, 2, 1, 3);
    ggml_tensor * k = ggml_view_3d(c, kc, D.hd, n_kv, D.n_head_kv, kc->nb[1], ggml_row_size(kc->type, D.hd), 0);
    ggml_tensor * v = ggml_view_3d(c, vc, D.hd, n_kv, D.n_head_kv, vc->nb[1], ggml_row_size(vc->type, D.hd), 0);
    ggml_tensor * kq = ggml_mul_mat(c, k, q);
    kq = ggml_soft_max_ext(c, kq, mask, 1.0f / 8.0f, 0.0f);
    if (g_layer_vtrans) v = ggml_view_3d(c, vc, n_kv, D.hd, D.n_head_kv, ggml_row_size(vc->type, D.n_ctx), ggml_row_size(vc->type, D.n_ctx * D.hd), 0);
    else v = ggml_cont(c, ggml_transpose(c, v));
    ggml_tensor * kqv = ggml_mul_mat(c, v, kq);
    cur = ggml_cont_2d(c, ggml_permute(c, kqv, 0, 2, 1, 3), D.n_embd, n_tokens);
    ggml_tensor * ffn_inp = ggml_add(c, biased(ggml_mul_mat(c, wo, cur), bo), x);
    cur = ggml_mul(c, ggml_rms_norm(c, ffn_inp, 1e-5f), fn);
    cur = ggml_swiglu_split(c, ggml_mul_mat(c, wg, cur), ggml_mul_mat(c, wu, cur));
    ggml_tensor * res = ggml_add(c, ggml_mul_mat(c, wd, cur), ffn_inp);
    ggml_cgraph * gf = ggml_new_graph(c);
    // node order as llama.cpp produces it (build_attn: q_cur, k_cur, v_cur are expanded first, then the cache writes,
    // src/llama-graph.cpp:1451-1463): q, rope(q), k, rope(k), v, set_rows k, set_rows v — the order the plugin's list hoisting,
    // scratch results and SET_ROWS pair are written for
    ggml_build_forward_expand(gf, Q);
    ggml_build_forward_expand(gf, K);
    ggml_build_forward_expand(gf, V);
    ggml_build_forward_expand(gf, k_store);
    ggml_build_forward_expand(gf, v_store);
    ggml_build_forward_expand(gf, res);
    // intermediate results are placed by ggml-alloc, as in llama.cpp: buffers are REUSED along the graph (a matmul's result may
    // take the place of a tensor that died a few nodes earlier), which is what the plugin's reordering has to respect
    for (ggml_tensor * t : { x, pos, idx, mask, idxv }) ggml_set_input(t);
    ggml_set_output(res);
    ggml_gallocr_t ga = ggml_gallocr_new(ggml_backend_get_default_buffer_type(be));
    int rc = ga && ggml_gallocr_alloc_graph(ga, gf) ? 0 : -4;
    if (rc == 0) {
        if (unsupported) {
            *unsupported = 0;
            for (int i = 0; i < ggml_graph_n_nodes(gf); ++i) {
                if (!ggml_backend_supports_op(be, ggml_graph_node(gf, i))) {
                    ++*unsupported;
                    fprintf(stderr, "pd_llama_layer: not supported: %s\n", ggml_op_desc(ggml_graph_node(gf, i)));
                }
            }
        }
        std::vector<float> xs((size_t) (D.n_embd * n_tokens));
        for (auto & f : xs) f = rng.next();
        std::vector<int32_t> ps((size_t) n_tokens);
        std::vector<int64_t> is((size_t) n_tokens);
        std::vector<float>   ms((size_t) (n_kv * n_tokens));
        for (int t = 0; t < n_tokens; ++t) {
            ps[(size_t) t] = n_past + t;
            is[(size_t) t] = n_past + t;
            for (int64_t j = 0; j < n_kv; ++j) ms[(size_t) (t * n_kv + j)] = j <= n_past + t ? 0.0f : -INFINITY;
        }
        ggml_backend_tensor_set(x, xs.data(), 0, ggml_nbytes(x));
        ggml_backend_tensor_set(pos, ps.data(), 0, ggml_nbytes(pos));
        ggml_backend_tensor_set(idx, is.data(), 0, ggml_nbytes(idx));
        ggml_backend_tensor_set(mask, ms.data(), 0, ggml_nbytes(mask));
        if (g_layer_vtrans) {
            std::vector<int64_t> iv((size_t) (kv_dim * n_tokens));
            for (int t = 0; t < n_tokens; ++t) for (int64_t e = 0; e < kv_dim; ++e) iv[(size_t) (t * kv_dim + e)] = e * D.n_ctx + n_past + t;
            ggml_backend_tensor_set(idxv, iv.data(), 0, ggml_nbytes(idxv));
        }
        if (!unsupported || *unsupported == 0) {
            if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -5;
            out.resize((size_t) (D.n_embd * n_tokens));
            if (rc == 0) ggml_backend_tensor_get(res, out.data(), 0, ggml_nbytes(res));
        }
    }
    if (ga) ggml_gallocr_free(ga);
    ggml_backend_buffer_free(wbuf);
    return rc;
}
}  // namespace

// The MoE feed-forward block of gpt-oss as src/llama-graph.cpp build_moe_ffn builds it (SOFTMAX_WEIGHT gating): router matmul
// (f32) + bias, top-k = view of the descending ARGSORT, GET_ROWS of the logits by the selection, soft_max over the selected,
// MUL_MAT_ID (up, gate) + ADD_ID biases, SWIGLU_OAI, MUL_MAT_ID (down) + ADD_ID, weighting, and the sum over the used experts.
namespace {
int run_moe(ggml_backend_t be, int n_tokens, std::vector<float> & out, int * unsupported) {
    const int64_t n_embd = 512, n_ff = 512, n_expert = 8, n_used = 4;
    ctx_scope wc(16), cc(96);
    ggml_tensor * gi  = ggml_new_tensor_2d(wc.ctx, GGML_TYPE_F32, n_embd, n_expert);
    ggml_tensor * gib = ggml_new_tensor_1d(wc.ctx, GGML_TYPE_F32, n_expert);
    ggml_tensor * up  = ggml_new_tensor_3d(wc.ctx, GGML_TYPE_MXFP4, n_embd, n_ff, n_expert);
    ggml_tensor * gt  = ggml_new_tensor_3d(wc.ctx, GGML_TYPE_MXFP4, n_embd, n_ff, n_expert);
    ggml_tensor * dn  = ggml_new_tensor_3d(wc.ctx, GGML_TYPE_MXFP4, n_ff, n_embd, n_expert);
    ggml_tensor * upb = ggml_new_tensor_2d(wc.ctx, GGML_TYPE_F32, n_ff, n_expert);
    ggml_tensor * gtb = ggml_new_tensor_2d(wc.ctx, GGML_TYPE_F32, n_ff, n_expert);
    ggml_tensor * dnb = ggml_new_tensor_2d(wc.ctx, GGML_TYPE_F32, n_embd, n_expert);
    ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, be);
    if (!wbuf) return -3;
    lcg rng{ 777u };
    auto fill = [&](ggml_tensor * t, float amp) {
        const int64_t n = ggml_nelements(t);
        std::vector<float> f((size_t) n);
        for (auto & v : f) v = rng.next() * amp;
        std::vector<uint8_t> q(ggml_nbytes(t));
        if (t->type == GGML_TYPE_F32) memcpy(q.data(), f.data(), q.size());
        else ggml_quantize_chunk(t->type, f.data(), q.data(), 0, n / t->ne[0], t->ne[0], nullptr);
        ggml_backend_tensor_set(t, q.data(), 0, q.size());
    };
    fill(gi, 0.5f); fill(gib, 0.1f); fill(up, 0.06f); fill(gt, 0.06f); fill(dn, 0.06f); fill(upb, 0.1f); fill(gtb, 0.1f); fill(dnb, 0.1f);
    ggml_context * c = cc.ctx;
    ggml_tensor * x = ggml_new_tensor_2d(c, GGML_TYPE_F32, n_embd, n_tokens);
    ggml_tensor * logits = ggml_add(c, ggml_mul_mat(c, gi, x), gib);                          // [n_expert, n_tokens]
    ggml_tensor * sel = ggml_top_k(c, logits, (int) n_used);                                  // [n_used, n_tokens] i32 (view)
    ggml_tensor * w = ggml_get_rows(c, ggml_reshape_3d(c, logits, 1, n_expert, n_tokens), sel);  // [1, n_used, n_tokens]
    w = ggml_reshape_3d(c, ggml_soft_max(c, ggml_reshape_2d(c, w, n_used, n_tokens)), 1, n_used, n_tokens);
    ggml_tensor * cur = ggml_reshape_3d(c, x, n_embd, 1, n_tokens);
    ggml_tensor * u = ggml_add_id(c, ggml_mul_mat_id(c, up, cur, sel), upb, sel);             // [n_ff, n_used, n_tokens]
    ggml_tensor * g = ggml_add_id(c, ggml_mul_mat_id(c, gt, cur, sel), gtb, sel);
    ggml_tensor * a = ggml_swiglu_oai(c, g, u, 1.702f, 7.0f);
    ggml_tensor * e = ggml_add_id(c, ggml_mul_mat_id(c, dn, a, sel), dnb, sel);               // [n_embd, n_used, n_tokens]
    e = ggml_mul(c, e, w);
    ggml_cgraph * gf = ggml_new_graph(c);
    ggml_tensor * parts[4];
    for (int i = 0; i < (int) n_used; ++i) {
        parts[i] = ggml_view_2d(c, e, n_embd, n_tokens, e->nb[2], (size_t) i * e->nb[1]);
        ggml_build_forward_expand(gf, parts[i]);
    }
    ggml_tensor * res = parts[0];
    for (int i = 1; i < (int) n_used; ++i) res = ggml_add(c, res, parts[i]);
    ggml_build_forward_expand(gf, res);
    // intermediate results are placed by ggml-alloc, as in llama.cpp: buffers are REUSED along the graph (a matmul's result may
    // take the place of a tensor that died a few nodes earlier), which is what the plugin's reordering has to respect
    for (ggml_tensor * t : { x }) ggml_set_input(t);
    ggml_set_output(res);
    ggml_gallocr_t ga = ggml_gallocr_new(ggml_backend_get_default_buffer_type(be));
    int rc = ga && ggml_gallocr_alloc_graph(ga, gf) ? 0 : -4;
    if (rc == 0) {
        if (unsupported) {
            *unsupported = 0;
            for (int i = 0; i < ggml_graph_n_nodes(gf); ++i) {
                if (!ggml_backend_supports_op(be, ggml_graph_node(gf, i))) {
                    ++*unsupported;
                    fprintf(stderr, "pd_moe_block: not supported: %s\n", ggml_op_desc(ggml_graph_node(gf, i)));
                }
            }
        }
        std::vector<float> xs((size_t) (n_embd * n_tokens));
        for (auto & f : xs) f = rng.next();
        ggml_backend_tensor_set(x, xs.data(), 0, ggml_nbytes(x));
        if (!unsupported || *unsupported == 0) {
            if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -5;
            out.resize((size_t) (n_embd * n_tokens));
            if (rc == 0) ggml_backend_tensor_get(res, out.data(), 0, ggml_nbytes(res));
        }
    }
    if (ga) ggml_gallocr_free(ga);
    ggml_backend_buffer_free(wbuf);
    return rc;
}
}  // namespace

int pd_moe_block(int dev, int n_tokens, double * nmse, int * unsupported) {
    if (!g_reg) return -1;
    ggml_backend_t be  = ggml_backend_dev_init(ggml_backend_reg_dev_get(g_reg, dev), nullptr);
    ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    if (!be || !cpu) return -2;
    std::vector<float> got, want;
    int rc = run_moe(be, n_tokens, got, unsupported);
    if (rc == 0) rc = run_moe(cpu, n_tokens, want, nullptr);
    *nmse = 1.0;
    if (rc == 0 && got.size() == want.size() && !got.empty()) {
        double num = 0.0, den = 0.0;
        for (size_t i = 0; i < got.size(); ++i) {
            const double d = (double) got[i] - (double) want[i];
            num += d * d;
            den += (double) want[i] * (double) want[i];
        }
        *nmse = den > 0.0 ? num / den : 1.0;
    }
    ggml_backend_free(cpu);
    ggml_backend_free(be);
    return rc;
}

int pd_llama_layer(int dev, int type, int n_tokens, int n_past, double * nmse, int * unsupported);
int pd_llama_layer_vtrans(int dev, int type, int n_tokens, int n_past, int bias, double * nmse, int * unsupported) {
    g_layer_vtrans = true;
    g_layer_bias   = bias != 0;
    const int rc = pd_llama_layer(dev, type, n_tokens, n_past, nmse, unsupported);
    g_layer_vtrans = g_layer_bias = false;
    return rc;
}
int pd_llama_layer_bias(int dev, int type, int n_tokens, int n_past, double * nmse, int * unsupported) {
    g_layer_bias = true;
    const int rc = pd_llama_layer(dev, type, n_tokens, n_past, nmse, unsupported);
    g_layer_bias = false;
    return rc;
}
int pd_llama_layer(int dev, int type, int n_tokens, int n_past, double * nmse, int * unsupported) {
    if (!g_reg) return -1;
    ggml_backend_t be  = ggml_backend_dev_init(ggml_backend_reg_dev_get(g_reg, dev), nullptr);
    ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
    if (!be || !cpu) return -2;
    std::vector<float> got, want;
    int rc = run_layer(be, type, n_tokens, n_past, got, unsupported);
    if (rc == 0) rc = run_layer(cpu, type, n_tokens, n_past, want, nullptr);
    if (rc == 0 && got.size() == want.size() && !got.empty()) {
        double num = 0.0, den = 0.0;
        for (size_t i = 0; i < got.size(); ++i) {
            const double d = (double) got[i] - (double) want[i];
            num += d * d;
            den += (double) want[i] * (double) want[i];
        }
        *nmse = den > 0.0 ? num / den : 1.0;
    } else if (rc == 0) {
        *nmse = 1.0;
    }
    ggml_backend_free(cpu);
    ggml_backend_free(be);
    return rc;
}

static double g_last_us = 0.0;
// microseconds per graph_compute + synchronize of the last pd_layer_replay call (mean over passes 4..reps)
double pd_last_us(void) { return g_last_us; }
int pd_layer_replay(int dev, int type, int64_t k, int layers, const void * wa, const void * wb, const float * x, int reps, float * out) {
    if (!g_reg) return -1;
    ggml_backend_dev_t d  = ggml_backend_reg_dev_get(g_reg, dev);
    ggml_backend_t     be = ggml_backend_dev_init(d, nullptr);
    if (!be) return -2;
    int rc = 0;
    {
        ctx_scope wc(4), cc(16 + 8 * layers);
        ggml_tensor * a = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, k);
        ggml_tensor * b = ggml_new_tensor_2d(wc.ctx, (ggml_type) type, k, k);
        ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(wc.ctx, be);
        if (!wbuf) { ggml_backend_free(be); return -3; }
        ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        ggml_backend_tensor_set(a, wa, 0, ggml_nbytes(a));
        ggml_backend_tensor_set(b, wb, 0, ggml_nbytes(b));
        ggml_tensor * xin = ggml_new_tensor_1d(cc.ctx, GGML_TYPE_F32, k);
        ggml_tensor * cur = xin;
        for (int l = 0; l < layers; ++l) {
            ggml_tensor * h = ggml_rms_norm(cc.ctx, cur, 1e-5f);
            h = ggml_mul_mat(cc.ctx, a, h);
            h = ggml_silu(cc.ctx, h);
            h = ggml_mul_mat(cc.ctx, b, h);
            cur = ggml_add(cc.ctx, h, cur);
        }
        ggml_cgraph * gf = ggml_new_graph(cc.ctx);
        ggml_build_forward_expand(gf, cur);
        ggml_backend_buffer_t cbuf = ggml_backend_alloc_ctx_tensors(cc.ctx, be);
        if (!cbuf) rc = -4;
        std::vector<float> xr((size_t) k);
        double     total_us      = 0.0;
        const bool out_last_only = reps > 16;  // timing runs keep one result row
        for (int rep = 0; rep < reps && rc == 0; ++rep) {
            for (int64_t i = 0; i < k; ++i) xr[(size_t) i] = x[i] * (float) (rep % 7 + 1);
            ggml_backend_tensor_set(xin, xr.data(), 0, ggml_nbytes(xin));
            ggml_backend_synchronize(be);
            const auto t0 = std::chrono::steady_clock::now();
            if (ggml_backend_graph_compute(be, gf) != GGML_STATUS_SUCCESS) rc = -5;
            ggml_backend_synchronize(be);
            if (rep >= 3) total_us += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
            if (rc == 0 && out) ggml_backend_tensor_get(cur, out + (size_t) (out_last_only ? 0 : rep) * k, 0, (size_t) k * sizeof(float));
        }
        g_last_us = reps > 3 ? total_us / (reps - 3) : 0.0;
        if (cbuf) ggml_backend_buffer_free(cbuf);
        ggml_backend_buffer_free(wbuf);
    }
    ggml_backend_free(be);
    return rc;
}

}  // extern "C"
