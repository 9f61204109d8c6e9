/*
 * b200q.h — the thin C-ABI of the B200-native quantized matmul kernels (libb200q.so).
 *
 * Plain pointers and sizes only: no ggml, torch or C++ types cross this boundary.  The ggml backend
 * plugin (include/ggml-b200.h, libggml-b200.so) is the C++ host code that sits above it; tests and
 * bench.py bind it with ctypes.
 *
 * Each entry point replaces one piece of the reference's CUDA hot path (paths relative to
 * /root/reference/ggml/src/ggml-cuda/ unless stated):
 *
 *   b200q_dequantize      <- dequantize_block_* / ggml_get_to_fp32_cuda       convert.cu:82-290,661-760
 *                            (oracle: dequantize_row_* ggml-quants.c:307-1800, bit-exact)
 *   b200q_get_rows        <- k_get_rows / ggml_cuda_op_get_rows               getrows.cu:5-67,206-252
 *   b200q_mul_mat         <- ggml_cuda_mul_mat (GEMV: mul_mat_vec_q mmvq.cu:139-226,500-570 +
 *                            quantize_q8_1 quantize.cu:4-48;  MMQ: mul_mat_q mmq.cuh:3134-3370 +
 *                            quantize_mmq_q8_1 quantize.cu:50-146)            ggml-cuda.cu:2002-2085
 *   b200q_mul_mat_id      <- ggml_cuda_mul_mat_id + host routing              ggml-cuda.cu:2087-2228, mmq.cu:145-227
 *   b200q_*_host          <- the set_tensor / graph_compute / get_tensor round trip a ggml caller
 *                            makes (ggml-backend.cpp:327-376) — used for end-to-end timing
 *   b200q_rowsplit_*      <- ggml_cuda_op_mul_mat row-split orchestration     ggml-cuda.cu:1436-1754
 *
 * Tensor conventions are ggml's (ggml/include/ggml.h:614-646): ne[0] is the contiguous dimension,
 * nb[] are byte strides.  `type` values are ggml_type numbers (ggml.h:371-411).
 * All pointers are DEVICE pointers unless the function name ends in _host.
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 * Every function returns 0 on success, a negative B200Q_E* code otherwise; b200q_last_error() gives text.
 * There is no CPU fallback anywhere behind this header: without a CUDA device the calls fail.
 */
#ifndef B200Q_H
#define B200Q_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define B200Q_API
#else
#define B200Q_API __attribute__((visibility("default")))
#endif

/* ggml_type numbers accepted as weight types */
enum b200q_type {
    B200Q_TYPE_F32   = 0,
    B200Q_TYPE_F16   = 1,   /* unquantized weights (k % 64 == 0) and operands of the elementwise operators */
    B200Q_TYPE_BF16  = 30,
    B200Q_TYPE_I32   = 26,  /* index / position tensors of the operators below */
    B200Q_TYPE_I64   = 27,
    B200Q_TYPE_Q4_0  = 2,
    B200Q_TYPE_Q4_1  = 3,
    B200Q_TYPE_Q5_0  = 6,
    B200Q_TYPE_Q5_1  = 7,
    B200Q_TYPE_Q8_0  = 8,
    B200Q_TYPE_Q2_K  = 10,
    B200Q_TYPE_Q3_K  = 11,
    B200Q_TYPE_Q4_K  = 12,
    B200Q_TYPE_Q5_K  = 13,
    B200Q_TYPE_Q6_K  = 14,
    B200Q_TYPE_IQ4_NL = 20,
    B200Q_TYPE_IQ4_XS = 23,
    B200Q_TYPE_MXFP4 = 39,
};

enum b200q_status {
    B200Q_OK            = 0,
    B200Q_EINVAL        = -1, /* bad argument / unsupported layout */
    B200Q_ENOTSUP       = -2, /* type or shape not supported by the kernels */
    B200Q_ECUDA         = -3, /* CUDA runtime error (see b200q_last_error) */
    B200Q_ENOMEM        = -4,
    B200Q_ENODEVICE     = -5,
    B200Q_EWORKSPACE    = -6, /* workspace too small */
};

/* b200q_mul_mat flags */
#define B200Q_FLAG_WEIGHTS_IMMUTABLE 1u /* src0 is never written by stream work => weight prefetch may start before the
                                            previous kernel finishes (programmatic dependent launch) */
#define B200Q_FLAG_FORCE_GEMV        2u /* debugging / measurement: always take the GEMV kernels (8 columns per pass) */
#define B200Q_FLAG_FORCE_MMQ         4u /* debugging / measurement: always take the tensor-core kernel */
#define B200Q_FLAG_NO_PDL            8u /* launch without programmatic stream serialization */
#define B200Q_FLAG_ACT_UNCHANGED    16u /* b200q_mul_mat: `act` holds the same values as in the previous b200q_mul_mat call that used
                                           this workspace on this stream (wq / wk / wv, ffn_gate / ffn_up of one vector): the
                                           tensor-core kernel may reuse the packed image it left in the workspace instead of
                                           packing again; ignored when the previous call's layout differs */

/* ---- library / device ---------------------------------------------------------------------- */
B200Q_API int          b200q_version(void);
B200Q_API const char * b200q_last_error(void);
B200Q_API int          b200q_device_count(void);
/* name may be NULL; any out pointer may be NULL */
B200Q_API int b200q_device_info(int device, char * name, size_t name_len, size_t * total_mem, size_t * free_mem,
                                int * cc_major, int * cc_minor, int * sm_count, size_t * l2_bytes);

/* ---- device memory / streams (thin wrappers so that hosts above this ABI need no CUDA headers; they replace the
 *      cudaMalloc / cudaMemcpyAsync / cudaStream calls of ggml-cuda.cu:551-664, 2525-2618) ------------------------- */
B200Q_API int b200q_set_device(int device);
B200Q_API int b200q_malloc(void ** ptr, size_t bytes);             /* B200Q_ENOMEM when the device is out of memory */
B200Q_API int b200q_free(void * ptr);
B200Q_API int b200q_memset(void * ptr, int value, size_t bytes, void * stream);
#define B200Q_H2D 1
#define B200Q_D2H 2
#define B200Q_D2D 3
B200Q_API int b200q_memcpy(void * dst, const void * src, size_t bytes, int kind, void * stream); /* asynchronous */
B200Q_API int b200q_stream_create(void ** stream);
B200Q_API int b200q_stream_destroy(void * stream);
B200Q_API int b200q_stream_synchronize(void * stream);
B200Q_API int b200q_host_alloc(void ** ptr, size_t bytes);         /* pinned host memory */
B200Q_API int b200q_host_free(void * ptr);
/* events, cross-stream ordering, peer mappings (cudaEvent* / cudaStreamWaitEvent / cudaDeviceEnablePeerAccess /
 * cudaMemcpyPeerAsync of ggml-cuda.cu:1358-1414, 2556-2605, 3038-3060): what the plugin's events, cpy_tensor_async and
 * single-process row split are built from */
/* ---- stream capture: the plugin records a decode step's launches once and replays them as ONE CUDA graph (the reference's
 *      cudaGraph path, ggml-cuda.cu:2876-3100).  begin/end bracket ordinary calls of this library on `stream`; calls that
 *      synchronise or allocate (a workspace that has to grow) are not capturable — run the step once eagerly first. -------- */
B200Q_API int b200q_graph_begin(void * stream);
B200Q_API int b200q_graph_end(void * stream, void ** exec_out);   /* instantiated graph; NULL + error if the capture broke */
B200Q_API int b200q_graph_launch(void * exec, void * stream);
B200Q_API int b200q_graph_destroy(void * exec);
B200Q_API int b200q_event_create(void ** event);                     /* on the current device, timing disabled */
B200Q_API int b200q_event_destroy(void * event);
B200Q_API int b200q_event_record(void * event, void * stream);
B200Q_API int b200q_event_synchronize(void * event);
B200Q_API int b200q_stream_wait_event(void * stream, void * event);  /* event may belong to another device */
B200Q_API int b200q_enable_peer_access(int device, int peer);        /* kernels on `device` may load / store peer's memory;
                                                                         B200Q_ENOTSUP when the link does not allow it */
B200Q_API int b200q_memcpy_peer(void * dst, int dst_device, const void * src, int src_device, size_t bytes, void * stream);

/* ---- type table (mirrors ggml_blck_size / ggml_type_size, ggml.c:586-876) -------------------- */
B200Q_API int64_t b200q_blck_size(int type);
B200Q_API size_t  b200q_type_size(int type);
B200Q_API size_t  b200q_row_size(int type, int64_t ne0);
B200Q_API int     b200q_type_supported(int type); /* 1 if usable as MUL_MAT/GET_ROWS weight type */
B200Q_API int64_t b200q_max_k(int type);          /* longest row every MUL_MAT(_ID) shape of `type` is served at (host-only
                                                      planner probe; what supports_op caps ne[0] with) */

/* ---- dequantize / GET_ROWS (bit-exact vs dequantize_row_*) ----------------------------------- */
/* contiguous: n_elems % blck_size == 0 */
B200Q_API int b200q_dequantize(int type, const void * src, float * dst, int64_t n_elems, void * stream);

typedef struct b200q_get_rows_args {
    int     type;              /* src0 type (quantized or F32) */
    int64_t ne00;              /* row length */
    int64_t ne10, ne11, ne12;  /* index tensor shape; dst = [ne00, ne10, ne11, ne12] */
    size_t  nb01, nb02, nb03;  /* src0 byte strides (row, plane, batch) */
    size_t  nb10, nb11, nb12;  /* ids byte strides */
    size_t  nb1, nb2, nb3;     /* dst byte strides (dst rows contiguous f32) */
} b200q_get_rows_args;
B200Q_API int b200q_get_rows(const b200q_get_rows_args * a, const void * src0, const int32_t * ids, float * dst,
                             void * stream);

/* ---- MUL_MAT --------------------------------------------------------------------------------- */
/* dst[i13][i12][j][i] = sum_l W[i13/r3][i12/r2][i][l] * act[i13][i12][j][l]   (ggml_mul_mat, ggml.c:3015-3030) */
typedef struct b200q_mul_mat_args {
    int      type;             /* weight type */
    int64_t  k, m, n;          /* ne00 (= ne10), ne01, ne11 */
    int64_t  ne02, ne03;       /* weight batch dims */
    int64_t  ne12, ne13;       /* activation batch dims (multiples of ne02 / ne03) */
    size_t   nb01, nb02, nb03; /* weight byte strides */
    size_t   nb11, nb12, nb13; /* activation byte strides (f32, nb10 == 4) */
    size_t   nb1, nb2, nb3;    /* dst byte strides (f32, nb0 == 4) */
    uint32_t flags;
} b200q_mul_mat_args;
B200Q_API size_t b200q_mul_mat_workspace(const b200q_mul_mat_args * a);
B200Q_API int    b200q_mul_mat(const b200q_mul_mat_args * a, const void * w, const float * act, float * dst,
                               void * workspace, size_t workspace_bytes, void * stream);

/* ---- MUL_MAT that allocates nothing.  No counterpart in the reference: added for
 *      callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *      Same arguments, kernels and results as b200q_mul_mat, but the tensor-core kernel's scratch (the packed bf16
 *      activation image, b200q_mul_mat_workspace bytes) is a file-scope device array inside the library of
 *      B200Q_STATIC_WORKSPACE_BYTES (16 MiB; every GPU that loads libb200q.so holds it until exit).  Shapes that need more
 *      fail with B200Q_EWORKSPACE and a message that says so (512 columns fit up to k = 16384; decode shapes need none).
 *      The call makes no cudaMalloc / cudaFree / allocating library call.  All calls share the one array, which keeps the
 *      last call's contents (nothing in it has to start at zero: the pack kernel rewrites every byte the matmul reads):
 *      calls that overlap on the GPU — different streams, or several host threads — race on it; serialise them on one
 *      stream or use b200q_mul_mat with separate workspaces. */
#define B200Q_STATIC_WORKSPACE_BYTES (16u << 20)
B200Q_API int b200q_mul_mat_static(const b200q_mul_mat_args * a, const void * w, const float * act, float * dst,
                                   void * stream);

/* ---- grouped decode MUL_MAT: up to 4 weight matrices [m_g][k] of one type that multiply the SAME activation vector
 *      (n = 1: wq/wk/wv, ffn_gate/ffn_up) in one launch.  Equivalent to n_mats calls of b200q_mul_mat (bit-identical
 *      results); replaces the reference's fused gate/up GEMV (ggml-cuda mmvq.cu, fusion args of mul_mat_vec_q).
 *      Rows dense (row pitch = b200q_row_size), all pointers 16-byte aligned. ------------------------------------ */
#define B200Q_MAX_GROUP 4
typedef struct b200q_mul_mat_group_args {
    int      type;
    int64_t  k;
    int      n_mats;
    int64_t  m[B200Q_MAX_GROUP];
    uint32_t flags;
} b200q_mul_mat_group_args;
B200Q_API int b200q_mul_mat_group(const b200q_mul_mat_group_args * a, const void * const * w, const float * act,
                                  float * const * dst, void * stream);

/* ---- row split over the GPUs of one box, one process per GPU (replaces the split buffer's cudaMemcpyPeerAsync of
 *      the dst slices, ggml-cuda.cu:1700-1740): the decode kernel stores every result row into the dst buffer of each
 *      peer GPU as well (NVLink peer stores in the epilogue), so no collective follows the matmul.  peer_delta[r] =
 *      (this process's mapping of peer r's dst buffer) - (this GPU's own dst buffer), in bytes; the buffers have the
 *      same layout.  The stores are complete when the kernel is: consumers on the peer need a stream-ordered
 *      cross-GPU fence (e.g. one small collective per step).  REUSE RULE: that fence orders the stores before the peers'
 *      reads only; a rank that has passed it may launch the next step, whose epilogues overwrite the same rows of the
 *      peers' buffers — a peer still reading the previous step's rows needs a second (consumer-done) fence before the
 *      buffer is reused, or one dst region per step.  Inside a persistent list (b200q_list_*, below) the rule holds by
 *      construction — every in-list reader of a vector has passed a barrier point before any rank can store into it
 *      again; kernels that read a list's results after the launch fall under the rule above (alternate two dst buffers).
 *      b200q_ipc_* exchange the mappings (cudaIpc). ------------------------------------------------------------------- */
B200Q_API int b200q_mul_mat_group_peers(const b200q_mul_mat_group_args * a, const void * const * w, const float * act,
                                        float * const * dst, int n_peers, const int64_t * peer_delta, void * stream);
B200Q_API int b200q_ipc_export(void * dev_ptr, void * handle64);          /* dev_ptr from b200q_malloc; 64-byte handle */
B200Q_API int b200q_ipc_import(const void * handle64, void ** dev_ptr);   /* maps the peer's buffer, enables peer access */
B200Q_API int b200q_ipc_close(void * dev_ptr);

/* ---- persistent decode list: a whole run of n = 1 matmuls (a decode token's layers) in ONE launch -------------------
 *      Replaces the per-matmul launch pair of the reference's decode path (quantize_q8_1 + mul_mat_vec_q,
 *      ggml-cuda/quantize.cu:148-160, mmvq.cu:500-570, called once per MUL_MAT node by ggml_cuda_compute_forward,
 *      ggml-cuda.cu:2230-2520): the CTAs never exit between matrices, the weight stream continues across matrix
 *      boundaries, and data dependencies are device-side barrier points instead of kernel boundaries.
 *      Weights must be immutable while a list exists and rows dense (pitch = b200q_row_size); w and act 16-byte aligned.
 *      Consecutive ops that read the same `act` share one quantization.  B200Q_LIST_DEP on an op says its `act` is written
 *      by earlier ops of this list (on this GPU or, in a row split, on its peers): the kernel then waits until EVERY
 *      earlier op is complete everywhere before it reads the vector.  A dst must not overlap the `act` of an earlier op
 *      unless a B200Q_LIST_DEP op lies between them.  Results are bit-reproducible and do not depend on how a list is
 *      composed, launched or sharded (fixed summation order per (type, k)). */
#define B200Q_LIST_DEP 1u
/* B200Q_LIST_SWIGLU: the vector this op multiplies is silu(act[i]) * act2[i] (ggml's SWIGLU of two tensors, unary-ops / glu.cu),
 * computed while the activations are quantized: ffn_gate, ffn_up and ffn_down of a layer are ONE launch with one barrier point
 * and the GLU node's own result is never written.  Normally together with B200Q_LIST_DEP (act and act2 are results of earlier
 * ops of the list).  The op's dst must not overlap act or act2.  silu uses the approximate exp and division (2^-21 relative,
 * far below the int8 step of the quantized vector).  Measured: every CTA evaluates the whole vector behind the barrier point,
 * which costs more (4.8 us at k = 14336) than a separate GLU launch — the plugin uses it only with GGML_B200_GLU_LIST=1. */
#define B200Q_LIST_SWIGLU 2u
typedef struct b200q_list_op {
    int           type;
    int64_t       m, k;   /* dst[m] = W[m][k] . act[k];  m may be 0 (a rank without rows of a small matrix) */
    const void *  w;
    const float * act;
    float *       dst;
    uint32_t      flags;
    const float * act2;   /* B200Q_LIST_SWIGLU only, else ignored (NULL) */
} b200q_list_op;
/* row split: every result row is also stored into the same place of each peer's dst buffer (peer_delta[i] = this
 * process's mapping of the i-th OTHER rank's dst buffer minus this GPU's own, ascending rank order), and every barrier
 * point — plus one at the end of every launch — is a cross-GPU one: sync[r] = device pointer to rank r's 256-byte,
 * zero-initialised sync block (own block at sync[rank], peers' mapped with b200q_ipc_import).  All ranks must create and
 * launch the same lists in the same order.  A list must not be launched again before every rank has consumed the dst
 * regions it overwrites (alternate two dst buffers, or fence). */
typedef struct b200q_list_peers {
    int             rank, world;
    const int64_t * peer_delta;  /* [world - 1] */
    void * const *  sync;        /* [world] */
} b200q_list_peers;
typedef struct b200q_list b200q_list;
B200Q_API int b200q_list_create(const b200q_list_op * ops, int n_ops, const b200q_list_peers * peers /* NULL: one GPU */,
                                b200q_list ** out);     /* B200Q_ENOTSUP: some op is not served by the list kernel */
B200Q_API int b200q_list_launch(b200q_list * list, void * stream);   /* asynchronous; capturable in a CUDA graph */
B200Q_API int b200q_list_destroy(b200q_list * list);
/* out[4] = kernel launches per b200q_list_launch, barrier points, activation generations, ring slots of the first launch */
B200Q_API int b200q_list_info(const b200q_list * list, int * out);

/* ---- MUL_MAT_ID (MoE) ------------------------------------------------------------------------ */
/* dst[t][s][:] = W[ids[t][s]] . b[t][s or 0][:]     (ggml_mul_mat_id, ggml.c:3056-3080) */
typedef struct b200q_mul_mat_id_args {
    int      type;
    int64_t  k, m;             /* expert matrix [m][k] */
    int64_t  n_expert;         /* as->ne[2] */
    int64_t  n_used;           /* ids->ne[0] */
    int64_t  n_tokens;         /* ids->ne[1] */
    int64_t  b_ne1;            /* 1 (broadcast) or n_used */
    size_t   nb01, nb02;       /* expert weights byte strides (row, expert) */
    size_t   nb11, nb12;       /* b byte strides (slot, token) */
    size_t   ids_nb1;          /* ids byte stride between tokens */
    size_t   nb1, nb2;         /* dst byte strides (slot, token) */
    uint32_t flags;
} b200q_mul_mat_id_args;
B200Q_API size_t b200q_mul_mat_id_workspace(const b200q_mul_mat_id_args * a);
B200Q_API int    b200q_mul_mat_id(const b200q_mul_mat_id_args * a, const void * as, const float * b,
                                  const int32_t * ids, float * dst, void * workspace, size_t workspace_bytes,
                                  void * stream);

/* MUL_MAT_ID + ADD_ID (dst[t][s][:] += bias[ids[t][s]][:], the experts' bias rows of gpt-oss) in one launch; decode only
 * (n_tokens == 1: the ring kernel's epilogue adds the bias) — B200Q_ENOTSUP otherwise, the caller then runs the two operators.
 * bias f32, bias_nb1 = bytes between the experts' rows. */
B200Q_API int    b200q_mul_mat_id_bias(const b200q_mul_mat_id_args * a, const void * as, const float * b, const int32_t * ids,
                                       const float * bias, size_t bias_nb1, float * dst, void * workspace, size_t workspace_bytes,
                                       void * stream);

/* ---- the operators around the matmuls of a decoder layer: they keep a layer's graph on the device.  Each one
 *      replaces the reference CUDA kernel named beside it (ggml/src/ggml-cuda/) and follows the arithmetic of the CPU operator
 *      (ggml/src/ggml-cpu/ops.cpp); tensors are described ggml's way.  B200Q_ENOTSUP = layout / type the kernel does not take
 *      (the plugin's supports_op declines exactly those). ------------------------------------------------------------------ */
typedef struct b200q_tensor {
    void *  data;   /* device pointer */
    int     type;   /* b200q_type */
    int64_t ne[4];  /* elements per dimension, ne[0] innermost */
    size_t  nb[4];  /* byte strides */
} b200q_tensor;
enum { B200Q_BIN_ADD = 0, B200Q_BIN_SUB, B200Q_BIN_MUL, B200Q_BIN_DIV };
enum { B200Q_UNARY_SILU = 0, B200Q_UNARY_GELU, B200Q_UNARY_GELU_QUICK, B200Q_UNARY_GELU_ERF, B200Q_UNARY_RELU, B200Q_UNARY_SIGMOID,
       B200Q_UNARY_TANH, B200Q_UNARY_NEG, B200Q_UNARY_ABS, B200Q_UNARY_EXP };
enum { B200Q_GLU_REGLU = 0, B200Q_GLU_GEGLU, B200Q_GLU_SWIGLU, B200Q_GLU_GEGLU_ERF, B200Q_GLU_GEGLU_QUICK, B200Q_GLU_SWIGLU_OAI };
enum { B200Q_ROPE_NORMAL = 0, B200Q_ROPE_NEOX = 2 };
typedef struct b200q_rope_params {
    int   n_dims, mode;
    float freq_base, freq_scale, ext_factor, attn_factor;
    float corr_dims[2];  /* ggml_rope_yarn_corr_dims(n_dims, n_ctx_orig, freq_base, beta_fast, beta_slow) — host arithmetic */
} b200q_rope_params;
B200Q_API int b200q_op_rms_norm(const b200q_tensor * src, const b200q_tensor * dst, float eps, void * stream);            /* norm.cu */
/* [ADD +] RMS_NORM + MUL in one launch (the reference fuses RMS_NORM + MUL (+ ADD) too, ggml-cuda.cu:2811-2845): x = a (+ b; the sum
 * is also written to sum_out when given), dst = rms_norm(x, eps) * w with w broadcast over rows; rows of at most 8192 values */
B200Q_API int b200q_op_norm_fused(const b200q_tensor * a, const b200q_tensor * b /* or NULL */, const b200q_tensor * sum_out /* or NULL */,
                                  const b200q_tensor * w, const b200q_tensor * dst, float eps, void * stream);
B200Q_API int b200q_op_bin(int op, const b200q_tensor * src0, const b200q_tensor * src1, const b200q_tensor * dst,
                           void * stream);                                                                                 /* binbcast.cu */
B200Q_API int b200q_op_scale(const b200q_tensor * src, const b200q_tensor * dst, float scale, float bias, void * stream); /* scale.cu */
B200Q_API int b200q_op_unary(int op, const b200q_tensor * src, const b200q_tensor * dst, void * stream);                   /* unary.cu */
/* src1 NULL: gate and value are the two halves of src0's rows (`swapped` picks the order); alpha, limit: SWIGLU_OAI only */
B200Q_API int b200q_op_glu(int op, const b200q_tensor * src0, const b200q_tensor * src1, const b200q_tensor * dst, int swapped,
                           float alpha, float limit, void * stream);                                                       /* unary.cu (glu) */
/* mask (f32 | f16, broadcast over dims 2, 3) and sinks (f32 [n_head]) may be NULL */
B200Q_API int b200q_op_soft_max(const b200q_tensor * src, const b200q_tensor * mask, const b200q_tensor * sinks,
                                const b200q_tensor * dst, float scale, float max_bias, void * stream);                     /* softmax.cu */
/* decode attention core in one launch (what llama.cpp emits without flash attention as MUL_MAT(k, q) -> SOFT_MAX -> MUL_MAT(v, kq) ->
 * PERMUTE -> CONT, llama-graph.cpp:1285-1330): q f32 [hd, n_tok, n_head], k f16 [hd, n_kv, n_head_kv], v f16 [n_kv, hd, n_head_kv]
 * (the transposed V cache), mask f32 | f16 [n_kv, >= n_tok, 1 | n_head] or NULL, sinks f32 [n_head] or NULL, dst f32 [hd, n_head, n_tok];
 * n_kv <= 8192, meant for a handful of tokens */
B200Q_API int b200q_op_attn_decode(const b200q_tensor * q, const b200q_tensor * k, const b200q_tensor * v, const b200q_tensor * mask,
                                   const b200q_tensor * sinks, const b200q_tensor * dst, float scale, void * stream);
B200Q_API int b200q_op_rope(const b200q_tensor * src, const b200q_tensor * pos, const b200q_tensor * freq_factors /* or NULL */,
                            const b200q_tensor * dst, const b200q_rope_params * p, void * stream);                         /* rope.cu */
B200Q_API int b200q_op_cpy(const b200q_tensor * src, const b200q_tensor * dst, void * stream);       /* cpy.cu: CPY / CONT / DUP */
B200Q_API int b200q_op_add_id(const b200q_tensor * a, const b200q_tensor * b, const b200q_tensor * ids, const b200q_tensor * dst,
                              void * stream);  /* dst[i0,i1,i2] = a[i0,i1,i2] + b[i0, ids[i1,i2]] (MoE expert bias)      add-id.cu */
B200Q_API int b200q_op_argsort(const b200q_tensor * src, const b200q_tensor * dst, int descending,
                               void * stream);  /* f32 rows (ne0 <= 1024) -> i32 indices; ggml_top_k is a view of it   argsort.cu */
B200Q_API int b200q_op_set_rows(const b200q_tensor * src, const b200q_tensor * idx, const b200q_tensor * dst,
                                void * stream);                                                                            /* set-rows.cu */

/* MoE combine in one launch: dst[i0, t] = sum_s experts[i0, s, t] * weights[0, s, t] (llama.cpp emits MUL + one VIEW per used expert + a
 * chain of ADDs, src/llama-graph.cpp build_moe_ffn); products rounded separately and added in slot order, as the separate operators do */
B200Q_API int b200q_op_moe_weighted_sum(const b200q_tensor * experts, const b200q_tensor * weights, const b200q_tensor * dst, void * stream);

/* the MoE router behind the logits matmul, as gpt-oss gates (src/llama-graph.cpp build_moe_ffn: router bias, top-k, soft_max over
 * the selected logits): biased = logits + bias; sorted = argsort(biased, descending) per token; weights (may be NULL)[s, t] =
 * soft_max over the first weights->ne[0] entries of sorted of biased — the nodes ADD, ARGSORT, (top-k view,) GET_ROWS, SOFT_MAX in
 * one launch.  logits / biased f32 [experts <= 1024, tokens], bias f32 [experts] or NULL, sorted i32 [experts, tokens], weights f32
 * [used, tokens].  llama.cpp's node order puts the expert matmuls between the top-k and the GET_ROWS, so the plugin runs the
 * first half here (weights = NULL) and the second half with b200q_op_moe_gate_weights: weights[q, t] = soft_max over q of
 * probs[0, ids[q, t], t] (GET_ROWS of single values + SOFT_MAX); probs f32 [1, experts, tokens], ids i32 [used, tokens]. */
B200Q_API int b200q_op_moe_route(const b200q_tensor * logits, const b200q_tensor * bias, const b200q_tensor * biased, const b200q_tensor * sorted,
                                 const b200q_tensor * weights, void * stream);
B200Q_API int b200q_op_moe_gate_weights(const b200q_tensor * probs, const b200q_tensor * ids, const b200q_tensor * weights, void * stream);

/* two independent SET_ROWS with one destination type in one launch (a layer's K and V cache writes are consecutive nodes) */
B200Q_API int b200q_op_set_rows_pair(const b200q_tensor * src0, const b200q_tensor * idx0, const b200q_tensor * dst0,
                                     const b200q_tensor * src1, const b200q_tensor * idx1, const b200q_tensor * dst1, void * stream);

/* what follows the q / k / v matmuls of a decode step, in one launch: ROPE(q_src) -> q_dst; ROPE(k_src) stored into the rows
 * k_idx of the K cache k_dst (the rotated k has no other reader: llama.cpp emits ROPE, RESHAPE to 2-D, SET_ROWS); v_src stored into
 * the rows v_idx of v_dst.  q / k f32 [head dim, heads, tokens], v f32 [width, tokens], both ROPEs with the same positions,
 * frequency factors and parameters; i64 indices [tokens]; caches [width, rows] of one type (f32 / f16 / bf16).  q_bias / k_bias /
 * v_bias (NULL: none): f32 vectors of a token's width added to the source first — the ADD nodes models with attention biases
 * (gpt-oss, Qwen2) put between the matmul and the ROPE / the store.  Values and rounding are those of the separate operators
 * (binbcast.cu, rope.cu, set-rows.cu). */
B200Q_API int b200q_op_rope_store(const b200q_tensor * q_src, const b200q_tensor * q_dst, const b200q_tensor * k_src, const b200q_tensor * pos,
                                  const b200q_tensor * freq_factors, const b200q_rope_params * rp, const b200q_tensor * k_idx,
                                  const b200q_tensor * k_dst, const b200q_tensor * v_src, const b200q_tensor * v_idx, const b200q_tensor * v_dst,
                                  const b200q_tensor * q_bias, const b200q_tensor * k_bias, const b200q_tensor * v_bias, void * stream);

/* ---- expert-parallel MUL_MAT_ID across the GPUs of one box (NOT in the reference, which asserts that expert tensors are
 *      never split, ggml-cuda.cu:2094).  Experts are sharded by id, activations and routing are replicated: every rank runs
 *      b200q_mul_mat_id on the (token, slot) pairs whose expert it owns and stores the rows it computed into every rank's result
 *      with this call (row r of src -> row row_index[r] of dst here and at dst + peer_delta[q] on every peer: NVLink stores, no
 *      all-to-all); a stream-ordered cross-GPU fence completes the step.  Host logic: b200q/expert_parallel.py. ------------------ */
B200Q_API int b200q_scatter_rows_peers(const float * src, const int32_t * row_index, float * dst, int64_t n_rows, int64_t m,
                                       int64_t dst_row_stride /* floats */, int n_peers, const int64_t * peer_delta /* bytes */, void * stream);

/* ---- experiment (no counterpart in the reference's shipped kernels): one CTA runs tcgen05.mma kind::mxf8f6f4.block_scale
 *      (e2m1 A x e4m3 B, UE8M0 block-32 scales) on a 128 x n x 128 problem laid out by the host exactly as the instruction reads
 *      it (csrc/mxf_probe.cu documents the images); `iters` repeats the four MMAs for an issue-rate measurement (cycles_dev:
 *      device pointer to one int64, or NULL).  tools/mxf_probe.py validates the layouts against a CPU emulation. ------------ */
B200Q_API int b200q_mxf_probe(const void * a_img, const void * b_img, const void * sfa_img, const void * sfb_img, float * d, int n,
                              int iters, long long * cycles_dev, void * stream);

/* ---- host-buffer round trip (what a ggml caller does: tensor_set -> graph_compute -> tensor_get) --- */
/* Weights stay resident on the device (w_dev); activations and results are HOST buffers (pinned or pageable).
 * act_host: n*k floats contiguous, dst_host: n*m floats contiguous.  Synchronous. */
B200Q_API int b200q_mul_mat_host(int type, const void * w_dev, const float * act_host, float * dst_host, int64_t m,
                                 int64_t k, int64_t n, void * act_dev_scratch, void * dst_dev_scratch,
                                 void * workspace, size_t workspace_bytes, void * stream);

#ifdef __cplusplus
}
#endif
#endif /* B200Q_H */
