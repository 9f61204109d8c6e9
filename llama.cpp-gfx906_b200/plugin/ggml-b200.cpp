// ggml-b200.cpp — the ggml backend plugin ("B200" backend): C++ host code behind the reference's own
// plugin API (ggml/src/ggml-backend-impl.h:17-207, GGML_BACKEND_API_VERSION 1) that routes the quantized
// MUL_MAT / MUL_MAT_ID / GET_ROWS hot path to the hand-written sm_100a kernels behind include/b200q.h.
//
// It plays the role ggml-cuda.cu plays in the reference (reg :3751-3790, device :3148-3649, buffer type
// :672-756, buffer :566-664, backend :2525-3077) but is NOT derived from it: nothing besides the hot path and the
// operators that bracket it in a decoder layer (csrc/ops.cu) is claimed (supports_op is exact, so the scheduler
// leaves everything else on ggml-cpu), buffers are plain device allocations with tail slack instead of per-row
// padding, MoE routing never touches the host, row split runs from one process with peer stores instead of gather
// copies, and all CUDA calls go through the single runtime inside libb200q.so.
//
// Loading: dlopen'ed by the reference's registry (GGML_BACKEND_PATH=.../libggml-b200.so or
// ggml_backend_load(), ggml-backend-reg.cpp:232-276,588-592).  The ggml-base symbols used here
// (ggml_backend_buffer_init, ggml_nbytes, ...) are resolved from the host process at load time — the plugin
// is deliberately not linked against any ggml library (and never against anything under oracle/).
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#include "ggml-backend.h"
#include "ggml.h"

#include "../../include/b200q.h"
#include "../../include/ggml-b200.h"

#include <array>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace {

[[noreturn]] void b200_abort(const char * what, int rc, const char * file, int line) {
    fprintf(stderr, "ggml-b200: %s failed (%d): %s  [%s:%d]\n", what, rc, b200q_last_error(), file, line);
    abort();  // same convention as CUDA_CHECK -> ggml_cuda_error -> GGML_ABORT (ggml-cuda.cu:85-96)
}
#define B200_CHECK(expr)                                           \
    do {                                                           \
        int rc_ = (expr);                                          \
        if (rc_ != 0) b200_abort(#expr, rc_, __FILE__, __LINE__);  \
    } while (0)

constexpr size_t B200_TENSOR_ALIGNMENT = 128;
constexpr size_t BUFFER_SLACK     = 256;  // the kernels' 16-byte aligned bulk windows may over-read up to 15 bytes

constexpr int    MAX_DEVICES  = 16;           // length of the tensor_split array llama passes (GGML_CUDA_MAX_DEVICES)
constexpr size_t STAGE_BYTES  = 16u << 20;    // pinned staging buffers of the load path (two per device)
constexpr size_t STAGE_MIN    = 1u << 20;     // uploads below this go straight through cudaMemcpyAsync
constexpr int64_t SPLIT_ROUNDING = 128;       // row-split boundaries are multiples of the tensor-core kernel's row tile
                                              // (the reference rounds to its mmq_y, ggml-cuda.cu:760-771)

struct b200_device_ctx {
    int         device;       // CUDA ordinal
    int         index;        // position in the registry (what tensor_split is indexed by)
    std::string name;         // "B2000", "B2001", ... (what `test-backend-ops -b` matches)
    std::string description;  // "NVIDIA B200"
    void *      xfer_stream = nullptr;  // stream for the synchronous buffer operations
    void *      stage[2]    = { nullptr, nullptr };  // pinned staging buffers (load path)
    void *      stage_ev[2] = { nullptr, nullptr };
    std::mutex  mu;
};

struct b200_buffer_ctx {
    int    device;
    void * base;
};

// what a backend (a stream on the main device) keeps on ANOTHER device for row-split matmuls
struct b200_split_peer {
    void * stream    = nullptr;
    void * done      = nullptr;  // event: this device's slice of dst has been stored into the main device's memory
    void * act       = nullptr;  // local copy of the activations
    size_t act_bytes = 0;
    void * workspace = nullptr;
    size_t ws_bytes  = 0;
};

struct b200_backend_ctx {
    int         device;
    std::string name;
    void *      stream    = nullptr;
    void *      workspace = nullptr;
    size_t      ws_bytes  = 0;
    void *      copy_event = nullptr;            // cpy_tensor_async between two backends
    void *      split_start = nullptr;           // event: the activations of a row-split matmul are ready on this stream
    b200_split_peer peers[MAX_DEVICES];          // indexed by registry device index
    // runs of decode matmuls go out as persistent lists (b200q_list_*), cached by their pointers and shapes: llama.cpp
    // rebuilds its graph for every token but ggml-alloc hands out the same addresses
    std::map<std::vector<int64_t>, b200q_list *> lists;
    std::vector<std::vector<int64_t>>            list_order;  // insertion order, for eviction
    // CUDA-graph replay of a whole ggml graph (decode steps): the launches of a graph whose nodes, shapes, addresses and
    // parameters equal the previous call's are captured on the second sighting and replayed from the third on
    // (ggml-cuda.cu:2876-3100 does the same with cudaGraphExecUpdate; here a changed graph is simply re-captured)
    // (several graphs alternate when the scheduler cuts a model into splits: a small cache keyed by the fingerprint)
    struct graph_entry { void * exec = nullptr; int seen = 0; };
    std::map<std::vector<uint64_t>, graph_entry> graphs;
    std::vector<std::vector<uint64_t>>           graph_order;  // insertion order, for eviction
    bool                                         graphs_off = false;  // a capture broke once: stay eager
    // the activation matrix the last tensor-core MUL_MAT of this graph packed into the workspace (nullptr: none / overwritten)
    const ggml_tensor * packed_act = nullptr;
    // Results of hoisted list members whose own buffer is still in use at the list's position (ggml-alloc reuses buffers: in
    // llama's graphs k's result takes q's buffer once rope has read q): the list writes them into the head of the workspace —
    // decode graphs leave it unused — and a device copy at the node's own position delivers them.  A bump allocator that
    // restarts with every graph, so the same graph gets the same addresses (cached lists and recorded graphs hold them).
    size_t hoist_used = 0;
    bool   hoist_ok   = false;
    struct hoist_copy { const void * src = nullptr; size_t bytes = 0; };
    std::vector<hoist_copy> hoist_copies;  // per node of the graph being dispatched
    // a scratch result whose ONLY reader is a ROPE / SET_ROWS a few nodes on is not copied at all: that operator reads the scratch
    std::vector<std::pair<const void *, const void *>> redirect;  // (the node's own address, where its values are)
};
constexpr size_t HOIST_BYTES = 1u << 20;
constexpr size_t MAX_CACHED_GRAPHS = 256;
constexpr size_t MAX_CACHED_LISTS = 512;

void * device_xfer_stream(b200_device_ctx * d) {
    std::lock_guard<std::mutex> lk(d->mu);
    if (!d->xfer_stream) {
        B200_CHECK(b200q_set_device(d->device));
        B200_CHECK(b200q_stream_create(&d->xfer_stream));
    }
    return d->xfer_stream;
}

// ------------------------------------------------------------------------------------------------ buffer
void buffer_free(ggml_backend_buffer_t buffer) {
    auto * ctx = (b200_buffer_ctx *) buffer->context;
    b200q_set_device(ctx->device);
    b200q_free(ctx->base);
    delete ctx;
}
void * buffer_get_base(ggml_backend_buffer_t buffer) { return ((b200_buffer_ctx *) buffer->context)->base; }

ggml_status buffer_init_tensor(ggml_backend_buffer_t, ggml_tensor *) { return GGML_STATUS_SUCCESS; }

b200_device_ctx * buffer_device(ggml_backend_buffer_t buffer) {
    return (b200_device_ctx *) buffer->buft->device->context;
}

void buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, uint8_t value, size_t offset, size_t size) {
    b200_device_ctx * d = buffer_device(buffer);
    void *            s = device_xfer_stream(d);
    B200_CHECK(b200q_set_device(d->device));
    B200_CHECK(b200q_memset((char *) tensor->data + offset, value, size, s));
    B200_CHECK(b200q_stream_synchronize(s));
}
// host -> device upload of the load path (the reference's set_tensor is one synchronous cudaMemcpyAsync per
// tensor from pageable memory, ggml-cuda.cu:601-607): large pageable sources (the mmap'ed GGUF) are cut into 16 MiB pieces
// that alternate between two pinned staging buffers, so the host-side memcpy of piece i + 1 overlaps the DMA of piece i
void upload(b200_device_ctx * d, void * dst, const void * src, size_t size) {
    void * s = device_xfer_stream(d);
    B200_CHECK(b200q_set_device(d->device));
    static const bool no_stage = getenv("GGML_B200_NO_STAGING") != nullptr;
    if (size < STAGE_MIN || no_stage) {
        B200_CHECK(b200q_memcpy(dst, src, size, B200Q_H2D, s));
        B200_CHECK(b200q_stream_synchronize(s));
        return;
    }
    std::lock_guard<std::mutex> lk(d->mu);
    for (int i = 0; i < 2; ++i) {
        if (!d->stage[i]) {
            if (b200q_host_alloc(&d->stage[i], STAGE_BYTES) != 0 || b200q_event_create(&d->stage_ev[i]) != 0) {
                d->stage[i] = nullptr;  // no pinned memory: plain copy
                B200_CHECK(b200q_memcpy(dst, src, size, B200Q_H2D, s));
                B200_CHECK(b200q_stream_synchronize(s));
                return;
            }
        }
    }
    size_t off = 0;
    for (int i = 0; off < size; i ^= 1) {
        const size_t n = size - off < STAGE_BYTES ? size - off : STAGE_BYTES;
        B200_CHECK(b200q_event_synchronize(d->stage_ev[i]));  // the DMA that last read this staging buffer is done
        memcpy(d->stage[i], (const char *) src + off, n);
        B200_CHECK(b200q_memcpy((char *) dst + off, d->stage[i], n, B200Q_H2D, s));
        B200_CHECK(b200q_event_record(d->stage_ev[i], s));
        off += n;
    }
    B200_CHECK(b200q_stream_synchronize(s));
}

void buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    upload(buffer_device(buffer), (char *) tensor->data + offset, data, size);
}
void buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    b200_device_ctx * d = buffer_device(buffer);
    void *            s = device_xfer_stream(d);
    B200_CHECK(b200q_set_device(d->device));
    B200_CHECK(b200q_memcpy(data, (const char *) tensor->data + offset, size, B200Q_D2H, s));
    B200_CHECK(b200q_stream_synchronize(s));
}
bool buffer_is_b200(ggml_backend_buffer_t buffer);

bool buffer_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    if (!src->buffer || !buffer_is_b200(src->buffer)) return false;
    auto * sctx = (b200_buffer_ctx *) src->buffer->context;
    auto * dctx = (b200_buffer_ctx *) buffer->context;
    if (sctx->device != dctx->device) return false;  // cross-device copies go through the host path of ggml-backend
    b200_device_ctx * d = buffer_device(buffer);
    void *            s = device_xfer_stream(d);
    B200_CHECK(b200q_set_device(d->device));
    B200_CHECK(b200q_memcpy(dst->data, src->data, ggml_nbytes(src), B200Q_D2D, s));
    B200_CHECK(b200q_stream_synchronize(s));
    return true;
}
void buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    b200_device_ctx * d = buffer_device(buffer);
    void *            s = device_xfer_stream(d);
    B200_CHECK(b200q_set_device(d->device));
    B200_CHECK(b200q_memset(((b200_buffer_ctx *) buffer->context)->base, value, buffer->size, s));
    B200_CHECK(b200q_stream_synchronize(s));
}

const ggml_backend_buffer_i buffer_iface = {
    /* .free_buffer   = */ buffer_free,
    /* .get_base      = */ buffer_get_base,
    /* .init_tensor   = */ buffer_init_tensor,
    /* .memset_tensor = */ buffer_memset_tensor,
    /* .set_tensor    = */ buffer_set_tensor,
    /* .get_tensor    = */ buffer_get_tensor,
    /* .cpy_tensor    = */ buffer_cpy_tensor,
    /* .clear         = */ buffer_clear,
    /* .reset         = */ nullptr,
};
bool buffer_is_b200(ggml_backend_buffer_t buffer) { return buffer->iface.free_buffer == buffer_free; }

// ------------------------------------------------------------------------------------------- buffer type
const char * buft_get_name(ggml_backend_buffer_type_t buft) {
    return ((b200_device_ctx *) buft->device->context)->name.c_str();
}
ggml_backend_buffer_t buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    auto * d = (b200_device_ctx *) buft->device->context;
    if (b200q_set_device(d->device) != 0) return nullptr;
    void * base = nullptr;
    if (b200q_malloc(&base, size + BUFFER_SLACK) != 0) {
        fprintf(stderr, "ggml-b200: allocating %.2f MiB on device %d failed: %s\n", size / 1048576.0, d->device,
                b200q_last_error());
        return nullptr;  // OOM -> NULL, as the reference does (ggml-cuda.cu:689-694)
    }
    return ggml_backend_buffer_init(buft, buffer_iface, new b200_buffer_ctx{ d->device, base }, size);
}
size_t buft_get_alignment(ggml_backend_buffer_type_t) { return B200_TENSOR_ALIGNMENT; }
size_t buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * tensor) {
    // quantized rows are read through 16-byte aligned windows: keep 16 bytes of slack behind every such tensor so a
    // window never touches another tensor's bytes that a concurrent copy might be writing
    return ggml_nbytes(tensor) + (ggml_is_quantized(tensor->type) ? 16 : 0);
}
bool buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i buft_iface = {
    /* .get_name       = */ buft_get_name,
    /* .alloc_buffer   = */ buft_alloc_buffer,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ buft_get_alloc_size,
    /* .is_host        = */ buft_is_host,
};

// ------------------------------------------------------------------------------- registry (forward decls)
struct b200_device;
int               registry_device_count();
b200_device_ctx * registry_device_ctx(int index);
ggml_backend_dev_t registry_device(int index);

// ------------------------------------------------------------------------------------- split buffer type
// Row split behind the plugin (the reference's `-sm row`: ggml_backend_cuda_split_buffer_type, ggml-cuda.cu:794-1089,
// looked up by llama through get_proc_address, src/llama-model.cpp:366-393).  A weight matrix allocated in this buffer
// type is cut by rows over the devices — device i owns rows [round128(m * split[i]), round128(m * split[i + 1])), split
// = cumulative normalised tensor_split, the last device takes the remainder (get_row_split, :773-786) — and every slice
// is uploaded to its own device by set_tensor.  MUL_MAT on such a weight runs from ONE process (compute_mul_mat_split).
struct b200_split_buft_ctx {
    int                            main_device;  // registry index
    std::array<float, MAX_DEVICES> split;        // cumulative fractions
    std::string                    name;
};
struct b200_split_tensor {  // tensor->extra
    void *  data[MAX_DEVICES]     = {};
    int64_t row_low[MAX_DEVICES]  = {};
    int64_t row_high[MAX_DEVICES] = {};
};
struct b200_split_buffer_ctx {
    std::vector<b200_split_tensor *> tensors;
};

void split_rows(const b200_split_buft_ctx * bc, int64_t nrows, int id, int64_t * lo, int64_t * hi) {
    const int n = registry_device_count();
    *lo = id == 0 ? 0 : (int64_t) (nrows * bc->split[id]);
    *lo -= *lo % SPLIT_ROUNDING;
    if (id == n - 1) {
        *hi = nrows;
    } else {
        *hi = (int64_t) (nrows * bc->split[id + 1]);
        *hi -= *hi % SPLIT_ROUNDING;
    }
    if (id >= n) *lo = *hi = nrows;
}

const char * split_buft_get_name(ggml_backend_buffer_type_t buft) { return ((b200_split_buft_ctx *) buft->context)->name.c_str(); }
bool buft_is_split(ggml_backend_buffer_type_t buft) { return buft->iface.get_name == split_buft_get_name; }

void split_buffer_free(ggml_backend_buffer_t buffer) {
    auto * ctx = (b200_split_buffer_ctx *) buffer->context;
    for (b200_split_tensor * t : ctx->tensors) {
        for (int id = 0; id < registry_device_count(); ++id) {
            if (t->data[id]) {
                b200q_set_device(registry_device_ctx(id)->device);
                b200q_free(t->data[id]);
            }
        }
        delete t;
    }
    delete ctx;
}
void * split_buffer_get_base(ggml_backend_buffer_t) { return (void *) 0x1000; }  // never dereferenced (as in the reference, :857-862)

ggml_status split_buffer_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * tensor) {
    GGML_ASSERT(tensor->view_src == nullptr);  // views of split tensors are not supported
    GGML_ASSERT(ggml_is_contiguous(tensor) && "split buffers only supported for contiguous tensors");
    auto * ctx = (b200_split_buffer_ctx *) buffer->context;
    auto * bc  = (b200_split_buft_ctx *) buffer->buft->context;
    auto * ext = new b200_split_tensor;
    ctx->tensors.push_back(ext);
    const size_t row_bytes = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int id = 0; id < registry_device_count(); ++id) {
        split_rows(bc, ggml_nrows(tensor), id, &ext->row_low[id], &ext->row_high[id]);
        const int64_t nr = ext->row_high[id] - ext->row_low[id];
        if (nr <= 0) continue;
        B200_CHECK(b200q_set_device(registry_device_ctx(id)->device));
        if (b200q_malloc(&ext->data[id], (size_t) nr * row_bytes + BUFFER_SLACK) != 0) {
            fprintf(stderr, "ggml-b200: split slice of %s on device %d: %s\n", tensor->name, id, b200q_last_error());
            return GGML_STATUS_ALLOC_FAILED;
        }
    }
    tensor->extra = ext;
    return GGML_STATUS_SUCCESS;
}
void upload(b200_device_ctx * d, void * dst, const void * src, size_t size);
void split_buffer_set_tensor(ggml_backend_buffer_t, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor));  // split tensors are set in one piece (:883-887)
    auto *       ext       = (b200_split_tensor *) tensor->extra;
    const size_t row_bytes = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int id = 0; id < registry_device_count(); ++id) {
        if (!ext->data[id]) continue;
        upload(registry_device_ctx(id), ext->data[id], (const char *) data + (size_t) ext->row_low[id] * row_bytes,
               (size_t) (ext->row_high[id] - ext->row_low[id]) * row_bytes);
    }
}
void split_buffer_get_tensor(ggml_backend_buffer_t, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor));
    auto *       ext       = (b200_split_tensor *) tensor->extra;
    const size_t row_bytes = ggml_row_size(tensor->type, tensor->ne[0]);
    for (int id = 0; id < registry_device_count(); ++id) {
        if (!ext->data[id]) continue;
        b200_device_ctx * d = registry_device_ctx(id);
        void *            s = device_xfer_stream(d);
        B200_CHECK(b200q_set_device(d->device));
        B200_CHECK(b200q_memcpy((char *) data + (size_t) ext->row_low[id] * row_bytes, ext->data[id],
                                (size_t) (ext->row_high[id] - ext->row_low[id]) * row_bytes, B200Q_D2H, s));
        B200_CHECK(b200q_stream_synchronize(s));
    }
}
void split_buffer_clear(ggml_backend_buffer_t, uint8_t) {}

const ggml_backend_buffer_i split_buffer_iface = {
    /* .free_buffer   = */ split_buffer_free,
    /* .get_base      = */ split_buffer_get_base,
    /* .init_tensor   = */ split_buffer_init_tensor,
    /* .memset_tensor = */ nullptr,
    /* .set_tensor    = */ split_buffer_set_tensor,
    /* .get_tensor    = */ split_buffer_get_tensor,
    /* .cpy_tensor    = */ nullptr,
    /* .clear         = */ split_buffer_clear,
    /* .reset         = */ nullptr,
};

ggml_backend_buffer_t split_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    // the slices are allocated per tensor in init_tensor (their sizes depend on the rounding); `size` is the cumulative
    // bound ggml-alloc enforces, as returned by get_alloc_size
    return ggml_backend_buffer_init(buft, split_buffer_iface, new b200_split_buffer_ctx, size);
}
size_t split_buft_get_alignment(ggml_backend_buffer_type_t) { return B200_TENSOR_ALIGNMENT; }
size_t split_buft_get_alloc_size(ggml_backend_buffer_type_t buft, const ggml_tensor * tensor) {
    auto * bc = (b200_split_buft_ctx *) buft->context;
    GGML_ASSERT(ggml_is_contiguous(tensor) && "split buffers only supported for contiguous tensors");
    size_t total = 0;
    for (int id = 0; id < registry_device_count(); ++id) {
        int64_t lo, hi;
        split_rows(bc, ggml_nrows(tensor), id, &lo, &hi);
        if (hi > lo) total += (size_t) (hi - lo) * ggml_row_size(tensor->type, tensor->ne[0]) + BUFFER_SLACK;
    }
    return total;
}
bool split_buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i split_buft_iface = {
    /* .get_name       = */ split_buft_get_name,
    /* .alloc_buffer   = */ split_buft_alloc_buffer,
    /* .get_alignment  = */ split_buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ split_buft_get_alloc_size,
    /* .is_host        = */ split_buft_is_host,
};

// ggml_backend_split_buffer_type_t (ggml-backend.h:185-186): tensor_split has MAX_DEVICES entries (proportions per
// device; all zero or NULL = equal shares)
ggml_backend_buffer_type_t split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mu;
    static std::map<std::pair<int, std::array<float, MAX_DEVICES>>, ggml_backend_buffer_type> bufts;
    std::lock_guard<std::mutex> lk(mu);
    const int n = registry_device_count();
    if (main_device < 0 || main_device >= n) return nullptr;
    std::array<float, MAX_DEVICES> cum = {};
    float sum = 0.0f;
    for (int i = 0; i < n && tensor_split; ++i) sum += tensor_split[i];
    float run = 0.0f;
    for (int i = 0; i < n; ++i) {
        cum[i] = sum > 0.0f ? run / sum : (float) i / n;
        if (sum > 0.0f) run += tensor_split[i];
    }
    for (int i = n; i < MAX_DEVICES; ++i) cum[i] = 1.0f;
    auto key = std::make_pair(main_device, cum);
    auto it  = bufts.find(key);
    if (it != bufts.end()) return &it->second;
    auto * ctx = new b200_split_buft_ctx{ main_device, cum, "B200" + std::to_string(main_device) + "_Split" };
    ggml_backend_buffer_type buft{ split_buft_iface, registry_device(main_device), ctx };
    return &bufts.emplace(key, buft).first->second;
}

// -------------------------------------------------------------------------------------- host buffer type
// pinned host memory for the tensors llama keeps on the CPU side of a partially offloaded model and for its input
// staging (ggml_backend_cuda_host_buffer_type, ggml-cuda.cu:1093-1160): the buffer behaves like a CPU buffer, only the
// allocation differs, so asynchronous copies from it really are asynchronous
const char * host_buft_get_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
void host_buffer_free(ggml_backend_buffer_t buffer) { b200q_host_free(buffer->context); }
ggml_backend_buffer_t host_buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    void * ptr = nullptr;
    if (getenv("GGML_B200_NO_PINNED") || b200q_host_alloc(&ptr, size) != 0 || !ptr) {
        return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);  // pageable host memory, as the reference
    }
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(ptr, size);
    buffer->buft              = buft;
    buffer->iface.free_buffer = host_buffer_free;
    return buffer;
}
size_t host_buft_get_alignment(ggml_backend_buffer_type_t) { return ggml_backend_buft_get_alignment(ggml_backend_cpu_buffer_type()); }
bool   host_buft_is_host(ggml_backend_buffer_type_t) { return true; }
const ggml_backend_buffer_type_i host_buft_iface = {
    /* .get_name       = */ host_buft_get_name,
    /* .alloc_buffer   = */ host_buft_alloc_buffer,
    /* .get_alignment  = */ host_buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ nullptr,  // ggml_nbytes
    /* .is_host        = */ host_buft_is_host,
};
ggml_backend_buffer_type_t host_buffer_type() {
    static ggml_backend_buffer_type buft{ host_buft_iface, registry_device(0), nullptr };
    return &buft;
}

// ------------------------------------------------------------------------------------------------ ops
bool type_ok(ggml_type t) { return b200q_type_supported((int) t) != 0; }
// unquantized weights go through the tensor-core kernel in 64-element chunks
bool dense_weight(ggml_type t) { return t == GGML_TYPE_F16 || t == GGML_TYPE_BF16 || t == GGML_TYPE_F32; }
bool weight_ok(const ggml_tensor * a) { return (type_ok(a->type) || a->type == GGML_TYPE_F32) && (!dense_weight(a->type) || a->ne[0] % 64 == 0); }
// MUL_MAT takes unquantized weights with any k (dense_mm.cu: the attention products), MUL_MAT_ID and split weights do not
bool mul_mat_weight_ok(const ggml_tensor * a) { return dense_weight(a->type) || weight_ok(a); }

bool rows_dense(const ggml_tensor * t) { return t->nb[0] == ggml_type_size(t->type); }

// exactly what the kernels take; anything else is declined so the scheduler keeps it on another backend
// (north_star: no CPU fallback for supported ops; cf. the reference's rules at ggml-cuda.cu:3265-3348)
bool tensor_is_split(const ggml_tensor * t) { return t->buffer && buft_is_split(t->buffer->buft); }

bool supports_mul_mat(const b200_device_ctx * dev, const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * b = op->src[1];
    // f16 activations are converted to f32 first (the reference converts src1 as well: ggml-cuda.cu:1237-1262, 1489-1505)
    if (!mul_mat_weight_ok(a) || (b->type != GGML_TYPE_F32 && b->type != GGML_TYPE_F16) || op->type != GGML_TYPE_F32) return false;
    if (tensor_is_split(a)) {
        // row-split weights: plain 2-D matrices only, and the device that runs the op must own rows (ggml-cuda.cu:3270-3283)
        if (!weight_ok(a) || b->type != GGML_TYPE_F32) return false;
        if (a->ne[2] > 1 || a->ne[3] > 1 || !ggml_is_contiguous(b) || !ggml_is_contiguous(op)) return false;
        int64_t lo, hi;
        split_rows((const b200_split_buft_ctx *) a->buffer->buft->context, ggml_nrows(a), dev->index, &lo, &hi);
        if (lo == hi) return false;
    }
    if (!rows_dense(a) || !rows_dense(b) || !rows_dense(op)) return false;  // dims 1..3 may be permuted / broadcast
    if (a->ne[0] != b->ne[0] || a->ne[0] % ggml_blck_size(a->type)) return false;
    if (b->ne[2] % a->ne[2] || b->ne[3] % a->ne[3]) return false;
    if (a->ne[0] > b200q_max_k((int) a->type)) return false;  // the kernels' shared-memory plans (host-side planner probe)
    return true;
}
bool supports_mul_mat_id(const ggml_tensor * op) {
    const ggml_tensor * as  = op->src[0];
    const ggml_tensor * b   = op->src[1];
    const ggml_tensor * ids = op->src[2];
    if (tensor_is_split(as)) return false;  // experts are not row-split (the reference asserts the same, ggml-cuda.cu:2094)
    if (!weight_ok(as) || as->ne[2] > 1024 || b->type != GGML_TYPE_F32 || ids->type != GGML_TYPE_I32 || op->type != GGML_TYPE_F32) return false;
    if (!rows_dense(as) || !rows_dense(b) || !rows_dense(op) || ids->nb[0] != sizeof(int32_t)) return false;
    if (as->ne[3] != 1 || b->ne[3] != 1 || ids->ne[2] != 1 || ids->ne[3] != 1) return false;
    if (as->ne[0] % ggml_blck_size(as->type) || as->ne[0] > b200q_max_k((int) as->type)) return false;
    if (b->ne[1] != 1 && b->ne[1] != ids->ne[0]) return false;
    return true;
}
bool supports_get_rows(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * r = op->src[1];
    if (!(type_ok(a->type) || a->type == GGML_TYPE_F32) || r->type != GGML_TYPE_I32 || op->type != GGML_TYPE_F32) return false;
    if (!rows_dense(a) || !rows_dense(op) || r->ne[3] != 1) return false;
    return true;
}

// ---- the operators around the matmuls (csrc/ops.cu): what the kernels take, stated once for supports_op and compute ------
b200q_tensor bt(const ggml_tensor * t) {
    b200q_tensor r{};
    r.data = t->data;
    r.type = (int) t->type;  // F32 0, F16 1, I32 26, I64 27, BF16 30: ggml's numbers are the C-ABI's
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; }
    return r;
}
bool is_f32(const ggml_tensor * t) { return t && t->type == GGML_TYPE_F32; }
bool float_type(ggml_type t) { return t == GGML_TYPE_F32 || t == GGML_TYPE_F16 || t == GGML_TYPE_BF16; }
int unary_code(const ggml_tensor * op) {
    switch (ggml_get_unary_op(op)) {
        case GGML_UNARY_OP_SILU:       return B200Q_UNARY_SILU;
        case GGML_UNARY_OP_GELU:       return B200Q_UNARY_GELU;
        case GGML_UNARY_OP_GELU_QUICK: return B200Q_UNARY_GELU_QUICK;
        case GGML_UNARY_OP_GELU_ERF:   return B200Q_UNARY_GELU_ERF;
        case GGML_UNARY_OP_RELU:       return B200Q_UNARY_RELU;
        case GGML_UNARY_OP_SIGMOID:    return B200Q_UNARY_SIGMOID;
        case GGML_UNARY_OP_TANH:       return B200Q_UNARY_TANH;
        case GGML_UNARY_OP_NEG:        return B200Q_UNARY_NEG;
        case GGML_UNARY_OP_ABS:        return B200Q_UNARY_ABS;
        case GGML_UNARY_OP_EXP:        return B200Q_UNARY_EXP;
        default:                       return -1;
    }
}
int glu_code(const ggml_tensor * op) {
    switch (ggml_get_glu_op(op)) {
        case GGML_GLU_OP_REGLU:       return B200Q_GLU_REGLU;
        case GGML_GLU_OP_GEGLU:       return B200Q_GLU_GEGLU;
        case GGML_GLU_OP_SWIGLU:      return B200Q_GLU_SWIGLU;
        case GGML_GLU_OP_GEGLU_ERF:   return B200Q_GLU_GEGLU_ERF;
        case GGML_GLU_OP_GEGLU_QUICK: return B200Q_GLU_GEGLU_QUICK;
        case GGML_GLU_OP_SWIGLU_OAI:  return B200Q_GLU_SWIGLU_OAI;
        default:                      return -1;
    }
}
int bin_code(ggml_op op) {
    return op == GGML_OP_ADD ? B200Q_BIN_ADD : op == GGML_OP_SUB ? B200Q_BIN_SUB : op == GGML_OP_MUL ? B200Q_BIN_MUL : B200Q_BIN_DIV;
}
bool supports_elementwise(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * b = op->src[1];
    switch (op->op) {
        case GGML_OP_RMS_NORM:
            return is_f32(a) && is_f32(op) && a->nb[0] == 4 && op->nb[0] == 4;
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV:
            return is_f32(a) && is_f32(b) && is_f32(op) && ggml_are_same_shape(a, op) && ggml_can_repeat(b, a);
        case GGML_OP_SCALE:
            return is_f32(a) && is_f32(op) && ggml_is_contiguous(a) && ggml_is_contiguous(op);
        case GGML_OP_UNARY:
            return is_f32(a) && is_f32(op) && ggml_is_contiguous(a) && ggml_is_contiguous(op) && unary_code(op) >= 0;
        case GGML_OP_GLU:
            return is_f32(a) && is_f32(op) && ggml_is_contiguous_1(a) && ggml_is_contiguous_1(op) && glu_code(op) >= 0 &&
                   (!b || (is_f32(b) && ggml_is_contiguous_1(b) && ggml_are_same_shape(a, b)));
        case GGML_OP_SOFT_MAX: {
            const ggml_tensor * sinks = op->src[2];
            if (!is_f32(a) || !is_f32(op) || a->nb[0] != 4 || !ggml_is_contiguous(op)) return false;
            if (b && ((b->type != GGML_TYPE_F32 && b->type != GGML_TYPE_F16) || b->nb[0] != ggml_type_size(b->type) || b->ne[0] != a->ne[0] ||
                      b->ne[1] < a->ne[1])) return false;
            if (sinks && (!is_f32(sinks) || sinks->ne[0] != a->ne[2] || !ggml_is_contiguous(sinks))) return false;
            return true;
        }
        case GGML_OP_ROPE: {
            const int mode = ((const int32_t *) op->op_params)[2];
            if (mode != 0 && mode != GGML_ROPE_TYPE_NEOX) return false;  // multimodal / vision rotary embeddings stay on the CPU backend
            if ((a->type != GGML_TYPE_F32 && a->type != GGML_TYPE_F16) || op->type != a->type) return false;
            if (a->nb[0] != ggml_type_size(a->type) || op->nb[0] != ggml_type_size(op->type) || a->ne[0] % 2) return false;
            if (!b || b->type != GGML_TYPE_I32 || b->nb[0] != 4 || b->ne[0] < a->ne[2]) return false;
            if (op->src[2] && (!is_f32(op->src[2]) || op->src[2]->nb[0] != 4)) return false;
            return true;
        }
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP:
            return float_type(a->type) && float_type(op->type) && ggml_nelements(a) == ggml_nelements(op);
        case GGML_OP_ARGSORT:  // router top-k (ggml_top_k = a view of the descending argsort); argsort.cu takes the same row cap
            return is_f32(a) && op->type == GGML_TYPE_I32 && a->nb[0] == 4 && op->nb[0] == 4 && a->ne[0] <= 1024;
        case GGML_OP_ADD_ID: {  // a [n, used, tokens] + b [n, experts] picked by ids [used, tokens] (add-id.cu; ggml-cuda.cu:3455)
            const ggml_tensor * ids = op->src[2];
            return is_f32(a) && is_f32(b) && ids && ids->type == GGML_TYPE_I32 && is_f32(op) && a->nb[0] == 4 && b->nb[0] == 4 &&
                   op->nb[0] == 4 && a->ne[3] == 1 && b->ne[0] == a->ne[0] && ids->ne[0] == a->ne[1] && ids->ne[1] == a->ne[2];
        }
        case GGML_OP_SET_ROWS:
            return is_f32(a) && b && b->type == GGML_TYPE_I64 && float_type(op->type) && a->ne[0] == op->ne[0] && a->ne[2] == op->ne[2] &&
                   a->ne[3] == op->ne[3] && b->ne[0] == a->ne[1] && a->ne[2] % b->ne[1] == 0 && a->ne[3] % b->ne[2] == 0;
        default:
            return false;
    }
}
// bt() of an operator's source, following a pending scratch redirect (see b200_backend_ctx::redirect)
b200q_tensor bt_src(b200_backend_ctx * ctx, const ggml_tensor * t) {
    b200q_tensor r = bt(t);
    for (size_t q = ctx->redirect.size(); q-- > 0;) {  // newest first; an entry serves its one reader and goes
        if (ctx->redirect[q].first == r.data) {
            r.data = (void *) ctx->redirect[q].second;
            ctx->redirect.erase(ctx->redirect.begin() + (long) q);
            break;
        }
    }
    return r;
}
b200q_rope_params rope_params_of(const ggml_tensor * op) {
    const int32_t *   ip = (const int32_t *) op->op_params;
    b200q_rope_params rp{};
    rp.n_dims = ip[1]; rp.mode = ip[2];
    const int n_ctx_orig = ip[4];
    float beta_fast, beta_slow;
    memcpy(&rp.freq_base, ip + 5, 4); memcpy(&rp.freq_scale, ip + 6, 4); memcpy(&rp.ext_factor, ip + 7, 4);
    memcpy(&rp.attn_factor, ip + 8, 4); memcpy(&beta_fast, ip + 9, 4); memcpy(&beta_slow, ip + 10, 4);
    ggml_rope_yarn_corr_dims(rp.n_dims, n_ctx_orig, rp.freq_base, beta_fast, beta_slow, rp.corr_dims);
    return rp;
}
int compute_elementwise(b200_backend_ctx * ctx, const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * b = op->src[1];
    const b200q_tensor  ta = (op->op == GGML_OP_ROPE || op->op == GGML_OP_SET_ROWS || op->op == GGML_OP_ADD) ? bt_src(ctx, a) : bt(a), td = bt(op);
    b200q_tensor        tb{}, tc{};
    if (b) tb = bt(b);
    if (op->src[2]) tc = bt(op->src[2]);
    const float *   fp = (const float *) op->op_params;
    const int32_t * ip = (const int32_t *) op->op_params;
    switch (op->op) {
        case GGML_OP_RMS_NORM: return b200q_op_rms_norm(&ta, &td, fp[0], ctx->stream);
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV:
            return b200q_op_bin(bin_code(op->op), &ta, &tb, &td, ctx->stream);
        case GGML_OP_SCALE: return b200q_op_scale(&ta, &td, fp[0], fp[1], ctx->stream);
        case GGML_OP_UNARY: return b200q_op_unary(unary_code(op), &ta, &td, ctx->stream);
        case GGML_OP_GLU:   return b200q_op_glu(glu_code(op), &ta, b ? &tb : nullptr, &td, ip[1], fp[2], fp[3], ctx->stream);
        case GGML_OP_SOFT_MAX:
            return b200q_op_soft_max(&ta, b ? &tb : nullptr, op->src[2] ? &tc : nullptr, &td, fp[0], fp[1], ctx->stream);
        case GGML_OP_ROPE: {
            const b200q_rope_params rp = rope_params_of(op);
            return b200q_op_rope(&ta, &tb, op->src[2] ? &tc : nullptr, &td, &rp, ctx->stream);
        }
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP: return b200q_op_cpy(&ta, &td, ctx->stream);
        case GGML_OP_SET_ROWS: return b200q_op_set_rows(&ta, &tb, &td, ctx->stream);
        case GGML_OP_ADD_ID:   return b200q_op_add_id(&ta, &tb, &tc, &td, ctx->stream);
        case GGML_OP_ARGSORT:  return b200q_op_argsort(&ta, &td, ip[0] == GGML_SORT_ORDER_DESC, ctx->stream);
        default: return B200Q_ENOTSUP;
    }
}

bool device_supports_op(ggml_backend_dev_t dev, const ggml_tensor * op) {
    // split buffers hold row slices on several devices and have no address: only MUL_MAT can read them (ggml-cuda.cu:3265-3283).
    // llama.cpp asks this for every weight (a norm weight under MUL, an embedding under GET_ROWS) to choose its buffer type.
    if (op->op != GGML_OP_MUL_MAT) {
        for (int i = 0; i < GGML_MAX_SRC; ++i) {
            if (op->src[i] && tensor_is_split(op->src[i])) return false;
        }
    }
    switch (op->op) {
        case GGML_OP_NONE:
        case GGML_OP_VIEW:
        case GGML_OP_RESHAPE:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;  // no-ops; the harness probes every tensor of the graph (test-backend-ops.cpp:1108-1115)
        case GGML_OP_MUL_MAT:    return supports_mul_mat((const b200_device_ctx *) dev->context, op);
        case GGML_OP_MUL_MAT_ID: return supports_mul_mat_id(op);
        case GGML_OP_GET_ROWS:   return supports_get_rows(op);
        default: return supports_elementwise(op);
    }
}

// recorded graphs hold the workspace and the persistent lists by address: when either goes away, so do the graphs
void drop_graphs(b200_backend_ctx * ctx) {
    bool any = false;
    for (auto & kv : ctx->graphs) any = any || kv.second.exec;
    if (any) B200_CHECK(b200q_stream_synchronize(ctx->stream));
    // (fingerprints that have only been sighted keep their count: nothing was recorded for them yet)
    std::vector<std::vector<uint64_t>> keep;
    for (auto & key : ctx->graph_order) {
        auto it = ctx->graphs.find(key);
        if (it == ctx->graphs.end()) continue;
        if (it->second.exec) {
            b200q_graph_destroy(it->second.exec);
            ctx->graphs.erase(it);
        } else {
            keep.push_back(key);
        }
    }
    ctx->graph_order.swap(keep);
}

// The first HOIST_BYTES of the workspace are the scratch of hoisted matmul results (compute_list_run) for the whole graph; what
// operators ask for (packed activations, MoE tables, f32 copies of f16 activations) starts behind it: ws_avail() bytes.
void * backend_workspace(b200_backend_ctx * ctx, size_t need) {
    need += HOIST_BYTES;
    if (need > ctx->ws_bytes) {
        drop_graphs(ctx);
        if (ctx->workspace) {
            B200_CHECK(b200q_stream_synchronize(ctx->stream));
            B200_CHECK(b200q_free(ctx->workspace));
        }
        const size_t bytes = need + need / 4;
        B200_CHECK(b200q_malloc(&ctx->workspace, bytes));
        ctx->ws_bytes = bytes;
    }
    return (char *) ctx->workspace + HOIST_BYTES;
}
size_t ws_avail(const b200_backend_ctx * ctx) { return ctx->ws_bytes > HOIST_BYTES ? ctx->ws_bytes - HOIST_BYTES : 0; }

uint32_t weight_flags(const ggml_tensor * w) {
    const ggml_tensor * root = w->view_src ? w->view_src : w;
    const bool immutable = root->buffer && ggml_backend_buffer_get_usage(root->buffer) == GGML_BACKEND_BUFFER_USAGE_WEIGHTS;
    return immutable ? B200Q_FLAG_WEIGHTS_IMMUTABLE : 0;
}

void * grow(void ** buf, size_t * have, size_t need, void * stream) {
    if (need > *have) {
        if (*buf) {
            B200_CHECK(b200q_stream_synchronize(stream));
            B200_CHECK(b200q_free(*buf));
        }
        const size_t bytes = need + need / 4;
        B200_CHECK(b200q_malloc(buf, bytes));
        *have = bytes;
    }
    return *buf;
}

// MUL_MAT on a row-split weight, from one process (replaces ggml_cuda_op_mul_mat, ggml-cuda.cu:1436-1754, which copies
// src1 to every device, runs the slices and gathers dst through cudaMemcpy2DAsync on the main device).  Here every other
// device gets the activations by ONE peer copy on its own stream, runs the ordinary kernel on its rows, and the kernel's
// epilogue stores the result rows straight into the main device's dst over NVLink (dst + row_low; the pitch is that of the
// whole tensor) — no gather copies, no collective; an event per device joins its stream back into the backend's stream.
int compute_mul_mat_split(b200_backend_ctx * ctx, const ggml_tensor * op) {
    const ggml_tensor * a   = op->src[0];
    const ggml_tensor * b   = op->src[1];
    auto *              ext = (b200_split_tensor *) a->extra;
    const int           n_dev = registry_device_count();
    const size_t        act_bytes = ggml_nbytes(b);
    int main_index = -1;
    for (int id = 0; id < n_dev; ++id) if (registry_device_ctx(id)->device == ctx->device) main_index = id;
    if (main_index < 0 || !ext) return B200Q_EINVAL;
    b200q_mul_mat_args q{};
    q.type = (int) a->type;
    q.k = a->ne[0]; q.n = b->ne[1];
    q.ne02 = q.ne03 = q.ne12 = q.ne13 = 1;
    q.nb01 = ggml_row_size(a->type, a->ne[0]);
    q.nb11 = b->nb[1];
    q.nb1  = op->nb[1];
    q.flags = weight_flags(a);
    bool forked = false;
    for (int id = 0; id < n_dev; ++id) {
        const int64_t lo = ext->row_low[id], nr = ext->row_high[id] - lo;
        if (nr <= 0 || id == main_index) continue;
        b200_device_ctx * d = registry_device_ctx(id);
        b200_split_peer & pe = ctx->peers[id];
        if (!forked) {
            if (!ctx->split_start) B200_CHECK(b200q_event_create(&ctx->split_start));
            B200_CHECK(b200q_event_record(ctx->split_start, ctx->stream));
            forked = true;
        }
        B200_CHECK(b200q_set_device(d->device));
        if (!pe.stream) {
            if (int rc = b200q_enable_peer_access(d->device, ctx->device)) return rc;  // its kernels store into the main device
            B200_CHECK(b200q_stream_create(&pe.stream));
            B200_CHECK(b200q_event_create(&pe.done));
        }
        B200_CHECK(b200q_stream_wait_event(pe.stream, ctx->split_start));
        void * act = grow(&pe.act, &pe.act_bytes, act_bytes, pe.stream);
        B200_CHECK(b200q_memcpy_peer(act, d->device, b->data, ctx->device, act_bytes, pe.stream));
        q.m = nr; q.nb02 = q.nb03 = q.nb01 * nr; q.nb12 = q.nb13 = q.nb11 * q.n; q.nb2 = q.nb3 = q.nb1 * q.n;
        const size_t need = b200q_mul_mat_workspace(&q);
        void *       ws   = need ? grow(&pe.workspace, &pe.ws_bytes, need, pe.stream) : nullptr;
        int rc = b200q_mul_mat(&q, ext->data[id], (const float *) act, (float *) op->data + lo, ws, pe.ws_bytes, pe.stream);
        if (rc == 0) rc = b200q_event_record(pe.done, pe.stream);
        B200_CHECK(b200q_set_device(ctx->device));
        if (rc) return rc;
    }
    {   // this device's own rows, on the backend's stream
        const int64_t lo = ext->row_low[main_index], nr = ext->row_high[main_index] - lo;
        q.m = nr; q.nb02 = q.nb03 = q.nb01 * nr; q.nb12 = q.nb13 = q.nb11 * q.n; q.nb2 = q.nb3 = q.nb1 * q.n;
        const size_t need = b200q_mul_mat_workspace(&q);
        void *       ws   = need ? backend_workspace(ctx, need) : nullptr;
        if (int rc = b200q_mul_mat(&q, ext->data[main_index], (const float *) b->data, (float *) op->data + lo, ws, ws_avail(ctx), ctx->stream)) return rc;
    }
    for (int id = 0; id < n_dev; ++id) {
        if (id != main_index && ext->row_high[id] > ext->row_low[id]) B200_CHECK(b200q_stream_wait_event(ctx->stream, ctx->peers[id].done));
    }
    return 0;
}

b200q_tensor bt(const ggml_tensor * t);
int compute_mul_mat(b200_backend_ctx * ctx, const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * b = op->src[1];
    if (tensor_is_split(a)) {
        ctx->packed_act = nullptr;
        return compute_mul_mat_split(ctx, op);
    }
    b200q_mul_mat_args  q{};
    q.type = (int) a->type;
    q.k = a->ne[0]; q.m = a->ne[1]; q.n = b->ne[1];
    q.ne02 = a->ne[2]; q.ne03 = a->ne[3]; q.ne12 = b->ne[2]; q.ne13 = b->ne[3];
    q.nb01 = a->nb[1]; q.nb02 = a->nb[2]; q.nb03 = a->nb[3];
    q.nb11 = b->nb[1]; q.nb12 = b->nb[2]; q.nb13 = b->nb[3];
    q.nb1 = op->nb[1]; q.nb2 = op->nb[2]; q.nb3 = op->nb[3];
    q.flags = weight_flags(a);
    if (ctx->packed_act == b && b->type == GGML_TYPE_F32) q.flags |= B200Q_FLAG_ACT_UNCHANGED;  // wq / wk / wv, gate / up: pack once
    ctx->packed_act = b->type == GGML_TYPE_F32 ? b : nullptr;
    size_t conv = 0;  // f16 activations: a contiguous f32 copy at the head of the workspace
    if (b->type == GGML_TYPE_F16) {
        conv   = ((size_t) ggml_nelements(b) * sizeof(float) + 255) & ~(size_t) 255;
        q.nb11 = (size_t) b->ne[0] * sizeof(float); q.nb12 = q.nb11 * b->ne[1]; q.nb13 = q.nb12 * b->ne[2];
    }
    const size_t need = b200q_mul_mat_workspace(&q);
    char *       ws   = need + conv ? (char *) backend_workspace(ctx, need + conv) : nullptr;
    const float * act = (const float *) b->data;
    if (conv) {
        const b200q_tensor src = bt(b);
        b200q_tensor       dst = src;
        dst.data = ws; dst.type = B200Q_TYPE_F32;
        dst.nb[0] = sizeof(float); dst.nb[1] = q.nb11; dst.nb[2] = q.nb12; dst.nb[3] = q.nb13;
        if (int rc = b200q_op_cpy(&src, &dst, ctx->stream)) return rc;
        act = (const float *) ws;
    }
    return b200q_mul_mat(&q, a->data, act, (float *) op->data, need ? ws + conv : nullptr, need ? ws_avail(ctx) - conv : 0, ctx->stream);
}
// `bias_node` != nullptr: the ADD_ID node behind this MUL_MAT_ID (same ids, the matmul's only consumer) — try both in one launch,
// writing into the ADD_ID's result; returns 1 when the pair has to go out separately
int compute_mul_mat_id(b200_backend_ctx * ctx, const ggml_tensor * op, const ggml_tensor * bias_node = nullptr) {
    const ggml_tensor *   as  = op->src[0];
    const ggml_tensor *   b   = op->src[1];
    const ggml_tensor *   ids = op->src[2];
    b200q_mul_mat_id_args q{};
    q.type = (int) as->type;
    q.k = as->ne[0]; q.m = as->ne[1]; q.n_expert = as->ne[2];
    q.n_used = ids->ne[0]; q.n_tokens = ids->ne[1]; q.b_ne1 = b->ne[1];
    q.nb01 = as->nb[1]; q.nb02 = as->nb[2];
    q.nb11 = b->nb[1]; q.nb12 = b->nb[2];
    q.ids_nb1 = ids->nb[1];
    q.nb1 = op->nb[1]; q.nb2 = op->nb[2];
    q.flags = weight_flags(as);
    const size_t need = b200q_mul_mat_id_workspace(&q);
    void *       ws   = need ? backend_workspace(ctx, need) : nullptr;
    if (bias_node) {
        const ggml_tensor * bias = bias_node->src[1];
        q.nb1 = bias_node->nb[1]; q.nb2 = bias_node->nb[2];
        const int rc = b200q_mul_mat_id_bias(&q, as->data, (const float *) b->data, (const int32_t *) ids->data, (const float *) bias->data,
                                             bias->nb[1], (float *) bias_node->data, ws, ws_avail(ctx), ctx->stream);
        return rc == B200Q_ENOTSUP ? 1 : rc;
    }
    return b200q_mul_mat_id(&q, as->data, (const float *) b->data, (const int32_t *) ids->data, (float *) op->data, ws,
                            ws_avail(ctx), ctx->stream);
}
int compute_get_rows(b200_backend_ctx * ctx, const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * r = op->src[1];
    b200q_get_rows_args q{};
    q.type = (int) a->type;
    q.ne00 = a->ne[0];
    q.ne10 = r->ne[0]; q.ne11 = r->ne[1]; q.ne12 = r->ne[2];
    q.nb01 = a->nb[1]; q.nb02 = a->nb[2]; q.nb03 = a->nb[3];
    q.nb10 = r->nb[0]; q.nb11 = r->nb[1]; q.nb12 = r->nb[2];
    q.nb1 = op->nb[1]; q.nb2 = op->nb[2]; q.nb3 = op->nb[3];
    return b200q_get_rows(&q, a->data, (const int32_t *) r->data, (float *) op->data, ctx->stream);
}

// ------------------------------------------------------------------------------------------------ backend
const char * backend_get_name(ggml_backend_t backend) { return ((b200_backend_ctx *) backend->context)->name.c_str(); }

void backend_free(ggml_backend_t backend) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    b200q_set_device(ctx->device);
    if (ctx->stream) {
        b200q_stream_synchronize(ctx->stream);
        if (ctx->workspace) b200q_free(ctx->workspace);
        if (ctx->copy_event) b200q_event_destroy(ctx->copy_event);
        if (ctx->split_start) b200q_event_destroy(ctx->split_start);
        b200q_stream_destroy(ctx->stream);
    }
    for (auto & kv : ctx->graphs) if (kv.second.exec) b200q_graph_destroy(kv.second.exec);
    for (auto & kv : ctx->lists) if (kv.second) b200q_list_destroy(kv.second);
    for (int id = 0; id < MAX_DEVICES; ++id) {
        b200_split_peer & pe = ctx->peers[id];
        if (!pe.stream) continue;
        b200q_set_device(registry_device_ctx(id)->device);
        b200q_stream_synchronize(pe.stream);
        if (pe.act) b200q_free(pe.act);
        if (pe.workspace) b200q_free(pe.workspace);
        b200q_event_destroy(pe.done);
        b200q_stream_destroy(pe.stream);
    }
    delete ctx;
    delete backend;
}
void backend_synchronize(ggml_backend_t backend) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    B200_CHECK(b200q_set_device(ctx->device));
    B200_CHECK(b200q_stream_synchronize(ctx->stream));
}
void backend_set_tensor_async(ggml_backend_t backend, ggml_tensor * tensor, const void * data, size_t offset, size_t size) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    B200_CHECK(b200q_set_device(ctx->device));
    B200_CHECK(b200q_memcpy((char *) tensor->data + offset, data, size, B200Q_H2D, ctx->stream));
}
void backend_get_tensor_async(ggml_backend_t backend, const ggml_tensor * tensor, void * data, size_t offset, size_t size) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    B200_CHECK(b200q_set_device(ctx->device));
    B200_CHECK(b200q_memcpy(data, (const char *) tensor->data + offset, size, B200Q_D2H, ctx->stream));
}

bool backend_is_b200(ggml_backend_t backend);

// device -> device copy between two streams of this plugin, ordered on both (the reference's semantics: copy on the source
// stream, the destination stream waits for it; ggml-cuda.cu:2556-2605)
bool backend_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const ggml_tensor * src, ggml_tensor * dst) {
    if (!backend_is_b200(backend_src) || !backend_is_b200(backend_dst)) return false;
    ggml_backend_buffer_t buf_src = src->view_src ? src->view_src->buffer : src->buffer;
    ggml_backend_buffer_t buf_dst = dst->view_src ? dst->view_src->buffer : dst->buffer;
    if (!buf_src || !buf_dst || !buffer_is_b200(buf_src) || !buffer_is_b200(buf_dst)) return false;
    auto * cs = (b200_backend_ctx *) backend_src->context;
    auto * cd = (b200_backend_ctx *) backend_dst->context;
    if (cs->device != ((b200_buffer_ctx *) buf_src->context)->device || cd->device != ((b200_buffer_ctx *) buf_dst->context)->device) return false;
    B200_CHECK(b200q_set_device(cs->device));
    if (backend_src == backend_dst) {
        B200_CHECK(b200q_memcpy(dst->data, src->data, ggml_nbytes(dst), B200Q_D2D, cs->stream));
        return true;
    }
    if (cs->device == cd->device) {
        B200_CHECK(b200q_memcpy(dst->data, src->data, ggml_nbytes(dst), B200Q_D2D, cs->stream));
    } else {
        B200_CHECK(b200q_memcpy_peer(dst->data, cd->device, src->data, cs->device, ggml_nbytes(dst), cs->stream));
    }
    if (!cs->copy_event) B200_CHECK(b200q_event_create(&cs->copy_event));
    B200_CHECK(b200q_event_record(cs->copy_event, cs->stream));
    B200_CHECK(b200q_set_device(cd->device));
    B200_CHECK(b200q_stream_wait_event(cd->stream, cs->copy_event));
    return true;
}
void backend_event_record(ggml_backend_t backend, ggml_backend_event_t event) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    B200_CHECK(b200q_event_record(event->context, ctx->stream));
}
void backend_event_wait(ggml_backend_t backend, ggml_backend_event_t event) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    B200_CHECK(b200q_set_device(ctx->device));
    B200_CHECK(b200q_stream_wait_event(ctx->stream, event->context));
}

// ---- runs of decode matmuls as ONE persistent launch (csrc/gemv_list.cu) -------------------------------------------------
// A MUL_MAT node qualifies when it is a plain n = 1 product on immutable, dense, unsplit weights; a run = consecutive such
// nodes (no-op view nodes between them are skipped).  wq / wk / wv and ffn_gate / ffn_up of a llama graph are such runs:
// they share one quantized activation image and one launch.  A node whose activations overlap the result of an earlier
// node of the run is flagged B200Q_LIST_DEP (device-side barrier point).  Results are bit-identical to the per-node path.
bool noop_node(const ggml_tensor * t) {
    return ggml_is_empty(t) || t->op == GGML_OP_NONE || t->op == GGML_OP_VIEW || t->op == GGML_OP_RESHAPE || t->op == GGML_OP_PERMUTE ||
           t->op == GGML_OP_TRANSPOSE;
}
bool list_candidate(const ggml_tensor * op) {
    if (op->op != GGML_OP_MUL_MAT) return false;
    const ggml_tensor * a = op->src[0];
    const ggml_tensor * b = op->src[1];
    if (!a || !b || tensor_is_split(a) || !(weight_flags(a) & B200Q_FLAG_WEIGHTS_IMMUTABLE)) return false;
    if (a->ne[2] != 1 || a->ne[3] != 1 || ggml_nrows(b) != 1 || b->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
    if (dense_weight(a->type) || !type_ok(a->type) || a->nb[1] != ggml_row_size(a->type, a->ne[0])) return false;
    if (b->nb[0] != 4 || op->nb[0] != 4 || ((uintptr_t) a->data & 15) || ((uintptr_t) b->data & 15)) return false;
    return true;
}
// Tries to run the decode matmuls that start at node `first` as ONE persistent list.  Members are the consecutive candidates
// and — llama.cpp emits q, rope(q), k, rope(k), v: matmuls of one vector separated by other nodes — candidates up to a few
// nodes further that multiply a vector an earlier member already multiplies.  Such a member runs EARLIER than its place in
// the graph, which is only legal if nothing in between touches its result's memory (ggml-alloc reuses buffers): its dst range
// must not overlap the dst or any source of a skipped node or of an earlier member.  Marks the members in `done` and returns
// how many there are (0: no list was launched, nothing marked).
bool ranges_overlap(const ggml_tensor * a, const ggml_tensor * b) {
    if (!a || !b || !a->data || !b->data) return false;
    const char * a0 = (const char *) a->data, * a1 = a0 + ggml_nbytes(a);
    const char * b0 = (const char *) b->data, * b1 = b0 + ggml_nbytes(b);
    return a0 < b1 && b0 < a1;
}
bool touches(const ggml_tensor * node, const ggml_tensor * t) {  // does `node` read or write memory of t's result?
    if (ranges_overlap(node, t)) return true;
    for (int j = 0; j < GGML_MAX_SRC; ++j) if (node->src[j] && ranges_overlap(node->src[j], t)) return true;
    return false;
}
int32_t node_use_count(const ggml_cgraph * cgraph, const ggml_tensor * t);
// how often each fusion of graph_dispatch fired ("ggml_backend_b200_fusion_stats": copies up to n counters, returns how many exist)
enum { FUSE_NORM, FUSE_ATTN, FUSE_ROPE_STORE, FUSE_SET_ROWS_PAIR, FUSE_MOE_SUM, FUSE_ID_BIAS, FUSE_LIST, FUSE_SCRATCH_READ, FUSE_SCRATCH_COPY, FUSE_GLU_LIST, FUSE_ROUTE, FUSE_N };
int64_t g_fusion_stats[FUSE_N] = {};
void    print_fusion_stats() {  // GGML_B200_STATS=1: at exit, what graph_dispatch did with the graphs of this process
    static const char * names[FUSE_N] = { "norm", "attn", "rope_store", "set_rows_pair", "moe_sum", "id_bias", "list", "scratch_read", "scratch_copy", "glu_list", "route" };
    fprintf(stderr, "ggml-b200 fusions:");
    for (int i = 0; i < FUSE_N; ++i) fprintf(stderr, " %s=%lld", names[i], (long long) g_fusion_stats[i]);
    fprintf(stderr, "\n");
}
int     fusion_stats(int64_t * out, int n) {
    for (int i = 0; i < n && i < FUSE_N; ++i) out[i] = g_fusion_stats[i];
    return FUSE_N;
}
int compute_list_run(b200_backend_ctx * ctx, ggml_cgraph * cgraph, int first, int n_nodes, std::vector<char> & done) {
    std::vector<b200q_list_op> ops;
    std::vector<int>           members, skipped, absorbed;
    std::vector<void *>        scratch_of;
    // measured (tools/ffn_list_bench.py, llama-bench): every CTA evaluating the whole 14336-value SWIGLU behind the barrier point costs
    // 4.8 us per layer, a separate GLU launch overlapped by programmatic dependent launch 3: tg128 409 -> 392 t/s.  Opt-in only.
    static const bool glu_list = getenv("GGML_B200_GLU_LIST") != nullptr && getenv("GGML_B200_NO_FUSE") == nullptr;
    const size_t               hoist_mark = ctx->hoist_used;
    static const bool no_hoist   = getenv("GGML_B200_NO_HOIST") != nullptr;
    static const bool no_scratch = getenv("GGML_B200_NO_SCRATCH") != nullptr;
    for (int i = first; i < n_nodes && ops.size() < 64; ++i) {
        ggml_tensor * t = ggml_graph_node(cgraph, i);
        if (noop_node(t) || done[(size_t) i]) continue;
        if (!list_candidate(t)) {
            if (no_hoist || skipped.size() >= 6) break;
            skipped.push_back(i);
            continue;
        }
        // ffn_down behind SWIGLU(ffn_gate, ffn_up), both of them members of this run: the GLU is computed while the activations
        // are quantized (B200Q_LIST_SWIGLU) and its node, which nobody else reads, never runs
        const ggml_tensor * glu = nullptr;
        if (glu_list && !skipped.empty() && cgraph->use_counts) {
            const ggml_tensor * g = ggml_graph_node(cgraph, skipped.back());
            const auto member_result = [&](const ggml_tensor * x) {
                for (size_t q = 0; x && q < members.size(); ++q) if (ggml_graph_node(cgraph, members[q]) == x && !scratch_of[q]) return true;
                return false;
            };
            if (g->op == GGML_OP_GLU && ggml_get_glu_op(g) == GGML_GLU_OP_SWIGLU && g->src[1] && t->src[1] == g && supports_elementwise(g) &&
                member_result(g->src[0]) && member_result(g->src[1]) && ggml_is_contiguous(g->src[0]) && ggml_is_contiguous(g->src[1]) &&
                node_use_count(cgraph, g) == 1 && !(g->flags & GGML_TENSOR_FLAG_OUTPUT) && !ranges_overlap(t, g->src[0]) && !ranges_overlap(t, g->src[1]) &&
                !(((uintptr_t) g->src[0]->data | (uintptr_t) g->src[1]->data) & 15)) {
                glu = g;
                absorbed.push_back(skipped.back());
                skipped.pop_back();
            }
        }
        void * scratch = nullptr;
        if (!skipped.empty()) {
            bool ok = false, aliased = false;
            ok = glu != nullptr;
            for (int mi : members) ok = ok || ggml_graph_node(cgraph, mi)->src[1]->data == t->src[1]->data;  // the vector is ready
            for (int si : skipped) ok = ok && !ranges_overlap(ggml_graph_node(cgraph, si), t->src[1]);      // ... and stays untouched
            for (int si : skipped) aliased = aliased || touches(ggml_graph_node(cgraph, si), t);
            for (int mi : members) aliased = aliased || touches(ggml_graph_node(cgraph, mi), t);
            if (!ok) break;
            if (aliased) {  // its buffer is somebody else's until its own place in the graph: compute into scratch, copy there
                const size_t bytes = ((size_t) t->src[0]->ne[1] * sizeof(float) + 255) & ~(size_t) 255;
                if (no_scratch || !ctx->hoist_ok || !ggml_is_contiguous(t) || ctx->ws_bytes < HOIST_BYTES || ctx->hoist_used + bytes > HOIST_BYTES) break;
                scratch = (char *) ctx->workspace + ctx->hoist_used;
                ctx->hoist_used += bytes;
            }
        }
        b200q_list_op o{};
        o.type = (int) t->src[0]->type; o.m = t->src[0]->ne[1]; o.k = t->src[0]->ne[0];
        o.w = t->src[0]->data; o.act = (const float *) t->src[1]->data; o.dst = scratch ? (float *) scratch : (float *) t->data;
        if (glu) {
            o.act = (const float *) glu->src[0]->data; o.act2 = (const float *) glu->src[1]->data;
            o.flags |= B200Q_LIST_SWIGLU;
        }
        scratch_of.push_back(scratch);
        const char * a0 = (const char *) o.act, * a1 = a0 + o.k * 4;
        for (const b200q_list_op & e : ops) {  // reads what an earlier op of the run writes?
            const char * d0 = (const char *) e.dst, * d1 = d0 + e.m * 4;
            if (a0 < d1 && d0 < a1) o.flags |= B200Q_LIST_DEP;
        }
        ops.push_back(o);
        members.push_back(i);
    }
    if (ops.size() < 2) {
        ctx->hoist_used = hoist_mark;
        return 0;
    }
    std::vector<int64_t> key;
    key.reserve(ops.size() * 6);
    for (const b200q_list_op & o : ops) {
        key.push_back(o.type); key.push_back(o.m); key.push_back(o.k);
        key.push_back((int64_t) (uintptr_t) o.w); key.push_back((int64_t) (uintptr_t) o.act);
        key.push_back((int64_t) (uintptr_t) o.dst * 2 + (o.flags & B200Q_LIST_DEP ? 1 : 0));
        if (o.flags & B200Q_LIST_SWIGLU) key.push_back((int64_t) (uintptr_t) o.act2);
    }
    auto it = ctx->lists.find(key);
    if (it == ctx->lists.end()) {
        b200q_list * l = nullptr;
        if (b200q_list_create(ops.data(), (int) ops.size(), nullptr, &l) != 0) {
            l = nullptr;  // some op is not served by the list kernel: remember that, the nodes go out one by one
        }
        if (ctx->lists.size() >= MAX_CACHED_LISTS) {  // evict the oldest entry
            auto old = ctx->lists.find(ctx->list_order.front());
            if (old != ctx->lists.end()) {
                if (old->second) {
                    drop_graphs(ctx);
                    B200_CHECK(b200q_stream_synchronize(ctx->stream));
                    b200q_list_destroy(old->second);
                }
                ctx->lists.erase(old);
            }
            ctx->list_order.erase(ctx->list_order.begin());
        }
        it = ctx->lists.emplace(key, l).first;
        ctx->list_order.push_back(key);
    }
    if (!it->second || b200q_list_launch(it->second, ctx->stream) != 0) {
        ctx->hoist_used = hoist_mark;
        return 0;
    }
    for (size_t q = 0; q < members.size(); ++q) {
        const int mi = members[q];
        done[(size_t) mi] = 1;
        if (scratch_of[q]) {
            done[(size_t) mi] = 2;  // delivered by a copy at its own position
            ctx->hoist_copies[(size_t) mi].src   = scratch_of[q];
            ctx->hoist_copies[(size_t) mi].bytes = (size_t) ggml_graph_node(cgraph, mi)->src[0]->ne[1] * sizeof(float);
        }
    }
    for (int ai : absorbed) done[(size_t) ai] = 1;
    g_fusion_stats[FUSE_GLU_LIST] += (int64_t) absorbed.size();
    return (int) members.size();
}

ggml_status graph_dispatch(b200_backend_ctx * ctx, ggml_cgraph * cgraph);
// statistics for tests and tuning (get_proc_address "ggml_backend_b200_graph_stats"): graphs recorded, graphs replayed
int64_t g_graph_stats[2] = { 0, 0 };
void    graph_stats(int64_t * recorded, int64_t * replayed) { *recorded = g_graph_stats[0]; *replayed = g_graph_stats[1]; }

// everything a launch bakes into its arguments: op, types, shapes, strides, addresses, parameters of every node and source
void graph_fingerprint(const ggml_cgraph * cgraph, std::vector<uint64_t> & key, bool * capturable) {
    const int n_nodes = ggml_graph_n_nodes((ggml_cgraph *) cgraph);
    key.clear();
    key.reserve((size_t) n_nodes * 24);
    *capturable = true;
    auto put_tensor = [&](const ggml_tensor * t) {
        key.push_back((uint64_t) (uintptr_t) t->data);
        key.push_back((uint64_t) t->type);
        for (int d = 0; d < 4; ++d) { key.push_back((uint64_t) t->ne[d]); key.push_back((uint64_t) t->nb[d]); }
    };
    for (int i = 0; i < n_nodes; ++i) {
        const ggml_tensor * node = ggml_graph_node((ggml_cgraph *) cgraph, i);
        key.push_back((uint64_t) node->op);
        put_tensor(node);
        uint64_t op[GGML_MAX_OP_PARAMS / sizeof(uint64_t)];
        memcpy(op, node->op_params, sizeof(op));
        for (uint64_t v : op) key.push_back(v);
        for (int j = 0; j < GGML_MAX_SRC; ++j) {
            if (!node->src[j]) { key.push_back(0); continue; }
            put_tensor(node->src[j]);
            if (tensor_is_split(node->src[j])) *capturable = false;  // other devices' streams and peer copies: eager
        }
    }
}

// asynchronous: kernels are enqueued on the backend's stream, the caller follows with synchronize()
// (ggml-backend.cpp:327-335).  Graph and tensor pointers are not retained (they may be stack temporaries).
ggml_status backend_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    auto * ctx = (b200_backend_ctx *) backend->context;
    B200_CHECK(b200q_set_device(ctx->device));
    // graphs of a few nodes (test harnesses, prompt chunks that change every call) are not worth a capture
    const bool try_graph = !ctx->graphs_off && ggml_graph_n_nodes(cgraph) >= 16 && getenv("GGML_B200_NO_GRAPHS") == nullptr;
    if (!try_graph) return graph_dispatch(ctx, cgraph);
    std::vector<uint64_t> key;
    bool capturable = true;
    graph_fingerprint(cgraph, key, &capturable);
    if (!capturable) return graph_dispatch(ctx, cgraph);
    auto it = ctx->graphs.find(key);
    if (it == ctx->graphs.end()) {  // first sighting: eager (grows workspaces, creates the persistent lists)
        if (ctx->graphs.size() >= MAX_CACHED_GRAPHS) {
            auto old = ctx->graphs.find(ctx->graph_order.front());
            if (old != ctx->graphs.end()) {
                if (old->second.exec) {
                    // the graph may still be running on the stream
                    B200_CHECK(b200q_stream_synchronize(ctx->stream));
                    b200q_graph_destroy(old->second.exec);
                }
                ctx->graphs.erase(old);
            }
            ctx->graph_order.erase(ctx->graph_order.begin());
        }
        b200_backend_ctx::graph_entry e;
        e.seen = 1;
        ctx->graphs.emplace(key, e);
        ctx->graph_order.push_back(std::move(key));
        return graph_dispatch(ctx, cgraph);
    }
    if (!it->second.exec) {  // second sighting: record
        if (b200q_graph_begin(ctx->stream) != 0) { ctx->graphs_off = true; return graph_dispatch(ctx, cgraph); }
        const ggml_status st = graph_dispatch(ctx, cgraph);
        void * exec = nullptr;
        const int rc = b200q_graph_end(ctx->stream, &exec);
        if (st != GGML_STATUS_SUCCESS || rc != 0 || !exec) {
            fprintf(stderr, "ggml-b200: CUDA-graph capture failed (%s); staying eager\n", b200q_last_error());
            if (exec) b200q_graph_destroy(exec);
            ctx->graphs_off = true;
            return graph_dispatch(ctx, cgraph);
        }
        it = ctx->graphs.find(key);  // (the dispatch may have dropped the cache)
        if (it == ctx->graphs.end()) {
            b200q_graph_destroy(exec);
            return graph_dispatch(ctx, cgraph);
        }
        it->second.exec = exec;
        ++g_graph_stats[0];
    } else {
        ++g_graph_stats[1];
    }
    ++it->second.seen;
    return b200q_graph_launch(it->second.exec, ctx->stream) == 0 ? GGML_STATUS_SUCCESS : GGML_STATUS_FAILED;
}

// [ADD ->] RMS_NORM -> MUL as one launch (b200q_op_norm_fused; the reference fuses the same run, ggml-cuda.cu:2811-2845,
// 3000-3040).  Returns the number of nodes consumed (0: no fusion here).  The norm's own output is never written, so it
// must have the MUL as its only consumer and be neither a view source nor a graph output (ggml_node_has_n_uses).
b200q_tensor bt_src(b200_backend_ctx * ctx, const ggml_tensor * t);
int try_fuse_norm(b200_backend_ctx * ctx, ggml_cgraph * cgraph, int i, int n_nodes, int * rc) {
    auto f32row = [](const ggml_tensor * t) { return t && t->type == GGML_TYPE_F32 && t->nb[0] == sizeof(float); };
    const ggml_tensor * add = nullptr;
    int j = i;
    if (ggml_graph_node(cgraph, j)->op == GGML_OP_ADD) {
        add = ggml_graph_node(cgraph, j);
        ++j;
    }
    if (j + 1 >= n_nodes) return 0;
    const ggml_tensor * norm = ggml_graph_node(cgraph, j);
    const ggml_tensor * mul  = ggml_graph_node(cgraph, j + 1);
    if (norm->op != GGML_OP_RMS_NORM || mul->op != GGML_OP_MUL || !cgraph->use_counts) return 0;
    if (add && norm->src[0] != add) return 0;
    if (mul->src[0] != norm && mul->src[1] != norm) return 0;
    if (!ggml_node_has_n_uses(cgraph, j, 1)) return 0;
    const ggml_tensor * x = norm->src[0];
    const ggml_tensor * w = mul->src[0] == norm ? mul->src[1] : mul->src[0];
    if (!f32row(x) || !f32row(norm) || !f32row(mul) || !f32row(w) || x->ne[0] > 8192 || w->ne[0] != x->ne[0]) return 0;
    if (!ggml_are_same_shape(x, norm) || !ggml_are_same_shape(norm, mul)) return 0;
    for (int d = 1; d < 4; ++d) if (x->ne[d] % w->ne[d]) return 0;
    if (tensor_is_split(w) || tensor_is_split(x)) return 0;
    float eps;
    memcpy(&eps, norm->op_params, sizeof(float));
    const b200q_tensor tw = bt(w), td = bt(mul);
    if (add) {
        const ggml_tensor * a = add->src[0];
        const ggml_tensor * b = add->src[1];
        if (!f32row(a) || !f32row(b) || !f32row(add) || !ggml_are_same_shape(a, add) || !ggml_are_same_shape(b, add)) return 0;
        const b200q_tensor ta = bt_src(ctx, a), tb = bt(b), ts = bt(add);  // a: possibly a hoisted matmul's result still in the scratch
        *rc = b200q_op_norm_fused(&ta, &tb, &ts, &tw, &td, eps, ctx->stream);
        return 3;
    }
    const b200q_tensor tx = bt(x);
    *rc = b200q_op_norm_fused(&tx, nullptr, nullptr, &tw, &td, eps, ctx->stream);
    return 2;
}

// The attention core of a decode step — MUL_MAT(k, q) -> SOFT_MAX -> MUL_MAT(v, kq) -> PERMUTE -> CONT as llama.cpp builds it
// without flash attention (src/llama-graph.cpp:1285-1330) — as ONE launch (b200q_op_attn_decode) when there are few tokens and a
// short cache: four ~3.5 us launches per layer otherwise.  Every intermediate result must have the next node of the run as
// its only consumer.  Returns the index behind the run (i: no fusion).
int32_t node_use_count(const ggml_cgraph * cgraph, const ggml_tensor * t) {
    const size_t pos = ggml_hash_find(&cgraph->visited_hash_set, t);
    if (pos == GGML_HASHSET_FULL || !ggml_bitset_get(cgraph->visited_hash_set.used, pos)) return -1;
    return cgraph->use_counts[pos];
}
int try_fuse_attn(b200_backend_ctx * ctx, ggml_cgraph * cgraph, int i, int n_nodes, int * rc) {
    if (!cgraph->use_counts) return i;
    auto next = [&](int j) {
        for (++j; j < n_nodes; ++j) {
            const ggml_tensor * t = ggml_graph_node(cgraph, j);
            if (t->op != GGML_OP_VIEW && t->op != GGML_OP_RESHAPE && t->op != GGML_OP_NONE) return j;  // PERMUTE / TRANSPOSE are matched below
        }
        return n_nodes;
    };
    const ggml_tensor * kq = ggml_graph_node(cgraph, i);
    const ggml_tensor * k = kq->src[0], * q = kq->src[1];
    if (!k || !q || k->type != GGML_TYPE_F16 || q->type != GGML_TYPE_F32 || kq->type != GGML_TYPE_F32 || k->nb[0] != 2 || q->nb[0] != 4) return i;
    if (q->ne[3] != 1 || k->ne[3] != 1 || q->ne[1] > 8 || k->ne[1] > 2048 || k->ne[0] != q->ne[0] || q->ne[0] > 512 || q->ne[2] % k->ne[2]) return i;
    if (tensor_is_split(k)) return i;
    const int j1 = next(i);
    if (j1 >= n_nodes) return i;
    const ggml_tensor * sm = ggml_graph_node(cgraph, j1);
    float scale, max_bias;
    memcpy(&scale, (const float *) sm->op_params + 0, sizeof(float));
    memcpy(&max_bias, (const float *) sm->op_params + 1, sizeof(float));
    if (sm->op != GGML_OP_SOFT_MAX || sm->src[0] != kq || max_bias != 0.0f || !ggml_is_contiguous(kq) || !ggml_is_contiguous(sm)) return i;
    const ggml_tensor * mask = sm->src[1], * sinks = sm->src[2];
    if (mask && ((mask->type != GGML_TYPE_F32 && mask->type != GGML_TYPE_F16) || mask->ne[0] < k->ne[1] || mask->ne[1] < q->ne[1] ||
                 mask->nb[0] != ggml_type_size(mask->type) || mask->ne[3] != 1 || q->ne[2] % mask->ne[2])) return i;
    if (sinks && (sinks->type != GGML_TYPE_F32 || sinks->ne[0] < q->ne[2] || sinks->nb[0] != 4)) return i;
    const int j2 = next(j1);
    if (j2 >= n_nodes) return i;
    const ggml_tensor * kqv = ggml_graph_node(cgraph, j2);
    const ggml_tensor * v = kqv->src[0];
    if (kqv->op != GGML_OP_MUL_MAT || kqv->src[1] != sm || !v || v->type != GGML_TYPE_F16 || v->nb[0] != 2 || v->ne[0] != k->ne[1] ||
        v->ne[1] != q->ne[0] || v->ne[2] != k->ne[2] || v->ne[3] != 1 || kqv->type != GGML_TYPE_F32 || !ggml_is_contiguous(kqv)) return i;
    const int j3 = next(j2);
    const int j4 = j3 < n_nodes ? next(j3) : n_nodes;
    if (j4 >= n_nodes) return i;
    const ggml_tensor * pm = ggml_graph_node(cgraph, j3);
    const ggml_tensor * ct = ggml_graph_node(cgraph, j4);
    if (pm->op != GGML_OP_PERMUTE || pm->src[0] != kqv || ct->op != GGML_OP_CONT || ct->src[0] != pm || ct->type != GGML_TYPE_F32 ||
        !ggml_is_contiguous(ct)) return i;
    // permute (0, 2, 1, 3): [hd, n_tok, n_head] -> [hd, n_head, n_tok]
    if (pm->ne[0] != kqv->ne[0] || pm->ne[1] != kqv->ne[2] || pm->ne[2] != kqv->ne[1] || pm->nb[1] != kqv->nb[2] || pm->nb[2] != kqv->nb[1] ||
        ggml_nelements(ct) != ggml_nelements(kqv)) return i;
    if (node_use_count(cgraph, kq) != 1 || node_use_count(cgraph, sm) != 1 || node_use_count(cgraph, kqv) != 1 || node_use_count(cgraph, pm) != 1) return i;
    for (const ggml_tensor * t : { kq, sm, kqv, pm }) if (t->flags & GGML_TENSOR_FLAG_OUTPUT) return i;
    const b200q_tensor tq = bt(q), tk = bt(k), tv = bt(v);
    b200q_tensor tm{}, ts{}, td{};
    if (mask) tm = bt(mask);
    if (sinks) ts = bt(sinks);
    td.data = ct->data; td.type = B200Q_TYPE_F32;
    td.ne[0] = q->ne[0]; td.ne[1] = q->ne[2]; td.ne[2] = q->ne[1]; td.ne[3] = 1;
    td.nb[0] = 4; td.nb[1] = (size_t) q->ne[0] * 4; td.nb[2] = td.nb[1] * (size_t) q->ne[2]; td.nb[3] = td.nb[2] * (size_t) q->ne[1];
    *rc = b200q_op_attn_decode(&tq, &tk, &tv, mask ? &tm : nullptr, sinks ? &ts : nullptr, &td, scale, ctx->stream);
    return j4 + 1;
}

// The MoE combine — MUL(experts [n, n_used, n_tok], weights [1, n_used, n_tok]), one VIEW per used expert, a chain of n_used - 1
// ADDs (src/llama-graph.cpp build_moe_ffn) — as ONE launch (b200q_op_moe_weighted_sum).  The MUL's result is never written: its
// only consumers must be the n_used views, every view and every ADD but the last must have exactly one consumer.
// Returns the index behind the run (i: no fusion).
int try_fuse_moe_sum(b200_backend_ctx * ctx, ggml_cgraph * cgraph, int i, int n_nodes, int * rc) {
    if (!cgraph->use_counts) return i;
    const ggml_tensor * mul = ggml_graph_node(cgraph, i);
    const ggml_tensor * e = mul->src[0], * w = mul->src[1];
    auto f32c = [](const ggml_tensor * t) { return t && t->type == GGML_TYPE_F32 && t->nb[0] == sizeof(float); };
    if (!f32c(e) || !f32c(w) || !f32c(mul) || !ggml_are_same_shape(e, mul) || !ggml_is_contiguous(mul)) return i;
    const int64_t n = e->ne[0], n_used = e->ne[1], n_tok = e->ne[2];
    if (n_used < 2 || n_used > 16 || e->ne[3] != 1 || w->ne[0] != 1 || w->ne[1] != n_used || w->ne[2] != n_tok || w->ne[3] != 1) return i;
    if (node_use_count(cgraph, mul) != (int32_t) n_used || (mul->flags & GGML_TENSOR_FLAG_OUTPUT)) return i;
    // the views (in slot order, possibly interleaved with other no-op nodes), then the adds
    std::vector<const ggml_tensor *> views;
    int j = i + 1;
    for (; j < n_nodes && (int64_t) views.size() < n_used; ++j) {
        const ggml_tensor * t = ggml_graph_node(cgraph, j);
        if (t->op != GGML_OP_VIEW) return i;
        const size_t slot = views.size();
        if (t->src[0] != mul || t->view_src != mul || t->view_offs != slot * mul->nb[1] || t->ne[0] != n || t->ne[1] != n_tok || t->ne[2] != 1 ||
            t->nb[1] != mul->nb[2] || node_use_count(cgraph, t) != 1) return i;
        views.push_back(t);
    }
    if ((int64_t) views.size() != n_used) return i;
    const ggml_tensor * prev = views[0];
    for (int64_t sidx = 1; sidx < n_used; ++sidx, ++j) {
        if (j >= n_nodes) return i;
        const ggml_tensor * a = ggml_graph_node(cgraph, j);
        if (a->op != GGML_OP_ADD || a->src[0] != prev || a->src[1] != views[(size_t) sidx] || !f32c(a) || !ggml_is_contiguous(a) ||
            a->ne[0] != n || a->ne[1] != n_tok) return i;
        if (sidx < n_used - 1 && (node_use_count(cgraph, a) != 1 || (a->flags & GGML_TENSOR_FLAG_OUTPUT) || a->view_src)) return i;
        prev = a;
    }
    const b200q_tensor te = bt(e), tw = bt(w), td = bt(prev);
    *rc = b200q_op_moe_weighted_sum(&te, &tw, &td, ctx->stream);
    return j;
}

// Is the result of node i read by exactly one operator, a ROPE or SET_ROWS within the next nodes, through a chain of
// single-use views that keep its address (RESHAPE / VIEW at offset 0)?  Then that operator can read the values from the
// scratch they were computed into and the delivery copy is dropped.
bool sole_reader_takes_scratch(const ggml_cgraph * cgraph, int i, int n_nodes) {
    if (!cgraph->use_counts) return false;
    const ggml_tensor * cur = ggml_graph_node((ggml_cgraph *) cgraph, i);
    const void *        data = cur->data;
    if (cur->flags & GGML_TENSOR_FLAG_OUTPUT) return false;
    for (int hops = 0; hops < 6; ++hops) {
        if (node_use_count(cgraph, cur) != 1) return false;
        const ggml_tensor * user = nullptr;
        for (int j = i + 1; j < n_nodes && j < i + 32 && !user; ++j) {
            const ggml_tensor * t = ggml_graph_node((ggml_cgraph *) cgraph, j);
            for (int q = 0; q < GGML_MAX_SRC; ++q) if (t->src[q] == cur) user = t;
        }
        if (!user) return false;
        if (user->op == GGML_OP_RESHAPE || user->op == GGML_OP_VIEW) {
            if (user->data != data || (user->flags & GGML_TENSOR_FLAG_OUTPUT)) return false;
            cur = user;
            continue;
        }
        return (user->op == GGML_OP_ROPE || user->op == GGML_OP_SET_ROWS || user->op == GGML_OP_ADD) && user->src[0] == cur && user->src[1] != cur &&
               supports_elementwise(user);
    }
    return false;
}

// The MoE router behind the logits matmul as gpt-oss gates (src/llama-graph.cpp build_moe_ffn, router bias +
// LLAMA_EXPERT_GATING_FUNC_TYPE_SOFTMAX_WEIGHT) is ADD(logits, bias) -> ARGSORT(desc) -> VIEW (top-k) ... GET_ROWS(RESHAPE(biased),
// top-k) -> RESHAPE -> SOFT_MAX, with the expert matmuls between the two halves in graph order.  Each half is one launch:
// ADD + ARGSORT (both results written), and GET_ROWS + SOFT_MAX (the GET_ROWS result, which only the SOFT_MAX reads, is not).
// Both return the index behind the run, or 0.
int try_fuse_moe_route(b200_backend_ctx * ctx, const ggml_cgraph * cgraph, int i, int n_nodes, int * rc) {
    const ggml_tensor * add = ggml_graph_node((ggml_cgraph *) cgraph, i);
    const ggml_tensor * lg = add->src[0], * bias = add->src[1];
    if (add->op != GGML_OP_ADD || !is_f32(add) || !is_f32(lg) || !is_f32(bias) || !ggml_is_contiguous(add) || !ggml_is_contiguous(lg) ||
        !ggml_is_contiguous(bias) || !ggml_are_same_shape(add, lg) || add->ne[0] > 1024 || add->ne[2] != 1 || add->ne[3] != 1 || bias->ne[0] != add->ne[0] ||
        ggml_nelements(bias) != bias->ne[0] || tensor_is_split(bias)) return 0;
    int j = i + 1;
    while (j < n_nodes && noop_node(ggml_graph_node((ggml_cgraph *) cgraph, j))) ++j;
    if (j >= n_nodes) return 0;
    const ggml_tensor * srt = ggml_graph_node((ggml_cgraph *) cgraph, j);
    if (srt->op != GGML_OP_ARGSORT || srt->src[0] != add || ((const int32_t *) srt->op_params)[0] != GGML_SORT_ORDER_DESC || !supports_elementwise(srt) ||
        !ggml_is_contiguous(srt) || ranges_overlap(srt, lg) || ranges_overlap(srt, add)) return 0;
    const b200q_tensor tl = bt_src(ctx, lg), tb = bt(bias), ta = bt(add), ts = bt(srt);
    *rc = b200q_op_moe_route(&tl, &tb, &ta, &ts, nullptr, ctx->stream);
    return j + 1;
}
int try_fuse_gate_weights(b200_backend_ctx * ctx, const ggml_cgraph * cgraph, int i, int n_nodes, int * rc) {
    if (!cgraph->use_counts) return 0;
    const ggml_tensor * gr = ggml_graph_node((ggml_cgraph *) cgraph, i);
    const ggml_tensor * probs = gr->src[0], * ids = gr->src[1];
    if (gr->op != GGML_OP_GET_ROWS || !is_f32(gr) || !is_f32(probs) || !ids || ids->type != GGML_TYPE_I32 || probs->ne[0] != 1 || probs->ne[3] != 1 ||
        ids->ne[2] != 1 || ids->ne[3] != 1 || ids->ne[1] != probs->ne[2] || ids->ne[0] > 1024 || gr->ne[0] != 1 || gr->ne[1] != ids->ne[0] ||
        gr->ne[2] != ids->ne[1] || node_use_count(cgraph, gr) != 1 || (gr->flags & GGML_TENSOR_FLAG_OUTPUT) || tensor_is_split(probs)) return 0;
    int j = i + 1;
    const ggml_tensor * view = nullptr;
    while (j < n_nodes && noop_node(ggml_graph_node((ggml_cgraph *) cgraph, j))) {
        const ggml_tensor * t = ggml_graph_node((ggml_cgraph *) cgraph, j);
        if (t->src[0] == gr) view = t;
        ++j;
    }
    if (j >= n_nodes || !view) return 0;
    const ggml_tensor * sm = ggml_graph_node((ggml_cgraph *) cgraph, j);
    float scale, max_bias;
    memcpy(&scale, sm->op_params, 4); memcpy(&max_bias, (const char *) sm->op_params + 4, 4);
    if (view->op != GGML_OP_RESHAPE || view->data != gr->data || node_use_count(cgraph, view) != 1 || (view->flags & GGML_TENSOR_FLAG_OUTPUT)) return 0;
    if (sm->op != GGML_OP_SOFT_MAX || sm->src[0] != view || sm->src[1] || sm->src[2] || scale != 1.0f || max_bias != 0.0f || !is_f32(sm) ||
        !ggml_is_contiguous(sm) || sm->ne[0] != ids->ne[0] || sm->ne[1] != ids->ne[1] || sm->ne[2] != 1 || sm->ne[3] != 1 || ranges_overlap(sm, probs) ||
        ranges_overlap(sm, ids)) return 0;
    const b200q_tensor tp = bt(probs), ti = bt(ids), tw = bt(sm);
    *rc = b200q_op_moe_gate_weights(&tp, &ti, &tw, ctx->stream);
    return j + 1;
}

// [ADD(bias) ->] ROPE(q) at node i, followed — with nothing in between but views and matmuls that already ran with an earlier
// list — by [ADD(bias) ->] ROPE(k) whose only reader is a SET_ROWS (the K cache write) and [ADD(bias) ->] a second SET_ROWS (the V
// cache write): llama.cpp's decode step behind the q / k / v matmuls (src/models/*.cpp: with biases in gpt-oss / Qwen2).  All of
// it goes out as one launch (b200q_op_rope_store) at the first node's position: every source is already computed, and the cache
// rows are written a few nodes early past nodes that do not execute.  Matmul results that wait in the scratch are read from
// there; the ADD results and the rotated k are never written (each has the next node of the run as its only reader).
// Returns the index behind the group, or 0.
int try_fuse_rope_store(b200_backend_ctx * ctx, const ggml_cgraph * cgraph, int i, int n_nodes, const std::vector<char> & done, int * rc) {
    if (!cgraph->use_counts) return 0;
    auto producer = [&](const ggml_tensor * t) {  // the node behind views that keep its address
        for (int hops = 0; hops < 4 && (t->op == GGML_OP_RESHAPE || t->op == GGML_OP_VIEW) && t->src[0] && t->src[0]->data == t->data; ++hops) {
            if (node_use_count(cgraph, t) != 1 || (t->flags & GGML_TENSOR_FLAG_OUTPUT)) return (const ggml_tensor *) nullptr;
            t = t->src[0];
        }
        return t;
    };
    auto bias_add = [&](const ggml_tensor * t) {  // [width, tokens] + [width]: its result has one reader and is not a graph output
        const ggml_tensor * a = t->src[0], * b = t->src[1];
        return t->op == GGML_OP_ADD && is_f32(t) && is_f32(a) && is_f32(b) && ggml_is_contiguous(a) && ggml_is_contiguous(t) && ggml_are_same_shape(a, t) &&
               ggml_is_contiguous(b) && b->ne[0] == a->ne[0] && b->ne[1] == 1 && b->ne[2] == 1 && b->ne[3] == 1 && a->ne[2] == 1 && a->ne[3] == 1 &&
               node_use_count(cgraph, t) == 1 && !(t->flags & GGML_TENSOR_FLAG_OUTPUT) && !tensor_is_split(b);
    };
    const ggml_tensor * rq = nullptr, * rk = nullptr, * s[2] = { nullptr, nullptr }, * adds[3] = { nullptr, nullptr, nullptr };
    const ggml_tensor * pending = nullptr;  // an ADD whose reader has not turned up yet
    int hoisted[4], n_hoisted = 0, last = -1;
    for (int j = i; j < n_nodes && j < i + 28 && last < 0; ++j) {
        const ggml_tensor * t = ggml_graph_node((ggml_cgraph *) cgraph, j);
        if (j > i && done[(size_t) j] == 2) {
            if (n_hoisted == 4) return 0;
            hoisted[n_hoisted++] = j;
            continue;
        }
        if ((j > i && done[(size_t) j]) || noop_node(t)) continue;
        if (t->op == GGML_OP_ADD && !pending && !s[0] && bias_add(t)) { pending = t; continue; }
        if (t->op == GGML_OP_ROPE && !s[0] && (!rq || !rk)) {
            if (pending && producer(t->src[0]) != pending) return 0;
            adds[rq ? 1 : 0] = pending;
            pending = nullptr;
            (rq ? rk : rq) = t;
            continue;
        }
        if (t->op == GGML_OP_SET_ROWS && rk && !s[0]) { s[0] = t; continue; }
        if (t->op == GGML_OP_SET_ROWS && s[0]) { s[1] = t; last = j; continue; }
        return 0;
    }
    if (last < 0 || !rq || !rk) return 0;
    if (!is_f32(rq) || rq->ne[3] != 1 || !supports_elementwise(rq)) return 0;
    // which of the two stores takes the rotated k?
    const ggml_tensor * sk = producer(s[0]->src[0]) == rk ? s[0] : producer(s[1]->src[0]) == rk ? s[1] : nullptr;
    if (!sk) return 0;
    const ggml_tensor * sv = sk == s[0] ? s[1] : s[0];
    const ggml_tensor * vp = producer(sv->src[0]);
    if (!vp || vp == rk || node_use_count(cgraph, rk) != 1 || (rk->flags & GGML_TENSOR_FLAG_OUTPUT)) return 0;
    if (pending) {  // the ADD in front of the V store
        if (vp != pending || !ggml_is_contiguous(sv->src[0])) return 0;
        adds[2] = pending;
    }
    if (!is_f32(rk) || !supports_elementwise(rk) || !ggml_is_contiguous(rk) || rk->ne[3] != 1 || rk->ne[0] != rq->ne[0] || rk->ne[2] != rq->ne[2] ||
        memcmp(rk->op_params, rq->op_params, sizeof(rq->op_params)) != 0 || rk->src[1] != rq->src[1] || rk->src[2] != rq->src[2]) return 0;
    if ((adds[0] && !ggml_is_contiguous(rq->src[0])) || (adds[1] && !ggml_is_contiguous(rk->src[0]))) return 0;
    auto store_ok = [&](const ggml_tensor * st, int64_t width, int64_t rows) {
        const ggml_tensor * a = st->src[0], * x = st->src[1];
        return supports_elementwise(st) && a->ne[0] == width && a->ne[1] == rows && a->ne[2] == 1 && a->ne[3] == 1 && a->nb[0] == 4 &&
               x->ne[0] == rows && x->ne[1] == 1 && x->ne[2] == 1 && st->ne[2] == 1 && st->ne[3] == 1;
    };
    if (!store_ok(sk, rk->ne[0] * rk->ne[1], rk->ne[2]) || !ggml_is_contiguous(sk->src[0]) || !store_ok(sv, sv->src[0]->ne[0], sv->src[0]->ne[1]) ||
        sk->type != sv->type || ranges_overlap(sk, sv) || ranges_overlap(rq, sk) || ranges_overlap(rq, sv)) return 0;
    // where the values really are: behind an ADD, the ADD's first source; a hoisted matmul of this group still in the scratch
    const ggml_tensor * qs = adds[0] ? adds[0]->src[0] : rq->src[0], * ks = adds[1] ? adds[1]->src[0] : rk->src[0], * vs = adds[2] ? adds[2]->src[0] : sv->src[0];
    b200q_tensor tq = bt(rq->src[0]), tk = bt(rk->src[0]), tv = bt(sv->src[0]);  // shapes and strides of the readers' views
    tq.data = qs->data; tk.data = ks->data; tv.data = vs->data;
    for (int h = 0; h < n_hoisted; ++h) {
        const ggml_tensor * m  = ggml_graph_node((ggml_cgraph *) cgraph, hoisted[h]);
        const auto &        hc = ctx->hoist_copies[(size_t) hoisted[h]];
        if (!sole_reader_takes_scratch(cgraph, hoisted[h], n_nodes)) return 0;  // somebody else waits for it: it needs its delivery copy
        if (m == producer(qs)) tq.data = (void *) hc.src;
        else if (m == producer(ks)) tk.data = (void *) hc.src;
        else if (m == producer(vs)) tv.data = (void *) hc.src;
        else return 0;
    }
    auto hits_q = [&](const b200q_tensor & t, const ggml_tensor * g) {
        const char * a0 = (const char *) t.data, * a1 = a0 + ggml_nbytes(g), * b0 = (const char *) rq->data, * b1 = b0 + ggml_nbytes(rq);
        return a0 < b1 && b0 < a1;
    };
    if (hits_q(tk, ks) || hits_q(tv, vs)) return 0;
    if (tq.data == qs->data) tq.data = bt_src(ctx, qs).data;  // q itself may be behind a redirect of an earlier position
    const b200q_tensor      td = bt(rq), tp = bt(rq->src[1]), kx = bt(sk->src[1]), kd = bt(sk), vx = bt(sv->src[1]), vd = bt(sv);
    b200q_tensor            tf{}, tb[3]{};
    if (rq->src[2]) tf = bt(rq->src[2]);
    for (int q = 0; q < 3; ++q) if (adds[q]) tb[q] = bt(adds[q]->src[1]);
    const b200q_rope_params rp = rope_params_of(rq);
    *rc = b200q_op_rope_store(&tq, &td, &tk, &tp, rq->src[2] ? &tf : nullptr, &rp, &kx, &kd, &tv, &vx, &vd, adds[0] ? &tb[0] : nullptr,
                              adds[1] ? &tb[1] : nullptr, adds[2] ? &tb[2] : nullptr, ctx->stream);
    return last + 1;
}

ggml_status graph_dispatch(b200_backend_ctx * ctx, ggml_cgraph * cgraph) {
    const int  n_nodes  = ggml_graph_n_nodes(cgraph);
    const bool no_fuse  = getenv("GGML_B200_NO_FUSE") != nullptr;
    std::vector<char> done((size_t) n_nodes, 0);  // nodes that already went out with an earlier list (2: result waits in scratch)
    ctx->packed_act = nullptr;
    ctx->hoist_used = 0;
    ctx->redirect.clear();
    ctx->hoist_copies.assign((size_t) n_nodes, b200_backend_ctx::hoist_copy{});
    // scratch results live in the head of the workspace, which no operator is given (backend_workspace)
    ctx->hoist_ok = getenv("GGML_B200_NO_LIST") == nullptr;
    static const bool stats_hook = getenv("GGML_B200_STATS") != nullptr && atexit(print_fusion_stats) == 0;
    (void) stats_hook;
    if (ctx->hoist_ok) backend_workspace(ctx, 0);
    const bool no_lists = getenv("GGML_B200_NO_LIST") != nullptr;
    for (int i = 0; i < n_nodes; ++i) {
        ggml_tensor * node = ggml_graph_node(cgraph, i);
        if (done[(size_t) i] == 2) {
            const auto & hc = ctx->hoist_copies[(size_t) i];
            if (ctx->packed_act && ranges_overlap(node, ctx->packed_act)) ctx->packed_act = nullptr;
            if (!no_fuse && sole_reader_takes_scratch(cgraph, i, n_nodes)) {
                ctx->redirect.emplace_back(node->data, hc.src);
                ++g_fusion_stats[FUSE_SCRATCH_READ];
                continue;
            }
            B200_CHECK(b200q_memcpy(node->data, hc.src, hc.bytes, B200Q_D2D, ctx->stream));
            ++g_fusion_stats[FUSE_SCRATCH_COPY];
            continue;
        }
        if (ggml_is_empty(node) || done[(size_t) i]) continue;
        // a node that writes over the packed activations (in-place operators) or takes the workspace for something else
        if (ctx->packed_act && (node->op == GGML_OP_MUL_MAT_ID || (node->op != GGML_OP_MUL_MAT && ranges_overlap(node, ctx->packed_act)))) {
            ctx->packed_act = nullptr;
        }
        if (!no_lists && list_candidate(node)) {
            if (compute_list_run(ctx, cgraph, i, n_nodes, done) > 0) {
                ++g_fusion_stats[FUSE_LIST];
                continue;
            }
        }
        int rc = 0;
        if (!no_fuse && (node->op == GGML_OP_ADD || node->op == GGML_OP_RMS_NORM)) {
            const int used = try_fuse_norm(ctx, cgraph, i, n_nodes, &rc);
            if (used > 0) {
                if (rc != 0) {
                    fprintf(stderr, "ggml-b200: fused norm (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                ++g_fusion_stats[FUSE_NORM];
                i += used - 1;
                continue;
            }
        }
        if (!no_fuse && node->op == GGML_OP_MUL_MAT_ID && i + 1 < n_nodes && node->src[2]->ne[1] == 1 && cgraph->use_counts) {
            // decode: the experts' bias rows (ADD_ID behind the matmul, gpt-oss) go into the matmul's epilogue
            const ggml_tensor * nx = ggml_graph_node(cgraph, i + 1);
            if (nx->op == GGML_OP_ADD_ID && nx->src[0] == node && nx->src[2] == node->src[2] && supports_elementwise(nx) &&
                ggml_is_contiguous(nx) && ggml_is_contiguous(node) && nx->src[1]->nb[0] == 4 && nx->src[1]->ne[0] == node->ne[0] &&
                node_use_count(cgraph, node) == 1 && !(node->flags & GGML_TENSOR_FLAG_OUTPUT) && !node->view_src) {
                rc = compute_mul_mat_id(ctx, node, nx);
                if (rc == 0) {
                    ++g_fusion_stats[FUSE_ID_BIAS];
                    ++i;
                    continue;
                }
                if (rc != 1) {
                    fprintf(stderr, "ggml-b200: MUL_MAT_ID + ADD_ID (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                rc = 0;  // not this shape: the two nodes go out one by one
            }
        }
        if (!no_fuse && node->op == GGML_OP_MUL && node->src[1] && node->src[1]->ne[0] == 1 && node->ne[1] > 1) {
            const int behind = try_fuse_moe_sum(ctx, cgraph, i, n_nodes, &rc);
            if (behind > i) {
                if (rc != 0) {
                    fprintf(stderr, "ggml-b200: fused MoE combine (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                ++g_fusion_stats[FUSE_MOE_SUM];
                i = behind - 1;
                continue;
            }
        }
        if (!no_fuse && (node->op == GGML_OP_ADD || node->op == GGML_OP_GET_ROWS)) {
            const int behind = node->op == GGML_OP_ADD ? try_fuse_moe_route(ctx, cgraph, i, n_nodes, &rc) : try_fuse_gate_weights(ctx, cgraph, i, n_nodes, &rc);
            if (behind > i) {
                if (rc != 0) {
                    fprintf(stderr, "ggml-b200: fused MoE router (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                ++g_fusion_stats[FUSE_ROUTE];
                i = behind - 1;
                continue;
            }
        }
        if (!no_fuse && (node->op == GGML_OP_ROPE || node->op == GGML_OP_ADD)) {
            const int behind = try_fuse_rope_store(ctx, cgraph, i, n_nodes, done, &rc);
            if (behind > i) {
                if (rc != 0) {
                    fprintf(stderr, "ggml-b200: fused rope + cache store (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                ++g_fusion_stats[FUSE_ROPE_STORE];
                i = behind - 1;
                continue;
            }
        }
        if (!no_fuse && node->op == GGML_OP_SET_ROWS) {  // the K and V cache writes of a layer: one launch
            int j = i + 1;
            while (j < n_nodes && noop_node(ggml_graph_node(cgraph, j))) ++j;
            const ggml_tensor * o = j < n_nodes ? ggml_graph_node(cgraph, j) : nullptr;
            if (o && o->op == GGML_OP_SET_ROWS && !done[(size_t) j] && supports_elementwise(node) && supports_elementwise(o) && o->type == node->type &&
                !touches(o, node) && !ranges_overlap(node, o)) {
                const b200q_tensor s0 = bt_src(ctx, node->src[0]), x0 = bt(node->src[1]), d0 = bt(node), s1 = bt_src(ctx, o->src[0]), x1 = bt(o->src[1]), d1 = bt(o);
                rc = b200q_op_set_rows_pair(&s0, &x0, &d0, &s1, &x1, &d1, ctx->stream);
                if (rc != 0) {
                    fprintf(stderr, "ggml-b200: set_rows pair (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                ++g_fusion_stats[FUSE_SET_ROWS_PAIR];
                i = j;
                continue;
            }
        }
        if (!no_fuse && node->op == GGML_OP_MUL_MAT && node->src[0] && node->src[0]->type == GGML_TYPE_F16) {
            const int behind = try_fuse_attn(ctx, cgraph, i, n_nodes, &rc);
            if (behind > i) {
                if (rc != 0) {
                    fprintf(stderr, "ggml-b200: fused attention (%s) failed (%d): %s\n", node->name, rc, b200q_last_error());
                    return GGML_STATUS_FAILED;
                }
                ++g_fusion_stats[FUSE_ATTN];
                ctx->packed_act = nullptr;
                i = behind - 1;
                continue;
            }
        }
        switch (node->op) {
            case GGML_OP_NONE:
            case GGML_OP_VIEW:
            case GGML_OP_RESHAPE:
            case GGML_OP_PERMUTE:
            case GGML_OP_TRANSPOSE:
                continue;
            case GGML_OP_MUL_MAT:    rc = compute_mul_mat(ctx, node); break;
            case GGML_OP_MUL_MAT_ID: rc = compute_mul_mat_id(ctx, node); break;
            case GGML_OP_GET_ROWS:   rc = compute_get_rows(ctx, node); break;
            case GGML_OP_RMS_NORM: case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV: case GGML_OP_SCALE:
            case GGML_OP_UNARY: case GGML_OP_GLU: case GGML_OP_SOFT_MAX: case GGML_OP_ROPE: case GGML_OP_CPY: case GGML_OP_CONT:
            case GGML_OP_DUP: case GGML_OP_SET_ROWS: case GGML_OP_ADD_ID: case GGML_OP_ARGSORT:
                rc = compute_elementwise(ctx, node);
                break;
            default:
                fprintf(stderr, "ggml-b200: op %s reached graph_compute but is not supported\n", ggml_op_name(node->op));
                return GGML_STATUS_FAILED;
        }
        if (rc != 0) {
            fprintf(stderr, "ggml-b200: %s (%s) failed (%d): %s\n", ggml_op_name(node->op), node->name, rc, b200q_last_error());
            return rc == B200Q_ENOMEM ? GGML_STATUS_ALLOC_FAILED : GGML_STATUS_FAILED;
        }
    }
    return GGML_STATUS_SUCCESS;
}

const ggml_backend_i backend_iface = {
    /* .get_name           = */ backend_get_name,
    /* .free               = */ backend_free,
    /* .set_tensor_async   = */ backend_set_tensor_async,
    /* .get_tensor_async   = */ backend_get_tensor_async,
    /* .cpy_tensor_async   = */ backend_cpy_tensor_async,
    /* .synchronize        = */ backend_synchronize,
    /* .graph_plan_create  = */ nullptr,
    /* .graph_plan_free    = */ nullptr,
    /* .graph_plan_update  = */ nullptr,
    /* .graph_plan_compute = */ nullptr,
    /* .graph_compute      = */ backend_graph_compute,
    /* .event_record       = */ backend_event_record,
    /* .event_wait         = */ backend_event_wait,
};
bool backend_is_b200(ggml_backend_t backend) { return backend && backend->iface.get_name == backend_get_name; }

ggml_guid_t backend_guid() {
    static ggml_guid guid = { 0xb2, 0x00, 0x71, 0x5e, 0x10, 0x0a, 0x4c, 0x0d, 0x9e, 0x51, 0x67, 0x67, 0x6d, 0x6c, 0x62, 0x32 };
    return &guid;
}

// ------------------------------------------------------------------------------------------------ device
struct b200_device {
    ggml_backend_device      dev;
    ggml_backend_buffer_type buft;
    b200_device_ctx          ctx;
};

const char * device_get_name(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->name.c_str(); }
const char * device_get_description(ggml_backend_dev_t dev) { return ((b200_device_ctx *) dev->context)->description.c_str(); }
void device_get_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    auto * d = (b200_device_ctx *) dev->context;
    *free = *total = 0;
    b200q_device_info(d->device, nullptr, 0, total, free, nullptr, nullptr, nullptr, nullptr);
}
enum ggml_backend_dev_type device_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }
void device_get_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    props->name        = device_get_name(dev);
    props->description = device_get_description(dev);
    props->type        = device_get_type(dev);
    device_get_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = { /* async */ true, /* host_buffer */ true, /* buffer_from_host_ptr */ false, /* events */ true };
}
ggml_backend_t device_init_backend(ggml_backend_dev_t dev, const char *) {
    auto * d   = (b200_device_ctx *) dev->context;
    auto * ctx = new b200_backend_ctx{ d->device, d->name };
    if (b200q_set_device(d->device) != 0 || b200q_stream_create(&ctx->stream) != 0) {
        fprintf(stderr, "ggml-b200: cannot initialise device %d: %s\n", d->device, b200q_last_error());
        delete ctx;
        return nullptr;
    }
    return new ggml_backend{ backend_guid(), backend_iface, dev, ctx };
}
ggml_backend_buffer_type_t device_get_buffer_type(ggml_backend_dev_t dev) {
    return &((b200_device *) ((char *) dev - offsetof(b200_device, dev)))->buft;
}
bool device_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    // its own device memory, or a split buffer whose main device it is (ggml-cuda.cu:3550-3556)
    return (buft->iface.get_name == buft_get_name || buft_is_split(buft)) && buft->device == dev;
}
ggml_backend_buffer_type_t device_get_host_buffer_type(ggml_backend_dev_t) { return host_buffer_type(); }
ggml_backend_event_t device_event_new(ggml_backend_dev_t dev) {
    auto * d = (b200_device_ctx *) dev->context;
    void * e = nullptr;
    if (b200q_set_device(d->device) != 0 || b200q_event_create(&e) != 0) return nullptr;
    return new ggml_backend_event{ dev, e };
}
void device_event_free(ggml_backend_dev_t, ggml_backend_event_t event) {
    b200q_event_destroy(event->context);
    delete event;
}
void device_event_synchronize(ggml_backend_dev_t, ggml_backend_event_t event) { B200_CHECK(b200q_event_synchronize(event->context)); }
bool device_offload_op(ggml_backend_dev_t, const ggml_tensor * op) {
    // weights living in host memory are worth shipping to the device for batches, as in the reference
    // (ggml-cuda.cu offload_op: batch >= 32)
    const int64_t batch = op->op == GGML_OP_MUL_MAT_ID ? op->ne[2] : op->ne[1];
    return (op->op == GGML_OP_MUL_MAT || op->op == GGML_OP_MUL_MAT_ID) && batch >= 32;
}

const ggml_backend_device_i device_iface = {
    /* .get_name             = */ device_get_name,
    /* .get_description      = */ device_get_description,
    /* .get_memory           = */ device_get_memory,
    /* .get_type             = */ device_get_type,
    /* .get_props            = */ device_get_props,
    /* .init_backend         = */ device_init_backend,
    /* .get_buffer_type      = */ device_get_buffer_type,
    /* .get_host_buffer_type = */ device_get_host_buffer_type,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ device_supports_op,
    /* .supports_buft        = */ device_supports_buft,
    /* .offload_op           = */ device_offload_op,
    /* .event_new            = */ device_event_new,
    /* .event_free           = */ device_event_free,
    /* .event_synchronize    = */ device_event_synchronize,
};

// ------------------------------------------------------------------------------------------------ registry
struct b200_reg_ctx {
    std::vector<b200_device *> devices;
};

const char * reg_get_name(ggml_backend_reg_t) { return "B200"; }
size_t       reg_get_device_count(ggml_backend_reg_t reg) { return ((b200_reg_ctx *) reg->context)->devices.size(); }
ggml_backend_dev_t reg_get_device(ggml_backend_reg_t reg, size_t index) {
    auto * ctx = (b200_reg_ctx *) reg->context;
    return index < ctx->devices.size() ? &ctx->devices[index]->dev : nullptr;
}

ggml_backend_feature * get_features(ggml_backend_reg_t) {
    static ggml_backend_feature features[] = {
        { "ARCHS", "1000a" }, { "TMA_BULK", "1" }, { "PDL", "1" }, { "DP4A", "1" }, { nullptr, nullptr },
    };
    return features;
}
void * reg_get_proc_address(ggml_backend_reg_t, const char * name) {
    if (strcmp(name, "ggml_backend_get_features") == 0) return (void *) get_features;
    if (strcmp(name, "ggml_backend_split_buffer_type") == 0) return (void *) split_buffer_type;  // llama's -sm row
    if (strcmp(name, "ggml_backend_b200_graph_stats") == 0) return (void *) graph_stats;
    if (strcmp(name, "ggml_backend_b200_fusion_stats") == 0) return (void *) fusion_stats;
    return nullptr;
}

const ggml_backend_reg_i reg_iface = {
    /* .get_name         = */ reg_get_name,
    /* .get_device_count = */ reg_get_device_count,
    /* .get_device       = */ reg_get_device,
    /* .get_proc_address = */ reg_get_proc_address,
};

b200_reg_ctx * g_reg_ctx = nullptr;
int               registry_device_count() { return g_reg_ctx ? (int) g_reg_ctx->devices.size() : 0; }
b200_device_ctx * registry_device_ctx(int index) { return &g_reg_ctx->devices[index]->ctx; }
ggml_backend_dev_t registry_device(int index) { return &g_reg_ctx->devices[index]->dev; }

int count_blackwell_devices(std::vector<int> * ordinals) {
    int n = b200q_device_count(), found = 0;
    for (int i = 0; i < n; ++i) {
        int major = 0, minor = 0;
        if (b200q_device_info(i, nullptr, 0, nullptr, nullptr, &major, &minor, nullptr, nullptr) == 0 && major == 10) {
            if (ordinals) ordinals->push_back(i);
            ++found;
        }
    }
    return found;
}

}  // namespace

extern "C" {

// named entry point for static registration (the reference's ggml_backend_cuda_reg, ggml-cuda.cu:3751-3790)
ggml_backend_reg_t ggml_backend_b200_reg(void) {
    static std::mutex       mu;
    static ggml_backend_reg reg;
    static bool             initialised = false;
    std::lock_guard<std::mutex> lk(mu);
    if (!initialised) {
        auto *           ctx = new b200_reg_ctx;
        std::vector<int> ordinals;
        count_blackwell_devices(&ordinals);
        for (size_t i = 0; i < ordinals.size(); ++i) {
            auto * d            = new b200_device;
            d->ctx.device       = ordinals[i];
            d->ctx.index        = (int) i;
            d->ctx.name         = "B200" + std::to_string(i);
            char name[128]      = "B200";
            b200q_device_info(ordinals[i], name, sizeof(name), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
            d->ctx.description  = name;
            d->dev              = ggml_backend_device{ device_iface, &reg, &d->ctx };
            d->buft             = ggml_backend_buffer_type{ buft_iface, &d->dev, nullptr };
            ctx->devices.push_back(d);
        }
        g_reg_ctx   = ctx;
        reg         = ggml_backend_reg{ GGML_BACKEND_API_VERSION, reg_iface, ctx };
        initialised = true;
    }
    return &reg;
}

// dlopen entry points (ggml-backend-impl.h:215-251; looked up by name in ggml-backend-reg.cpp:232-276)
ggml_backend_reg_t ggml_backend_init(void) { return ggml_backend_b200_reg(); }
int                ggml_backend_score(void) { return count_blackwell_devices(nullptr) > 0 ? 100 : 0; }

}  // extern "C"
