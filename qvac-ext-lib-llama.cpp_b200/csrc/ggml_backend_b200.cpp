// ggml_backend_b200.cpp — the drop-in: a ggml backend ("B200") that implements the reference's five backend vtables
// (/root/reference/ggml/src/ggml-backend-impl.h:16-204, API version 2) for GGML_OP_MUL_MAT and GGML_OP_MUL_MAT_ID on
// q4_0 / q8_0 / q4_K / q5_K / q6_K weights x F32 activations, the element-wise nodes the fused mat-vec epilogue absorbs (GLU, ADD,
// MUL, ADD_ID), and - SURVEY 8f rank 3, first slice - the ops between the matmuls of a Llama decode graph (GET_ROWS, RMS_NORM, ROPE,
// SET_ROWS, CPY / CONT / DUP, FLASH_ATTN_EXT) with the planner that fuses them (run_nodes). All device work goes through the C-ABI of
// include/ggml_b200.h (libb200mm.so); this file contains no CUDA and no CPU compute.
//
// Loaded out of tree: GGML_BACKEND_PATH=/path/libggml-b200.so (ggml-backend-reg.cpp:711-715) or ggml_backend_load();
// the exported symbols are ggml_backend_init / ggml_backend_score (ggml-backend-impl.h:215-251).
//
// What it mirrors in ggml-cuda (paths relative to /root/reference/ggml/src/ggml-cuda/ggml-cuda.cu):
//   buffer + buffer type :564-769, split buffer type :773-1102 + split MUL_MAT :1447-1765, pinned host buffer type :1104-1168,
//   backend (stream) iface :2748-2862, 3679-3754,
//   device iface :4160-4677, reg + get_proc_address :4683-4830.

#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#include "ggml.h"

#include "ggml_b200.h"

#include <cstdint>
#include <cstdio>
#include <algorithm>
#include <array>
#include <atomic>
#include <cstring>
#include <chrono>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#define B200_NAME "B200"
#define B200_MAX_DEVICES 16

namespace {

struct dev_ctx {
    int         device;
    std::string name, description, pci_id;
    ggml_backend_buffer_type buft;       // device buffer type singleton
    ggml_backend_buffer_type host_buft;  // pinned host buffer type singleton
};

struct buf_ctx {
    int    device;
    void * base;
};

// a split MUL_MAT runs one slice on every device: the backend keeps a context (stream), an event and scratch for the replicated
// src1 and the dst slice on each of the other devices (created on first use)
struct split_peer {
    b200_ctx * ctx     = nullptr;
    void *     event   = nullptr;
    void *     x       = nullptr;
    size_t     x_size  = 0;
    void *     d       = nullptr;
    size_t     d_size  = 0;
    bool       direct  = false;   // peer access works both ways: the slice is stored straight into the main device's dst
};
struct be_ctx {
    int        device;
    b200_ctx * ctx;
    std::string name;
    bool       fuse;                                // shared-src1 peephole (q/k/v, gate/up in one launch)
    bool       fuse_epilogue = true;               // K7 peephole: GLU / bias epilogues of the one-column mat-vec
    bool       weight_cache  = true;               // prefill operand expansions are kept per weight tensor (GGML_B200_WEIGHT_CACHE=0: rebuilt per call)
    void *     split_event = nullptr;              // "src1 is ready" on the main stream
    void *     copy_event  = nullptr;              // cpy_tensor_async: "the copy issued on this backend's stream is done"
    split_peer peers[B200_MAX_DEVICES];
    // CUDA-graph replay of graph_compute (ggml-cuda.cu:3679-3754): a cgraph whose node properties match one seen before is
    // captured on its second appearance and replayed from then on
    bool       use_graphs      = true;             // GGML_B200_DISABLE_GRAPHS=1 switches it off
    int        graph_min_nodes = 4;                // GGML_B200_GRAPH_MIN_NODES: shorter cgraphs are launched directly
    bool       capturing       = false;
    struct cached_graph {
        std::vector<uint64_t> key;                 // the properties of every node, flattened
        b200_graph *          exec  = nullptr;
        int                   seen  = 0;
        bool                  bad   = false;       // capture failed once: never again for this key
        uint64_t              stamp = 0;           // last use (LRU)
    };
    std::map<uint64_t, std::vector<cached_graph>> graphs;  // hash -> entries
    size_t     n_graphs    = 0;
    uint64_t   last_hash   = 0;                    // bucket of the cgraph computed last
    bool       last_hash_valid = false;
    uint64_t   graph_clock = 0;
    uint64_t   graph_replays = 0, graph_captures = 0;
    // try_fuse_attn_block -> the FLASH_ATTN_EXT right behind it: the only KV-cache rows in flight are the token's slot (b200_flash_attn_ext_decode)
    int        kv_hint_node = -1;
    const void * kv_hint_k = nullptr, * kv_hint_v = nullptr;
    const ggml_tensor * kv_hint_idx = nullptr;
    void *     blk_scratch = nullptr;              // try_fuse_attn_block: q / k / v projections of the layer in flight
    size_t     blk_scratch_size = 0;
    std::vector<char> done;                        // run_nodes: nodes already executed ahead of their position (hoisted into a fused launch)
    bool       fuse_f3 = true;                     // GGML_B200_DISABLE_F3_FUSION=1: RMS_NORM+MUL, ROPE/SET_ROWS and the q/k/v hoist run node by node
    // GGML_B200_TIMING=1: where a graph_compute goes (printed when the backend is freed)
    bool       timing = false;
    struct tstat { uint64_t n = 0, key_ns = 0, run_ns = 0, gpu_n = 0; double gpu_ms = 0; } t_replay, t_direct;
    void *     t_ev0 = nullptr, * t_ev1 = nullptr;
    bool       t_pending = false, t_pending_replay = false;
};

ggml_backend_reg_t b200_reg();

[[noreturn]] void die(const char * what) {
    GGML_LOG_ERROR("ggml-b200: %s failed: %s\n", what, b200_last_error());
    GGML_ABORT("ggml-b200 fatal error");  // same convention as CUDA_CHECK -> ggml_cuda_error (ggml-cuda.cu:82-93)
}
#define CHECK(call) do { if ((call) != B200_OK) die(#call); } while (0)
// inside graph_compute: a failure while the stream is being captured (a scratch buffer that would have to grow, a weight
// expansion that is not cached yet) is not fatal - the capture is abandoned and the nodes are launched directly
struct capture_failed {};
#define CHECK_RUN(c, call) do { if ((call) != B200_OK) { if ((c)->capturing) throw capture_failed{}; die(#call); } } while (0)

// ------------------------------------------------------------------------------------------------ device buffer

// every host-side write into a buffer (and its release) is reported to the prefill weight cache, which keeps int8
// expansions keyed by weight pointer: the whole tensor range is dropped, so chunked uploads (src/llama-model-loader.cpp
// :1072-1090) and per-expert matrices inside a 3-D tensor are covered
void drop_cached(ggml_backend_buffer_t buffer, const ggml_tensor * t) {
    if (t && t->data && ggml_is_quantized(t->type)) b200_weight_cache_invalidate(((buf_ctx *) buffer->context)->device, t->data, ggml_nbytes(t));
}
void buf_free(ggml_backend_buffer_t buffer) {
    buf_ctx * c = (buf_ctx *) buffer->context;
    b200_weight_cache_invalidate(c->device, c->base, buffer->size);
    b200_free(c->device, c->base);
    delete c;
}
void * buf_get_base(ggml_backend_buffer_t buffer) { return ((buf_ctx *) buffer->context)->base; }
ggml_status buf_init_tensor(ggml_backend_buffer_t, ggml_tensor *) {
    // nothing to do: the kernels never read past a row's own bytes, so the reference's 512-element row padding and its
    // zero fill (ggml-cuda.cu:593-612, 720-734) are not needed
    return GGML_STATUS_SUCCESS;
}
void buf_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor * t, uint8_t value, size_t offset, size_t size) {
    drop_cached(buffer, t);
    CHECK(b200_memset(((buf_ctx *) buffer->context)->device, (char *) t->data + offset, value, size));
}
void buf_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor * t, const void * data, size_t offset, size_t size) {
    drop_cached(buffer, t);
    CHECK(b200_memcpy_h2d(((buf_ctx *) buffer->context)->device, (char *) t->data + offset, data, size));
}
void buf_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * t, void * data, size_t offset, size_t size) {
    CHECK(b200_memcpy_d2h(((buf_ctx *) buffer->context)->device, data, (const char *) t->data + offset, size));
}
bool buft_is_b200(ggml_backend_buffer_type_t buft);
bool buf_cpy_tensor(ggml_backend_buffer_t buffer, const ggml_tensor * src, ggml_tensor * dst) {
    ggml_backend_buffer_t sb = src->view_src ? src->view_src->buffer : src->buffer;
    if (!sb || !buft_is_b200(sb->buft)) return false;
    buf_ctx * s = (buf_ctx *) sb->context, * d = (buf_ctx *) buffer->context;
    drop_cached(buffer, dst);
    if (s->device == d->device) CHECK(b200_memcpy_d2d(d->device, dst->data, src->data, ggml_nbytes(src)));
    else CHECK(b200_memcpy_peer(d->device, dst->data, s->device, src->data, ggml_nbytes(src)));  // synchronous on return (ggml-cuda.cu:648-651)
    return true;
}
void buf_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    buf_ctx * c = (buf_ctx *) buffer->context;
    b200_weight_cache_invalidate(c->device, c->base, buffer->size);
    CHECK(b200_memset(c->device, c->base, value, buffer->size));
}

const ggml_backend_buffer_i buf_iface = {
    /* .free_buffer   = */ buf_free,
    /* .get_base      = */ buf_get_base,
    /* .init_tensor   = */ buf_init_tensor,
    /* .memset_tensor = */ buf_memset_tensor,
    /* .set_tensor    = */ buf_set_tensor,
    /* .get_tensor    = */ buf_get_tensor,
    /* .cpy_tensor    = */ buf_cpy_tensor,
    /* .clear         = */ buf_clear,
    /* .reset         = */ nullptr,
};

// ------------------------------------------------------------------------------------------------ device buffer type

const char * buft_name(ggml_backend_buffer_type_t buft) { return ((dev_ctx *) buft->context)->name.c_str(); }
ggml_backend_buffer_t buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    dev_ctx * d = (dev_ctx *) buft->context;
    void * p = nullptr;
    if (b200_malloc(d->device, size, &p) != B200_OK) {  // soft failure, as ggml-cuda.cu:700-707
        GGML_LOG_ERROR("ggml-b200: allocating %.2f MiB on device %d failed: %s\n", size / 1048576.0, d->device, b200_last_error());
        return nullptr;
    }
    return ggml_backend_buffer_init(buft, buf_iface, new buf_ctx{ d->device, p }, size);
}
size_t buft_alignment(ggml_backend_buffer_type_t) { return 256; }  // every tensor starts 16-byte aligned for the bulk copies
size_t buft_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * t) { return ggml_nbytes(t); }

const ggml_backend_buffer_type_i buft_iface = {
    /* .get_name       = */ buft_name,
    /* .alloc_buffer   = */ buft_alloc,
    /* .get_alignment  = */ buft_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ buft_alloc_size,
    /* .is_host        = */ nullptr,
};
bool buft_is_b200(ggml_backend_buffer_type_t buft) { return buft->iface.get_name == buft_name; }

// ------------------------------------------------------------------------------------------------ pinned host buffer type

const char * host_buft_name(ggml_backend_buffer_type_t) { return B200_NAME "_Host"; }
void host_buf_free(ggml_backend_buffer_t buffer) { b200_host_free(buffer->context); }
ggml_backend_buffer_t host_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    void * p = nullptr;
    if (b200_host_malloc(size, &p) != B200_OK) {
        // fall back to pageable memory like ggml-cuda.cu:1140-1146
        return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);
    }
    ggml_backend_buffer_t b = ggml_backend_cpu_buffer_from_ptr(p, size);
    b->buft = buft;
    b->iface.free_buffer = host_buf_free;
    return b;
}
size_t host_buft_alignment(ggml_backend_buffer_type_t) { return 64; }
bool   host_buft_is_host(ggml_backend_buffer_type_t) { return true; }
const ggml_backend_buffer_type_i host_buft_iface = {
    /* .get_name       = */ host_buft_name,
    /* .alloc_buffer   = */ host_buft_alloc,
    /* .get_alignment  = */ host_buft_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ nullptr,
    /* .is_host        = */ host_buft_is_host,
};

// ------------------------------------------------------------------------------------------------ split buffer type
//
// ggml_backend_split_buffer_type(main_device, tensor_split) (ggml-backend.h:199-200; the reference's version: ggml-cuda.cu:773-1102):
// a 2-D weight tensor placed in such a buffer is stored as one row slice per device of the process - rows [low, high) by the
// get_row_split rule (:786-799: cumulative fractions, boundaries rounded down, the last device takes the rest). There is no single
// base address: tensor->extra carries the per-device pointers (the reference keeps them there too, common.cuh:943-946).

const int64_t SPLIT_ROUNDING = 128;  // rows: the tile height of the tensor-core kernels (the reference rounds to its mmq_y, :773-784)

struct split_buft_ctx {
    int         main_device;
    std::array<float, B200_MAX_DEVICES> tensor_split;  // cumulative start fraction of every device
    std::string name;
};
struct split_extra {
    void *  data[B200_MAX_DEVICES]     = {};
    int64_t row_low[B200_MAX_DEVICES]  = {};
    int64_t row_high[B200_MAX_DEVICES] = {};
    size_t  bytes[B200_MAX_DEVICES]    = {};
};
struct split_buf_ctx {
    std::vector<split_extra *> extras;
};

int split_dev_count() { return std::min(b200_device_count(), B200_MAX_DEVICES); }

void split_rows(const split_buft_ctx * c, int64_t nrows, int id, int64_t * low, int64_t * high) {
    const int n = split_dev_count();
    *low = id == 0 ? 0 : (int64_t) (nrows * c->tensor_split[id]);
    *low -= *low % SPLIT_ROUNDING;
    if (id == n - 1) {
        *high = nrows;
    } else {
        *high = (int64_t) (nrows * c->tensor_split[id + 1]);
        *high -= *high % SPLIT_ROUNDING;
    }
}

const char * split_buft_name(ggml_backend_buffer_type_t buft) { return ((split_buft_ctx *) buft->context)->name.c_str(); }
bool buft_is_split(ggml_backend_buffer_type_t buft) { return buft->iface.get_name == split_buft_name; }

void split_buf_free(ggml_backend_buffer_t buffer) {
    split_buf_ctx * c = (split_buf_ctx *) buffer->context;
    for (split_extra * e : c->extras) {
        for (int id = 0; id < B200_MAX_DEVICES; id++) {
            if (!e->data[id]) continue;
            b200_weight_cache_invalidate(id, e->data[id], e->bytes[id]);
            b200_free(id, e->data[id]);
        }
        delete e;
    }
    delete c;
}
void * split_buf_get_base(ggml_backend_buffer_t) { return (void *) 0x1000; }  // never dereferenced (the reference returns the same dummy, :833-837)
ggml_status split_buf_init_tensor(ggml_backend_buffer_t buffer, ggml_tensor * t) {
    GGML_ASSERT(t->view_src == nullptr);  // views of split tensors are not supported (:846-848)
    GGML_ASSERT(ggml_is_contiguous(t) && "split buffers only support contiguous tensors");
    split_buf_ctx *  c  = (split_buf_ctx *) buffer->context;
    split_buft_ctx * bc = (split_buft_ctx *) buffer->buft->context;
    split_extra * e = new split_extra();
    c->extras.push_back(e);
    const int64_t nrows = ggml_nrows(t);
    const size_t  nb1   = t->nb[1];
    for (int id = 0; id < split_dev_count(); id++) {
        split_rows(bc, nrows, id, &e->row_low[id], &e->row_high[id]);
        const int64_t n = e->row_high[id] - e->row_low[id];
        if (n <= 0) continue;
        if (b200_malloc(id, (size_t) n * nb1, &e->data[id]) != B200_OK) {
            GGML_LOG_ERROR("ggml-b200: allocating a %.2f MiB slice of %s on device %d failed: %s\n", n * nb1 / 1048576.0, t->name, id, b200_last_error());
            return GGML_STATUS_ALLOC_FAILED;
        }
        e->bytes[id] = (size_t) n * nb1;
    }
    t->extra = e;
    return GGML_STATUS_SUCCESS;
}
void split_buf_set_tensor(ggml_backend_buffer_t, ggml_tensor * t, const void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(t));  // split tensors are set in one piece (:908-909)
    split_extra * e = (split_extra *) t->extra;
    const size_t nb1 = t->nb[1];
    for (int id = 0; id < split_dev_count(); id++) {
        const int64_t n = e->row_high[id] - e->row_low[id];
        if (n <= 0) continue;
        b200_weight_cache_invalidate(id, e->data[id], (size_t) n * nb1);
        CHECK(b200_memcpy_h2d(id, e->data[id], (const char *) data + e->row_low[id] * nb1, (size_t) n * nb1));
    }
}
void split_buf_get_tensor(ggml_backend_buffer_t, const ggml_tensor * t, void * data, size_t offset, size_t size) {
    GGML_ASSERT(offset == 0 && size == ggml_nbytes(t));
    const split_extra * e = (const split_extra *) t->extra;
    const size_t nb1 = t->nb[1];
    for (int id = 0; id < split_dev_count(); id++) {
        const int64_t n = e->row_high[id] - e->row_low[id];
        if (n > 0) CHECK(b200_memcpy_d2h(id, (char *) data + e->row_low[id] * nb1, e->data[id], (size_t) n * nb1));
    }
}
void split_buf_clear(ggml_backend_buffer_t, uint8_t) {}  // as the reference (:960-963)

const ggml_backend_buffer_i split_buf_iface = {
    /* .free_buffer   = */ split_buf_free,
    /* .get_base      = */ split_buf_get_base,
    /* .init_tensor   = */ split_buf_init_tensor,
    /* .memset_tensor = */ nullptr,
    /* .set_tensor    = */ split_buf_set_tensor,
    /* .get_tensor    = */ split_buf_get_tensor,
    /* .cpy_tensor    = */ nullptr,
    /* .clear         = */ split_buf_clear,
    /* .reset         = */ nullptr,
};

ggml_backend_buffer_t split_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    // the slices are allocated tensor by tensor in init_tensor; `size` only has to be what get_alloc_size promised
    return ggml_backend_buffer_init(buft, split_buf_iface, new split_buf_ctx(), size);
}
size_t split_buft_alignment(ggml_backend_buffer_type_t) { return 256; }
size_t split_buft_alloc_size(ggml_backend_buffer_type_t buft, const ggml_tensor * t) {
    const split_buft_ctx * c = (const split_buft_ctx *) buft->context;
    GGML_ASSERT(ggml_is_contiguous(t) && "split buffers only support contiguous tensors");
    size_t total = 0;
    for (int id = 0; id < split_dev_count(); id++) {
        int64_t lo, hi;
        split_rows(c, ggml_nrows(t), id, &lo, &hi);
        if (hi > lo) total += (size_t) (hi - lo) * t->nb[1];
    }
    return total;
}
bool split_buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i split_buft_iface = {
    /* .get_name       = */ split_buft_name,
    /* .alloc_buffer   = */ split_buft_alloc,
    /* .get_alignment  = */ split_buft_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ split_buft_alloc_size,
    /* .is_host        = */ split_buft_is_host,
};

// ------------------------------------------------------------------------------------------------ op geometry

bool is_noop(const ggml_tensor * t) {
    return ggml_is_empty(t) || t->op == GGML_OP_NONE || t->op == GGML_OP_RESHAPE || t->op == GGML_OP_VIEW || t->op == GGML_OP_PERMUTE ||
           t->op == GGML_OP_TRANSPOSE;
}
bool weight_type_ok(ggml_type t) { return t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0 || t == GGML_TYPE_Q4_K || t == GGML_TYPE_Q5_K || t == GGML_TYPE_Q6_K; }

bool fill_mul_mat(const ggml_tensor * op, b200_mul_mat_desc * d) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    if (!a || !b || !weight_type_ok(a->type) || b->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
    if (a->nb[0] != ggml_type_size(a->type) || b->nb[0] != sizeof(float) || op->nb[0] != sizeof(float)) return false;
    d->type = (int32_t) a->type;
    d->ne00 = a->ne[0]; d->ne01 = a->ne[1]; d->ne02 = a->ne[2]; d->ne03 = a->ne[3];
    d->nb01 = a->nb[1]; d->nb02 = a->nb[2]; d->nb03 = a->nb[3];
    d->ne11 = b->ne[1]; d->ne12 = b->ne[2]; d->ne13 = b->ne[3];
    d->nb11 = b->nb[1]; d->nb12 = b->nb[2]; d->nb13 = b->nb[3];
    d->nb1 = op->nb[1]; d->nb2 = op->nb[2]; d->nb3 = op->nb[3];
    if (b->ne[0] != a->ne[0] || op->ne[0] != a->ne[1] || op->ne[1] != b->ne[1] || op->ne[2] != b->ne[2] || op->ne[3] != b->ne[3]) return false;
    return b200_mul_mat_supported(d) != 0;
}

bool fill_mul_mat_id(const ggml_tensor * op, b200_mul_mat_id_desc * d) {
    const ggml_tensor * a = op->src[0], * b = op->src[1], * ids = op->src[2];
    if (!a || !b || !ids || !weight_type_ok(a->type) || b->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || ids->type != GGML_TYPE_I32) return false;
    if (a->nb[0] != ggml_type_size(a->type) || b->nb[0] != sizeof(float) || op->nb[0] != sizeof(float) || ids->nb[0] != sizeof(int32_t)) return false;
    if (a->ne[3] != 1 || b->ne[3] != 1 || op->ne[3] != 1 || ids->ne[2] != 1 || ids->ne[3] != 1) return false;
    if (b->ne[0] != a->ne[0] || op->ne[0] != a->ne[1] || op->ne[1] != ids->ne[0] || op->ne[2] != b->ne[2] || ids->ne[1] != b->ne[2]) return false;
    if (ids->ne[0] % b->ne[1] != 0) return false;  // broadcast rule of ggml_mul_mat_id (ggml.c:3194-3225)
    d->type = (int32_t) a->type;
    d->ne00 = a->ne[0]; d->ne01 = a->ne[1]; d->n_as = a->ne[2];
    d->nb01 = a->nb[1]; d->nb02 = a->nb[2];
    d->ne11 = b->ne[1]; d->n_tok = b->ne[2];
    d->nb11 = b->nb[1]; d->nb12 = b->nb[2];
    d->n_used = ids->ne[0]; d->ids_nb1 = ids->nb[1];
    d->nb1 = op->nb[1]; d->nb2 = op->nb[2];
    return b200_mul_mat_id_supported(d) != 0;
}

// ------------------------------------------------------------------------------------------------ backend (stream)

const char * be_name(ggml_backend_t backend) { return ((be_ctx *) backend->context)->name.c_str(); }
void timing_collect(be_ctx * c);
void be_free(ggml_backend_t backend) {
    be_ctx * c = (be_ctx *) backend->context;
    for (int id = 0; id < B200_MAX_DEVICES; id++) {
        split_peer & sp = c->peers[id];
        if (!sp.ctx) continue;
        b200_ctx_synchronize(sp.ctx);
        if (sp.x) b200_free(id, sp.x);
        if (sp.d) b200_free(id, sp.d);
        if (sp.event) b200_event_destroy(id, sp.event);
        b200_ctx_destroy(sp.ctx);
    }
    if (c->blk_scratch) b200_free(c->device, c->blk_scratch);
    if (c->split_event) b200_event_destroy(c->device, c->split_event);
    if (c->copy_event) b200_event_destroy(c->device, c->copy_event);
    b200_ctx_synchronize(c->ctx);
    if (c->timing) {
        timing_collect(c);
        for (const be_ctx::tstat * ts : { &c->t_replay, &c->t_direct }) {
            fprintf(stderr, "ggml-b200 timing: %s: %llu graph_compute calls; host per call: key %.1f us, launch %.1f us; device per call: %.1f us\n",
                    ts == &c->t_replay ? "graph replays" : "direct launches", (unsigned long long) ts->n, ts->n ? ts->key_ns / 1e3 / ts->n : 0.0,
                    ts->n ? ts->run_ns / 1e3 / ts->n : 0.0, ts->gpu_n ? ts->gpu_ms * 1e3 / ts->gpu_n : 0.0);
        }
        if (c->t_ev0) { b200_event_destroy(c->device, c->t_ev0); b200_event_destroy(c->device, c->t_ev1); }
    }
    for (auto & kv : c->graphs) for (auto & g : kv.second) b200_graph_destroy(g.exec);
    b200_ctx_destroy(c->ctx);
    delete c;
    delete backend;
}
void be_set_tensor_async(ggml_backend_t backend, ggml_tensor * t, const void * data, size_t offset, size_t size) {
    if (t->data && ggml_is_quantized(t->type)) b200_weight_cache_invalidate(((be_ctx *) backend->context)->device, t->data, ggml_nbytes(t));
    CHECK(b200_memcpy_h2d_async(((be_ctx *) backend->context)->ctx, (char *) t->data + offset, data, size));
}
void be_get_tensor_async(ggml_backend_t backend, const ggml_tensor * t, void * data, size_t offset, size_t size) {
    CHECK(b200_memcpy_d2h_async(((be_ctx *) backend->context)->ctx, data, (const char *) t->data + offset, size));
}
const char * be_name(ggml_backend_t backend);
// copy between two B200 backends (streams), possibly on different devices, without a host bounce
// (ggml_backend_cuda_cpy_tensor_async, ggml-cuda.cu:2773-2826): issued on the SOURCE backend's stream, which has produced src; the
// destination stream then waits for an event. Anything else returns false and the caller falls back to a synchronous copy
// (ggml-backend.cpp:408-429).
bool be_cpy_tensor_async(ggml_backend_t backend_src, ggml_backend_t backend_dst, const ggml_tensor * src, ggml_tensor * dst) {
    if (backend_src->iface.get_name != be_name || backend_dst->iface.get_name != be_name) return false;
    ggml_backend_buffer_t sb = src->view_src ? src->view_src->buffer : src->buffer;
    ggml_backend_buffer_t db = dst->view_src ? dst->view_src->buffer : dst->buffer;
    if (!sb || !db || !buft_is_b200(sb->buft) || !buft_is_b200(db->buft)) return false;
    be_ctx * cs = (be_ctx *) backend_src->context, * cd = (be_ctx *) backend_dst->context;
    buf_ctx * bs = (buf_ctx *) sb->context, * bd = (buf_ctx *) db->context;
    if (bs->device != cs->device || bd->device != cd->device) return false;  // a backend only touches buffers of its own device (:2797-2803)
    if (ggml_is_quantized(dst->type)) b200_weight_cache_invalidate(bd->device, dst->data, ggml_nbytes(dst));
    if (backend_src == backend_dst) {
        CHECK(b200_memcpy_d2d_async(cs->ctx, dst->data, src->data, ggml_nbytes(dst)));
        return true;
    }
    if (cs->device != cd->device) {
        b200_enable_peer_access(cs->device, cd->device);  // best effort: the driver stages the copy otherwise
        b200_enable_peer_access(cd->device, cs->device);
    }
    CHECK(b200_memcpy_peer_async(cs->ctx, dst->data, cd->device, src->data, cs->device, ggml_nbytes(dst)));
    if (!cs->copy_event) CHECK(b200_event_create(cs->device, &cs->copy_event));
    CHECK(b200_event_record(cs->ctx, cs->copy_event));
    CHECK(b200_event_wait(cd->ctx, cs->copy_event));
    return true;
}
void be_synchronize(ggml_backend_t backend) { CHECK(b200_ctx_synchronize(((be_ctx *) backend->context)->ctx)); }

bool ranges_overlap(const ggml_tensor * a, const ggml_tensor * b) {
    const char * a0 = (const char *) a->data, * b0 = (const char *) b->data;
    return a0 < b0 + ggml_nbytes(b) && b0 < a0 + ggml_nbytes(a);
}

bool grow(int device, void ** buf, size_t * cur, size_t need) {
    if (*cur >= need) return true;
    if (*buf) b200_free(device, *buf);
    *buf = nullptr; *cur = 0;
    if (b200_malloc(device, need, buf) != B200_OK) return false;
    *cur = need;
    return true;
}

// MUL_MAT whose src0 lives in a split buffer (ggml_cuda_op_mul_mat for split tensors, ggml-cuda.cu:1447-1765): every device
// multiplies its row slice by the whole src1. The main device (this backend's) works in place: its slice of dst is written straight
// into node->data with the full column stride. For every other device src1 is copied over once the main stream has produced it,
// the slice is computed into scratch [N][rows] on that device's own stream, copied back into its rows of dst as a 2-D copy
// (:1721-1734) and the main stream waits for that copy.
ggml_status split_mul_mat(be_ctx * c, ggml_tensor * node, const b200_mul_mat_desc & full) {
    const ggml_tensor * a = node->src[0], * b = node->src[1];
    const split_extra * e = (const split_extra *) a->extra;
    const split_buft_ctx * bc = (const split_buft_ctx *) a->buffer->buft->context;
    GGML_ASSERT(e && bc->main_device == c->device && "a split tensor is computed by the backend of its main device");
    const int64_t N = full.ne11, K = full.ne00;
    const size_t x_bytes = (size_t) N * K * sizeof(float);
    if (!c->split_event) CHECK(b200_event_create(c->device, &c->split_event));
    CHECK(b200_event_record(c->ctx, c->split_event));
    for (int id = 0; id < split_dev_count(); id++) {
        const int64_t rows = e->row_high[id] - e->row_low[id];
        if (rows <= 0) continue;
        b200_mul_mat_desc d = full;
        d.ne01 = rows;
        if (id == c->device) {
            CHECK(b200_mul_mat(c->ctx, &d, e->data[id], (const float *) b->data, (float *) node->data + e->row_low[id]));
            continue;
        }
        split_peer & sp = c->peers[id];
        if (!sp.ctx) {
            if (b200_ctx_create(id, nullptr, &sp.ctx) != B200_OK) {
                GGML_LOG_ERROR("ggml-b200: cannot create a context on device %d for a split tensor: %s\n", id, b200_last_error());
                return GGML_STATUS_FAILED;
            }
            b200_ctx_set_weight_cache(sp.ctx, c->weight_cache);
            CHECK(b200_event_create(id, &sp.event));
            // with peer access the other device's kernels write their rows of dst where they belong - in the main device's memory,
            // through NVLink - instead of into scratch that is then copied over (the reference's cudaMemcpy2DPeerAsync per device,
            // ggml-cuda.cu:1721-1734); without it the copies are staged by the driver
            const bool to_main = b200_enable_peer_access(id, c->device) == B200_OK;
            const bool from_main = b200_enable_peer_access(c->device, id) == B200_OK;
            const char * nd = getenv("GGML_B200_SPLIT_NO_DIRECT");
            sp.direct = to_main && from_main && !(nd && atoi(nd));
        }
        if (!grow(id, &sp.x, &sp.x_size, x_bytes)) return GGML_STATUS_ALLOC_FAILED;
        // (prefill-sized results go through scratch + one coalesced 2-D copy: 4-byte epilogue stores over NVLink ran 2-3x slower)
        const bool direct = sp.direct && N <= 8;
        if (!direct && !grow(id, &sp.d, &sp.d_size, (size_t) N * rows * sizeof(float))) return GGML_STATUS_ALLOC_FAILED;
        CHECK(b200_event_wait(sp.ctx, c->split_event));
        // src1 rows may be strided on the main device: pack them into [N][K] on the peer
        CHECK(b200_memcpy2d_async(sp.ctx, sp.x, (size_t) K * sizeof(float), b->data, full.nb11, (size_t) K * sizeof(float), (size_t) N));
        d.nb11 = (size_t) K * sizeof(float); d.nb12 = d.nb11 * N; d.nb13 = d.nb12;
        if (direct) {   // dst keeps the full tensor's strides: row slice [row_low, row_high) of every column, on the main device
            CHECK(b200_mul_mat(sp.ctx, &d, e->data[id], (const float *) sp.x, (float *) node->data + e->row_low[id]));
        } else {
            d.nb1 = (size_t) rows * sizeof(float); d.nb2 = d.nb1 * N; d.nb3 = d.nb2;
            CHECK(b200_mul_mat(sp.ctx, &d, e->data[id], (const float *) sp.x, (float *) sp.d));
            CHECK(b200_memcpy2d_async(sp.ctx, (float *) node->data + e->row_low[id], full.nb1, sp.d, (size_t) rows * sizeof(float),
                                      (size_t) rows * sizeof(float), (size_t) N));
        }
        CHECK(b200_event_record(sp.ctx, sp.event));
    }
    for (int id = 0; id < split_dev_count(); id++) {
        if (id != c->device && e->row_high[id] > e->row_low[id]) CHECK(b200_event_wait(c->ctx, c->peers[id].event));
    }
    return GGML_STATUS_SUCCESS;
}

// ------------------------------------------------------------------------------------------------ K7: GLU / ADD / ADD_ID nodes

std::atomic<uint64_t> g_fused_epilogues{ 0 };  // fused K7 launches of this process (proc address "ggml_backend_b200_fused_epilogues": tests)
uint64_t fused_epilogues() { return g_fused_epilogues.load(std::memory_order_relaxed); }

bool f32_rows(const ggml_tensor * t) { return t && t->type == GGML_TYPE_F32 && t->nb[0] == sizeof(float); }

// GGML_OP_GLU as ggml-cpu takes it (ops.cpp:2934-3290): F32, rows contiguous_1, split (src1) or single-tensor halves
bool glu_ok(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    const int g = ggml_get_op_params_i32(op, 0);
    if (g < 0 || g > B200_GLU_GEGLU_QUICK) return false;
    if (!f32_rows(a) || !f32_rows(op) || !ggml_is_contiguous_1(a) || !ggml_is_contiguous_1(op)) return false;
    if (b && (!f32_rows(b) || !ggml_is_contiguous_1(b) || !ggml_are_same_shape(a, b))) return false;
    return op->ne[0] == (b ? a->ne[0] : a->ne[0] / 2) && ggml_nrows(op) == ggml_nrows(a);
}
bool add_ok(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    return f32_rows(a) && f32_rows(b) && f32_rows(op) && ggml_are_same_shape(a, op) && ggml_can_repeat(b, a);
}
bool add_id_ok(const ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1], * ids = op->src[2];
    if (!f32_rows(a) || !f32_rows(b) || !f32_rows(op) || !ids || ids->type != GGML_TYPE_I32 || ids->nb[0] != sizeof(int32_t)) return false;
    if (!ggml_are_same_shape(a, op) || a->ne[3] != 1 || b->ne[0] != a->ne[0] || b->ne[2] != 1 || b->ne[3] != 1) return false;
    return ids->ne[0] == a->ne[1] && ids->ne[1] == a->ne[2] && ids->ne[2] == 1 && ids->ne[3] == 1;
}
bool add_like_ok(const ggml_tensor * t) { return t->op == GGML_OP_ADD ? add_ok(t) : add_id_ok(t); }

void run_glu(be_ctx * c, ggml_tensor * node) {
    const ggml_tensor * a = node->src[0], * b = node->src[1];
    const int64_t nc      = node->ne[0];
    const bool    swapped = ggml_get_op_params_i32(node, 1) != 0;
    const float * gate = (const float *) a->data, * up = b ? (const float *) b->data : (const float *) a->data;
    if (!b) (swapped ? gate : up) += nc;  // single tensor: [gate | up] halves of every row, or [up | gate] when swapped
    CHECK_RUN(c, b200_glu(c->ctx, ggml_get_op_params_i32(node, 0), ggml_get_op_params_f32(node, 2), ggml_get_op_params_f32(node, 3), gate, a->nb[1], up,
                       b ? b->nb[1] : a->nb[1], (float *) node->data, node->nb[1], nc, ggml_nrows(node)));
}
void run_add(be_ctx * c, ggml_tensor * node) {
    const ggml_tensor * a = node->src[0], * b = node->src[1];
    CHECK_RUN(c, (node->op == GGML_OP_MUL ? b200_mul : b200_add)(c->ctx, node->ne, (const float *) a->data, a->nb, (const float *) b->data, b->ne, b->nb, (float *) node->data, node->nb));
}
void run_add_id(be_ctx * c, ggml_tensor * node) {
    const ggml_tensor * a = node->src[0], * b = node->src[1], * ids = node->src[2];
    CHECK_RUN(c, b200_add_id(c->ctx, (const float *) a->data, a->nb[1], a->nb[2], (const float *) b->data, b->nb[1], b->ne[1], (const int32_t *) ids->data,
                      ids->nb[1], (float *) node->data, node->nb[1], node->nb[2], a->ne[0], a->ne[1], a->ne[2]));
}

// the bias operand of the ADD / ADD_ID that follows MUL_MAT(_ID) `mm` (either operand of an ADD; src1 of an ADD_ID), else nullptr
const ggml_tensor * bias_of(const ggml_tensor * add, const ggml_tensor * mm) {
    if (add->op == GGML_OP_ADD) return add->src[0] == mm ? add->src[1] : (add->src[1] == mm ? add->src[0] : nullptr);
    return add->src[0] == mm && add->src[2] == mm->src[2] ? add->src[1] : nullptr;
}

// K7 peephole (the reference's: ggml_cuda_can_fuse + ggml_cuda_should_fuse_mul_mat, ggml-cuda.cu:2020-2104, 3087-3115, 3404-3600):
// node i is a one-column MUL_MAT / one-token MUL_MAT_ID. Recognises {mm, add, mm, add, GLU}, {mm, mm, GLU} and {mm, add} whose
// intermediate results have no other consumer and runs them as ONE launch. Returns the number of nodes consumed (0: no fusion).
// an RMS_NORM (+ MUL) whose only readers are the MUL_MATs about to be launched: they can take the un-normalized row and normalize it
// themselves (b200_mul_mat_*_norm)
struct norm_src {
    const ggml_tensor * x0;   // the norm's input
    b200_norm           n;
};

int try_fuse_epilogue(be_ctx * c, ggml_cgraph * cgraph, int i, const norm_src * nf = nullptr) {
    ggml_tensor * n0 = cgraph->nodes[i];
    const ggml_op mm = n0->op, bias_op = mm == GGML_OP_MUL_MAT ? GGML_OP_ADD : GGML_OP_ADD_ID;
    if (mm == GGML_OP_MUL_MAT ? n0->ne[1] != 1 : n0->ne[2] != 1) return 0;
    ggml_tensor * gate = nullptr, * up = nullptr, * glu = nullptr, * last = nullptr;
    const ggml_tensor * gate_add = nullptr, * up_add = nullptr;
    int count = 0;
    if (ggml_can_fuse_subgraph(cgraph, i, { mm, bias_op, mm, bias_op, GGML_OP_GLU }, { i + 4 })) {
        glu = cgraph->nodes[i + 4];
        ggml_tensor * a = cgraph->nodes[i], * b = cgraph->nodes[i + 2];
        if (!glu->src[0] || !glu->src[1]) return 0;
        if (bias_of(glu->src[0], a) && bias_of(glu->src[1], b)) { gate = a; up = b; }
        else if (bias_of(glu->src[0], b) && bias_of(glu->src[1], a)) { gate = b; up = a; }
        else return 0;
        gate_add = glu->src[0]; up_add = glu->src[1]; last = glu; count = 5;
    } else if (ggml_can_fuse_subgraph(cgraph, i, { mm, mm, GGML_OP_GLU }, { i + 2 })) {
        glu = cgraph->nodes[i + 2];
        ggml_tensor * a = cgraph->nodes[i], * b = cgraph->nodes[i + 1];
        if (glu->src[0] == a && glu->src[1] == b) { gate = a; up = b; }
        else if (glu->src[0] == b && glu->src[1] == a) { gate = b; up = a; }
        else return 0;
        last = glu; count = 3;
    } else if (ggml_can_fuse(cgraph, i, { mm, bias_op })) {
        up = n0; up_add = cgraph->nodes[i + 1];
        if (!bias_of(up_add, up)) return 0;
        last = cgraph->nodes[i + 1]; count = 2;
    } else {
        return 0;
    }
    const ggml_tensor * w = up->src[0], * x = up->src[1];
    if (w->buffer && buft_is_split(w->buffer->buft)) return 0;
    b200_fusion f = {};
    f.glu_op = B200_GLU_NONE;
    if (glu) {
        const ggml_tensor * wg = gate->src[0];
        if (!glu_ok(glu) || ggml_get_op_params_i32(glu, 1) != 0) return 0;  // swapped operands: not fused (as ggml-cuda.cu:2090-2092)
        if (wg->type != w->type || !ggml_are_same_shape(wg, w) || !ggml_are_same_stride(wg, w) || gate->src[1] != x || gate->src[2] != up->src[2]) return 0;
        if (wg->buffer && buft_is_split(wg->buffer->buft)) return 0;
        f.glu_op = ggml_get_op_params_i32(glu, 0);
        f.alpha  = ggml_get_op_params_f32(glu, 2);
        f.limit  = ggml_get_op_params_f32(glu, 3);
    }
    const ggml_tensor * biases[2] = { up_add ? bias_of(up_add, up) : nullptr, gate_add ? bias_of(gate_add, gate) : nullptr };
    for (const ggml_tensor * b : biases) {
        if (!b) continue;
        if (!f32_rows(b) || ranges_overlap(last, b)) return 0;
        if (mm == GGML_OP_MUL_MAT ? !ggml_are_same_shape(b, up) : (b->ne[0] != up->ne[0] || b->ne[1] != w->ne[2] || b->ne[2] != 1 || b->ne[3] != 1)) return 0;  // no repeating adds
    }
    if ((up_add && !add_like_ok(up_add)) || (gate_add && !add_like_ok(gate_add))) return 0;
    if (biases[0] && biases[1] && (biases[0]->nb[1] != biases[1]->nb[1] || biases[0]->nb[2] != biases[1]->nb[2] || biases[0]->nb[3] != biases[1]->nb[3])) return 0;
    // the fused launch writes `last` while other CTAs still read src1 / ids: the allocator may have given it their memory
    if (ranges_overlap(last, x) || (up->src[2] && ranges_overlap(last, up->src[2])) || !f32_rows(last)) return 0;
    const ggml_tensor * bn = biases[0] ? biases[0] : biases[1];
    f.up_bias   = biases[0] ? (const float *) biases[0]->data : nullptr;
    f.gate_bias = biases[1] ? (const float *) biases[1]->data : nullptr;
    b200_status st;
    if (mm == GGML_OP_MUL_MAT) {
        b200_mul_mat_desc d;
        if (!fill_mul_mat(up, &d)) return 0;
        d.nb1 = last->nb[1]; d.nb2 = last->nb[2]; d.nb3 = last->nb[3];
        if (bn) { f.bias_nb2 = bn->nb[2]; f.bias_nb3 = bn->nb[3]; }
        if (!b200_mul_mat_fused_supported(c->ctx, &d)) return 0;
        if (nf) {
            if (!glu || ranges_overlap(last, nf->x0)) return 0;
            st = b200_mul_mat_fused_norm(c->ctx, &d, w->data, gate->src[0]->data, (const float *) nf->x0->data, &nf->n, &f, (float *) last->data);
        } else
        st = b200_mul_mat_fused(c->ctx, &d, w->data, gate ? gate->src[0]->data : nullptr, (const float *) x->data, &f, (float *) last->data);
    } else {
        b200_mul_mat_id_desc d;
        if (!fill_mul_mat_id(up, &d)) return 0;
        d.nb1 = last->nb[1]; d.nb2 = last->nb[2];
        if (bn) f.bias_nb1 = bn->nb[1];
        if (nf || !b200_mul_mat_id_fused_supported(c->ctx, &d)) return 0;
        st = b200_mul_mat_id_fused(c->ctx, &d, w->data, gate ? gate->src[0]->data : nullptr, (const float *) x->data, (const int32_t *) up->src[2]->data, &f,
                                   (float *) last->data);
    }
    if (st == B200_ERR_UNSUPPORTED) return 0;  // e.g. a weight pointer the TMA-staged kernel cannot take: run the nodes one by one
    CHECK_RUN(c, st);
    return count;
}

// ------------------------------------------------------------------------------------------------ f3: the ops between the matmuls
// SURVEY 8(f) rank 3, first slice (ops.cu): GET_ROWS, RMS_NORM, ROPE, SET_ROWS, CPY / CONT / DUP, FLASH_ATTN_EXT. With them a Llama
// decode graph is ONE split on this backend. GGML_B200_F3=0 declines them all (the scheduler then leaves them to ggml-cpu as before).
bool exact_mode();
bool f3_on() {
    static const bool on = [] { const char * s = getenv("GGML_B200_F3"); return !(s && atoi(s) == 0); }();
    return on;
}
b200_tensor as_b200(const ggml_tensor * t) {
    b200_tensor r;
    r.data = t->data;
    r.type = (int32_t) t->type;
    for (int i = 0; i < 4; i++) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; }
    return r;
}
b200_rope_params rope_params_of(const ggml_tensor * op) {
    b200_rope_params rp;
    rp.n_dims      = ggml_get_op_params_i32(op, 1);
    rp.mode        = ggml_get_op_params_i32(op, 2);
    rp.n_ctx_orig  = ggml_get_op_params_i32(op, 4);
    rp.freq_base   = ggml_get_op_params_f32(op, 5);
    rp.freq_scale  = ggml_get_op_params_f32(op, 6);
    rp.ext_factor  = ggml_get_op_params_f32(op, 7);
    rp.attn_factor = ggml_get_op_params_f32(op, 8);
    rp.beta_fast   = ggml_get_op_params_f32(op, 9);
    rp.beta_slow   = ggml_get_op_params_f32(op, 10);
    return rp;
}
// what dev_supports_op answers for the f3 ops; graph_compute runs exactly what this accepts
bool f3_ok(const ggml_tensor * op) {
    if (!f3_on() || exact_mode()) return false;
    const b200_tensor d = as_b200(op);
    b200_tensor s[5];
    for (int i = 0; i < 5; i++) if (op->src[i]) s[i] = as_b200(op->src[i]);
    switch (op->op) {
        case GGML_OP_RMS_NORM:
            return op->src[0] && b200_rms_norm_supported(&s[0], &d);
        case GGML_OP_GET_ROWS:
            return op->src[0] && op->src[1] && b200_get_rows_supported(&s[0], &s[1], &d);
        case GGML_OP_ROPE: {
            if (!op->src[0] || !op->src[1]) return false;
            const b200_rope_params rp = rope_params_of(op);
            return b200_rope_supported(&s[0], &s[1], op->src[2] ? &s[2] : nullptr, &d, &rp);
        }
        case GGML_OP_SET_ROWS:
            return op->src[0] && op->src[1] && b200_set_rows_supported(&s[0], &s[1], &d);
        case GGML_OP_CPY:   // dst = src[1]'s layout (the node is a view of it); CONT / DUP: the node's own
        case GGML_OP_CONT:
        case GGML_OP_DUP:
            return op->src[0] && b200_cpy_supported(&s[0], &d);
        case GGML_OP_FLASH_ATTN_EXT:
            return op->src[0] && op->src[1] && op->src[2] &&
                   b200_flash_attn_ext_supported(&s[0], &s[1], &s[2], op->src[3] ? &s[3] : nullptr, op->src[4] ? &s[4] : nullptr, &d);
        default:
            return false;
    }
}
bool is_f3_op(ggml_op op) {
    return op == GGML_OP_RMS_NORM || op == GGML_OP_GET_ROWS || op == GGML_OP_ROPE || op == GGML_OP_SET_ROWS || op == GGML_OP_CPY || op == GGML_OP_CONT ||
           op == GGML_OP_DUP || op == GGML_OP_FLASH_ATTN_EXT;
}
void run_f3(be_ctx * c, ggml_tensor * op, const ggml_tensor * hint_idx = nullptr) {
    const b200_tensor d = as_b200(op);
    b200_tensor s[5];
    for (int i = 0; i < 5; i++) if (op->src[i]) s[i] = as_b200(op->src[i]);
    switch (op->op) {
        case GGML_OP_RMS_NORM:
            CHECK_RUN(c, b200_rms_norm(c->ctx, &s[0], &d, ggml_get_op_params_f32(op, 0)));
            break;
        case GGML_OP_GET_ROWS:
            CHECK_RUN(c, b200_get_rows(c->ctx, &s[0], &s[1], &d));
            break;
        case GGML_OP_ROPE: {
            const b200_rope_params rp = rope_params_of(op);
            CHECK_RUN(c, b200_rope(c->ctx, &s[0], &s[1], op->src[2] ? &s[2] : nullptr, &d, &rp));
            break;
        }
        case GGML_OP_SET_ROWS:
            CHECK_RUN(c, b200_set_rows(c->ctx, &s[0], &s[1], &d));
            break;
        case GGML_OP_CPY:
        case GGML_OP_CONT:
        case GGML_OP_DUP:
            CHECK_RUN(c, b200_cpy(c->ctx, &s[0], &d));
            break;
        case GGML_OP_FLASH_ATTN_EXT:
            if (hint_idx) {
                const b200_tensor hi = as_b200(hint_idx);
                CHECK_RUN(c, b200_flash_attn_ext_decode(c->ctx, &s[0], &s[1], &s[2], op->src[3] ? &s[3] : nullptr, op->src[4] ? &s[4] : nullptr, &d,
                                                        ggml_get_op_params_f32(op, 0), ggml_get_op_params_f32(op, 1), ggml_get_op_params_f32(op, 2), &hi));
                break;
            }
            CHECK_RUN(c, b200_flash_attn_ext(c->ctx, &s[0], &s[1], &s[2], op->src[3] ? &s[3] : nullptr, op->src[4] ? &s[4] : nullptr, &d,
                                             ggml_get_op_params_f32(op, 0), ggml_get_op_params_f32(op, 1), ggml_get_op_params_f32(op, 2)));
            break;
        default:
            break;
    }
}

// next node from `from` on that still has work to do (-1: none)
int next_live(const be_ctx * c, const ggml_cgraph * cgraph, int from) {
    for (int j = from; j < cgraph->n_nodes; j++) if (!c->done[j] && !is_noop(cgraph->nodes[j])) return j;
    return -1;
}

bool try_fuse_attn_block(be_ctx * c, ggml_cgraph * cgraph, int i, const norm_src * nf = nullptr);
int32_t use_count(const ggml_cgraph * cgraph, const ggml_tensor * t);

// RMS_NORM at node i followed by MUL with a per-channel weight: one launch (build_norm's pair) - or none at all, when the only readers
// of the result are the one-token MUL_MATs that follow (q/k/v of the attention block; the gate / up pair of a GLU FFN), which then
// normalize the row themselves. Returns true when node i has been dealt with (the nodes consumed are marked done).
bool try_fuse_norm_mul(be_ctx * c, ggml_cgraph * cgraph, int i) {
    if (!c->fuse_f3 || i + 1 >= cgraph->n_nodes) return false;
    ggml_tensor * norm = cgraph->nodes[i], * mul = cgraph->nodes[i + 1];
    if (mul->op != GGML_OP_MUL || c->done[i + 1] || !ggml_can_fuse(cgraph, i, { GGML_OP_RMS_NORM, GGML_OP_MUL })) return false;
    const ggml_tensor * w = mul->src[0] == norm ? mul->src[1] : (mul->src[1] == norm ? mul->src[0] : nullptr);
    if (!w || w == norm || !ggml_are_same_shape(mul, norm)) return false;
    const ggml_tensor * x = norm->src[0];
    if (ranges_overlap(mul, w) || (ranges_overlap(mul, x) && mul->data != x->data)) return false;
    const b200_tensor bx = as_b200(x), bw = as_b200(w), bd = as_b200(mul);
    if (!b200_rms_norm_mul_supported(&bx, &bw, &bd)) return false;
    // the readers of the normalized row
    const int jm = next_live(c, cgraph, i + 2);
    if (jm >= 0 && c->fuse && cgraph->nodes[jm]->op == GGML_OP_MUL_MAT && cgraph->nodes[jm]->src[1] == mul && mul->ne[1] == 1 && ggml_is_contiguous(mul) &&
        ggml_is_contiguous(x) && ggml_is_contiguous(w) && ggml_nelements(w) == mul->ne[0] && !(mul->flags & GGML_TENSOR_FLAG_OUTPUT) && mul->data != x->data) {
        norm_src nf;
        nf.x0 = x;
        nf.n.w = (const float *) w->data; nf.n.eps = ggml_get_op_params_f32(norm, 0); nf.n.y = (float *) mul->data; nf.n.y_nb1 = mul->nb[1];
        const int32_t uses = use_count(cgraph, mul);
        if (uses == 3 && try_fuse_attn_block(c, cgraph, jm, &nf)) {
            c->done[i + 1] = c->done[jm] = 1;
            return true;
        }
        if (uses == 2 && c->fuse_epilogue) {
            const int fused = try_fuse_epilogue(c, cgraph, jm, &nf);
            if (fused > 0) {
                g_fused_epilogues.fetch_add(1, std::memory_order_relaxed);
                c->done[i + 1] = 1;
                for (int k = 0; k < fused; k++) c->done[jm + k] = 1;
                return true;
            }
        }
    }
    CHECK_RUN(c, b200_rms_norm_mul(c->ctx, &bx, &bw, &bd, ggml_get_op_params_f32(norm, 0)));
    c->done[i + 1] = 1;
    return true;
}

// ROPE(q) at node i, then - as the next nodes with work - ROPE(k), SET_ROWS(k cache <- rope(k)), SET_ROWS(v cache <- v): the attention
// inputs of a layer as build_attn lays them out (src/llama-graph.cpp:1593-1606) in one launch. Marks the three followers done.
bool try_fuse_attn_prep(be_ctx * c, ggml_cgraph * cgraph, int i) {
    if (!c->fuse_f3) return false;
    ggml_tensor * rq = cgraph->nodes[i];
    const int j1 = next_live(c, cgraph, i + 1), j2 = j1 < 0 ? -1 : next_live(c, cgraph, j1 + 1), j3 = j2 < 0 ? -1 : next_live(c, cgraph, j2 + 1);
    if (j3 < 0) return false;
    ggml_tensor * rk = cgraph->nodes[j1], * sk = cgraph->nodes[j2], * sv = cgraph->nodes[j3];
    if (rk->op != GGML_OP_ROPE || sk->op != GGML_OP_SET_ROWS || sv->op != GGML_OP_SET_ROWS) return false;
    if (memcmp(rq->op_params, rk->op_params, sizeof(rq->op_params)) != 0 || rq->src[1] != rk->src[1] || rq->src[2] != rk->src[2]) return false;
    const ggml_tensor * ks = sk->src[0];   // must be rope(k)'s output seen as [n_embd_k_gqa, n_tokens]
    if (!ks || ks->data != rk->data || ks->type != GGML_TYPE_F32 || ks->ne[0] != rk->ne[0] * rk->ne[1] || ks->ne[1] != rk->ne[2] || ks->ne[2] != 1 ||
        ks->ne[3] != 1 || ks->nb[1] != rk->nb[2] || rk->ne[3] != 1) return false;
    const ggml_tensor * vs = sv->src[0];
    if (!vs || ranges_overlap(vs, rq) || ranges_overlap(vs, rk) || ranges_overlap(rk->src[0], rq) || ranges_overlap(rq, rk)) return false;
    if (ranges_overlap(sk, sv) || ranges_overlap(sk, rq) || ranges_overlap(sk, rk) || ranges_overlap(sv, rq) || ranges_overlap(sv, rk)) return false;
    if (ranges_overlap(sk, rq->src[0]) || ranges_overlap(sk, rk->src[0]) || ranges_overlap(sv, rq->src[0]) || ranges_overlap(sv, rk->src[0]) || ranges_overlap(sk, vs)) return false;
    const b200_tensor q = as_b200(rq->src[0]), qd = as_b200(rq), k = as_b200(rk->src[0]), kd = as_b200(rk), pos = as_b200(rq->src[1]);
    b200_tensor ff;
    if (rq->src[2]) ff = as_b200(rq->src[2]);
    const b200_tensor ki = as_b200(sk->src[1]), kc = as_b200(sk), v = as_b200(vs), vi = as_b200(sv->src[1]), vc = as_b200(sv);
    const b200_rope_params rp = rope_params_of(rq);
    if (!b200_attn_prep_supported(&q, &qd, &k, &kd, &pos, rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc)) return false;
    CHECK_RUN(c, b200_attn_prep(c->ctx, &q, &qd, &k, &kd, &pos, rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc));
    c->done[j1] = c->done[j2] = c->done[j3] = 1;
    return true;
}

int32_t use_count(const ggml_cgraph * cgraph, const ggml_tensor * t) {   // ggml_node_get_use_count by tensor
    const size_t pos = ggml_hash_find(&cgraph->visited_hash_set, t);
    return ggml_bitset_get(cgraph->visited_hash_set.used, pos) ? cgraph->use_counts[pos] : 0;
}
// t is `root` seen through reshapes / views that nobody else reads, same bytes, contiguous: what a kernel reading t would read from
// any other contiguous copy of root's result
bool sole_contiguous_view_of(const ggml_cgraph * cgraph, const ggml_tensor * t, const ggml_tensor * root) {
    if (!t || t->data != root->data || !ggml_is_contiguous(t) || ggml_nelements(t) != ggml_nelements(root) || t->type != root->type) return false;
    for (int depth = 0; t != root; depth++) {
        if (depth > 4 || !(t->op == GGML_OP_RESHAPE || t->op == GGML_OP_VIEW) || !t->src[0] || use_count(cgraph, t) != 1 || (t->flags & GGML_TENSOR_FLAG_OUTPUT)) return false;
        t = t->src[0];
    }
    return use_count(cgraph, root) == 1 && !(root->flags & GGML_TENSOR_FLAG_OUTPUT);
}

// The attention inputs of a layer as build_attn lays them out (src/models/llama.cpp:36-75, src/llama-graph.cpp:1593-1606):
//   MUL_MAT(wq), ROPE(q), MUL_MAT(wv), MUL_MAT(wk), ROPE(k), SET_ROWS(k cache), SET_ROWS(v cache)      (reshapes / views in between)
// run as TWO launches: the three projections of the shared activations into a scratch buffer of the backend, then b200_attn_prep
// from there. The graph allocator gives Kcur the memory of Qcur (dead after ROPE(q)), so the projections cannot simply be hoisted
// into one launch at their graph addresses; nothing but the two ROPEs / the V store reads them (checked), so they never need to exist
// there. Returns true when the seven nodes have been run (the six followers are marked done).
bool try_fuse_attn_block(be_ctx * c, ggml_cgraph * cgraph, int i, const norm_src * nf) {
    if (!c->fuse || !c->fuse_f3 || !f3_on() || exact_mode()) return false;
    int j[7] = { i };
    for (int k = 1; k < 7; k++) { j[k] = next_live(c, cgraph, j[k - 1] + 1); if (j[k] < 0) return false; }
    ggml_tensor * mq = cgraph->nodes[j[0]], * rq = cgraph->nodes[j[1]], * ma = cgraph->nodes[j[2]], * mb = cgraph->nodes[j[3]], * rk = cgraph->nodes[j[4]],
                * sk = cgraph->nodes[j[5]], * sv = cgraph->nodes[j[6]];
    if (rq->op != GGML_OP_ROPE || ma->op != GGML_OP_MUL_MAT || mb->op != GGML_OP_MUL_MAT || rk->op != GGML_OP_ROPE || sk->op != GGML_OP_SET_ROWS ||
        sv->op != GGML_OP_SET_ROWS) return false;
    ggml_tensor * mk = rk->src[0] && rk->src[0]->data == ma->data ? ma : mb, * mv = mk == ma ? mb : ma;
    ggml_tensor * mm[3] = { mq, mk, mv };
    b200_mul_mat_desc d[3];
    for (int k = 0; k < 3; k++) {
        if (mm[k]->src[1] != mq->src[1] || mm[k]->src[0]->op != GGML_OP_NONE || !fill_mul_mat(mm[k], &d[k])) return false;
        if (mm[k]->src[0]->buffer && buft_is_split(mm[k]->src[0]->buffer->buft)) return false;
        if (d[k].type != d[0].type || d[k].ne00 != d[0].ne00 || d[k].ne02 != 1 || d[k].ne03 != 1 || d[k].ne12 != 1 || d[k].ne13 != 1 || !ggml_is_contiguous(mm[k])) return false;
    }
    if (!f3_ok(rq) || !f3_ok(rk) || !f3_ok(sk) || !f3_ok(sv)) return false;
    if (memcmp(rq->op_params, rk->op_params, sizeof(rq->op_params)) != 0 || rq->src[1] != rk->src[1] || rq->src[2] != rk->src[2]) return false;
    if (!sole_contiguous_view_of(cgraph, rq->src[0], mq) || !sole_contiguous_view_of(cgraph, rk->src[0], mk) || !sole_contiguous_view_of(cgraph, sv->src[0], mv)) return false;
    const ggml_tensor * ks = sk->src[0];   // rope(k)'s output seen as [n_embd_k_gqa, n_tokens]
    if (!ks || ks->data != rk->data || ks->type != GGML_TYPE_F32 || ks->ne[0] != rk->ne[0] * rk->ne[1] || ks->ne[1] != rk->ne[2] || ks->ne[2] != 1 ||
        ks->ne[3] != 1 || ks->nb[1] != rk->nb[2] || rk->ne[3] != 1) return false;
    // outputs that do land at their graph addresses: rope(q), rope(k), the cache rows. They are written at ROPE(q)'s position in the
    // order, after every MUL_MAT of the block has completed: the block's inputs (src1, positions, indices) must not be among them
    const ggml_tensor * outs[4] = { rq, rk, sk, sv };
    const ggml_tensor * ins[6] = { mq->src[1], rq->src[1], rq->src[2], sk->src[1], sv->src[1], nf ? nf->x0 : nullptr };
    for (const ggml_tensor * o : outs) {
        for (const ggml_tensor * in : ins) if (in && ranges_overlap(o, in)) return false;
        for (const ggml_tensor * o2 : outs) if (o != o2 && ranges_overlap(o, o2)) return false;
    }
    const int64_t N = d[0].ne11;
    size_t off[3], total = 0;
    for (int k = 0; k < 3; k++) { off[k] = total; total += ((size_t) d[k].ne01 * N * sizeof(float) + 255) / 256 * 256; }
    if (total > c->blk_scratch_size) {
        if (c->capturing) return false;   // sized by the uncaptured first run
        b200_ctx_synchronize(c->ctx);
        if (!grow(c->device, &c->blk_scratch, &c->blk_scratch_size, total)) return false;
    }
    const void * w[3];
    float *      o[3];
    for (int k = 0; k < 3; k++) {
        w[k] = mm[k]->src[0]->data;
        o[k] = (float *) ((char *) c->blk_scratch + off[k]);
        d[k].nb1 = (size_t) d[k].ne01 * sizeof(float); d[k].nb2 = d[k].nb3 = d[k].nb1 * (size_t) N;
    }
    // the descriptors b200_attn_prep would get node by node, with the three projections read from the scratch copies
    b200_tensor q = as_b200(rq->src[0]), k = as_b200(rk->src[0]), v = as_b200(sv->src[0]);
    q.data = o[0]; k.data = o[1]; v.data = o[2];
    const b200_tensor qd = as_b200(rq), kd = as_b200(rk), pos = as_b200(rq->src[1]), ki = as_b200(sk->src[1]), kc = as_b200(sk), vi = as_b200(sv->src[1]), vc = as_b200(sv);
    b200_tensor ff;
    if (rq->src[2]) ff = as_b200(rq->src[2]);
    const b200_rope_params rp = rope_params_of(rq);
    if (!b200_attn_prep_supported(&q, &qd, &k, &kd, &pos, rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc)) return false;
    // one token: the FLASH_ATTN_EXT that follows (behind, in the first layer, the CPY that casts the mask) joins the second launch
    int j_cpy = -1, j_fa = -1;
    b200_tensor fq, fk, fv, fm, fs, fd;
    ggml_tensor * fa = nullptr;
    // (opt-in, GGML_B200_ATTN_DECODE=1: measured 3 % SLOWER than attn_prep + FLASH_ATTN_EXT on the 8B step - every CTA of the attention
    // kernel repeats the rotation, which costs more than the launch it saves)
    static const bool attn_decode_on = [] { const char * e = getenv("GGML_B200_ATTN_DECODE"); return e && atoi(e) != 0; }();
    if (N == 1 && attn_decode_on) {
        int jn = next_live(c, cgraph, j[6] + 1);
        if (jn >= 0 && (cgraph->nodes[jn]->op == GGML_OP_CPY || cgraph->nodes[jn]->op == GGML_OP_CONT) && f3_ok(cgraph->nodes[jn])) {
            const ggml_tensor * cp = cgraph->nodes[jn];
            bool indep = true;   // it must not read anything this block writes
            for (const ggml_tensor * o2 : outs) indep = indep && !ranges_overlap(cp->src[0], o2) && !ranges_overlap(cp, o2);
            for (int k2 = 0; k2 < 3; k2++) indep = indep && !ranges_overlap(cp->src[0], mm[k2]);
            if (indep) { j_cpy = jn; jn = next_live(c, cgraph, jn + 1); }
        }
        if (jn >= 0 && cgraph->nodes[jn]->op == GGML_OP_FLASH_ATTN_EXT && f3_ok(cgraph->nodes[jn])) {
            fa = cgraph->nodes[jn];
            const ggml_tensor * aq = fa->src[0];
            bool ok = aq && aq->data == rq->data && fa->src[1] && fa->src[2];
            for (const ggml_tensor * o2 : outs) ok = ok && !ranges_overlap(fa, o2);
            for (const ggml_tensor * in : ins) ok = ok && (!in || !ranges_overlap(fa, in));
            for (int k2 = 3; k2 < 5; k2++) ok = ok && (!fa->src[k2] || !ranges_overlap(fa, fa->src[k2]));   // mask, sinks
            if (ok) {
                fq = as_b200(aq); fk = as_b200(fa->src[1]); fv = as_b200(fa->src[2]); fd = as_b200(fa);
                if (fa->src[3]) fm = as_b200(fa->src[3]);
                if (fa->src[4]) fs = as_b200(fa->src[4]);
                if (b200_attn_decode_supported(&q, &qd, &k, &kd, &pos, rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc, &fq, &fk, &fv,
                                               fa->src[3] ? &fm : nullptr, fa->src[4] ? &fs : nullptr, &fd)) j_fa = jn;
            }
        }
    }
    if (j_fa >= 0 && j_cpy >= 0) { run_f3(c, cgraph->nodes[j_cpy]); c->done[j_cpy] = 1; }
    if (j_fa < 0) {
        // projections + ROPEs + cache rows: one launch for one token where the mat-vec kernel can fold them into its store (b200_qkv_rope),
        // the projections followed by b200_attn_prep otherwise
        CHECK_RUN(c, b200_qkv_rope(c->ctx, d, w, (const float *) (nf ? nf->x0->data : mq->src[1]->data), nf ? &nf->n : nullptr, o, &q, &qd, &k, &kd, &pos,
                                   rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc));
        for (int k2 = 1; k2 < 7; k2++) c->done[j[k2]] = 1;
        if (N == 1 && sk->src[1] == sv->src[1]) {   // one token, one slot index for both caches: the attention directly behind may read the rest early
            c->kv_hint_node = next_live(c, cgraph, j[6] + 1);
            c->kv_hint_k = sk->data; c->kv_hint_v = sv->data; c->kv_hint_idx = sk->src[1];
        }
        return true;
    }
    if (nf) CHECK_RUN(c, b200_mul_mat_shared_src1_norm(c->ctx, 3, d, w, (const float *) nf->x0->data, &nf->n, o));
    else    CHECK_RUN(c, b200_mul_mat_shared_src1(c->ctx, 3, d, w, (const float *) mq->src[1]->data, o));
    if (j_fa >= 0) {
        CHECK_RUN(c, b200_attn_decode(c->ctx, &q, &qd, &k, &kd, &pos, rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc, &fq, &fk, &fv,
                                      fa->src[3] ? &fm : nullptr, fa->src[4] ? &fs : nullptr, &fd, ggml_get_op_params_f32(fa, 0),
                                      ggml_get_op_params_f32(fa, 1), ggml_get_op_params_f32(fa, 2)));
        c->done[j_fa] = 1;
    } else {
        CHECK_RUN(c, b200_attn_prep(c->ctx, &q, &qd, &k, &kd, &pos, rq->src[2] ? &ff : nullptr, &rp, &ki, &kc, &v, &vi, &vc));
    }
    for (int k2 = 1; k2 < 7; k2++) c->done[j[k2]] = 1;
    return true;
}

ggml_status run_nodes(be_ctx * c, ggml_cgraph * cgraph) {
    c->done.assign((size_t) cgraph->n_nodes, 0);
    c->kv_hint_node = -1;
    for (int i = 0; i < cgraph->n_nodes; i++) {
        ggml_tensor * node = cgraph->nodes[i];
        if (is_noop(node) || c->done[i]) continue;  // as ggml-cuda.cu:3333-3335; done: ran ahead of its position in a fused launch
        if (c->fuse_epilogue && (node->op == GGML_OP_MUL_MAT || node->op == GGML_OP_MUL_MAT_ID)) {
            const int fused = try_fuse_epilogue(c, cgraph, i);
            if (fused > 0) {
                g_fused_epilogues.fetch_add(1, std::memory_order_relaxed);
                i += fused - 1;
                continue;
            }
        }
        if (node->op == GGML_OP_MUL_MAT && try_fuse_attn_block(c, cgraph, i)) continue;
        if (node->op == GGML_OP_MUL_MAT) {
            b200_mul_mat_desc descs[4];
            const void *      w[4];
            float *           o[4];
            if (!fill_mul_mat(node, &descs[0])) {
                GGML_LOG_ERROR("ggml-b200: unsupported MUL_MAT reached graph_compute (%s)\n", node->name);
                return GGML_STATUS_FAILED;
            }
            if (node->src[0]->buffer && buft_is_split(node->src[0]->buffer->buft)) {
                const ggml_status st = split_mul_mat(c, node, descs[0]);
                if (st != GGML_STATUS_SUCCESS) return st;
                continue;
            }
            w[0] = node->src[0]->data;
            o[0] = (float *) node->data;
            int n = 1;
            // peephole: MUL_MATs reading the same src1 (q/k/v, gate/up) share one activation quantization and one launch; the
            // reference re-quantizes per op (mmvq.cu:667-674). llama's graphs put ROPE(q) and reshapes between the three attention
            // projections (src/llama-graph.cpp:1593-1595), so the followers are looked for in a window and run AHEAD of their position
            // - allowed when nothing that stays in between reads or writes what a hoisted MUL_MAT writes, or writes src1
            int idx[4] = { i, -1, -1, -1 };
            const ggml_tensor * between[16];
            int n_between = 0;
            const int window = c->fuse_f3 ? 12 : 1;
            for (int j = i + 1; c->fuse && n < 4 && j < cgraph->n_nodes && j <= i + window && n_between < 16; j++) {
                ggml_tensor * nx = cgraph->nodes[j];
                if (c->done[j] || is_noop(nx)) continue;
                bool take = nx->op == GGML_OP_MUL_MAT && nx->src[1] == node->src[1] && nx->src[0]->op == GGML_OP_NONE && fill_mul_mat(nx, &descs[n]) &&
                            !(nx->src[0]->buffer && buft_is_split(nx->src[0]->buffer->buft)) && descs[n].type == descs[0].type &&
                            descs[n].ne00 == descs[0].ne00 && descs[0].ne12 == 1 && descs[0].ne13 == 1;
                if (take) {
                    bool clash = ranges_overlap(nx, node->src[1]);
                    for (int k = 0; k < n; k++) clash = clash || ranges_overlap(nx, cgraph->nodes[idx[k]]) || nx->src[0] == cgraph->nodes[idx[k]];
                    for (int b = 0; b < n_between && !clash; b++) {
                        clash = ranges_overlap(nx, between[b]);
                        for (int k = 0; k < GGML_MAX_SRC && !clash; k++) clash = between[b]->src[k] && ranges_overlap(nx, between[b]->src[k]);
                    }
                    take = !clash;
                }
                if (take) {
                    w[n] = nx->src[0]->data;
                    o[n] = (float *) nx->data;
                    idx[n++] = j;
                    continue;
                }
                if (ranges_overlap(nx, node->src[1])) break;   // src1 changes here: nothing behind it may be hoisted
                between[n_between++] = nx;
            }
            if (n == 1) CHECK_RUN(c, b200_mul_mat(c->ctx, &descs[0], w[0], (const float *) node->src[1]->data, o[0]));
            else CHECK_RUN(c, b200_mul_mat_shared_src1(c->ctx, n, descs, w, (const float *) node->src[1]->data, o));
            for (int k = 1; k < n; k++) c->done[idx[k]] = 1;
        } else if (node->op == GGML_OP_MUL_MAT_ID) {
            b200_mul_mat_id_desc d;
            if (!fill_mul_mat_id(node, &d)) {
                GGML_LOG_ERROR("ggml-b200: unsupported MUL_MAT_ID reached graph_compute (%s)\n", node->name);
                return GGML_STATUS_FAILED;
            }
            CHECK_RUN(c, b200_mul_mat_id(c->ctx, &d, node->src[0]->data, (const float *) node->src[1]->data, (const int32_t *) node->src[2]->data,
                                  (float *) node->data));
        } else if (node->op == GGML_OP_GLU && glu_ok(node)) {
            run_glu(c, node);
        } else if ((node->op == GGML_OP_ADD || node->op == GGML_OP_MUL) && add_ok(node)) {
            run_add(c, node);
        } else if (node->op == GGML_OP_ADD_ID && add_id_ok(node)) {
            run_add_id(c, node);
        } else if (is_f3_op(node->op) && f3_ok(node)) {
            if (node->op == GGML_OP_RMS_NORM) {
                if (try_fuse_norm_mul(c, cgraph, i)) continue;
            } else if (node->op == GGML_OP_ROPE && try_fuse_attn_prep(c, cgraph, i)) {
                continue;
            }
            const ggml_tensor * hint = nullptr;
            if (node->op == GGML_OP_FLASH_ATTN_EXT && c->kv_hint_node == i && node->src[1]->data == c->kv_hint_k && node->src[2]->data == c->kv_hint_v &&
                node->src[0]->ne[1] == 1) hint = c->kv_hint_idx;
            run_f3(c, node, hint);
        } else {
            GGML_LOG_ERROR("ggml-b200: op %s is not implemented by this backend (%s)\n", ggml_op_name(node->op), node->name);
            return GGML_STATUS_FAILED;
        }
    }
    return GGML_STATUS_SUCCESS;
}

// ---- CUDA-graph replay. The key of a cgraph is everything a launch depends on: per node the op, its parameters, shape, strides
// and address, and shape, strides and address of every source (ggml_graph_node_properties, ggml-cuda.cu:2870-2935 does the same).
// Weight CONTENTS are not part of it: kernels read them through the same pointers on every replay.
// (runs once per graph_compute, i.e. once per token: raw stores into a buffer that never shrinks, ~50 words per node)
constexpr size_t KEY_WORDS_PER_NODE = 1 + 10 + GGML_MAX_OP_PARAMS / sizeof(uint64_t) + GGML_MAX_SRC * 10;
size_t graph_key(const ggml_cgraph * cgraph, std::vector<uint64_t> & buf, int * n_run, bool * capturable) {
    *n_run = 0;
    *capturable = true;
    const size_t need = (size_t) cgraph->n_nodes * KEY_WORDS_PER_NODE;
    if (buf.size() < need) buf.resize(need);
    uint64_t * o = buf.data();
    auto put_tensor = [&](const ggml_tensor * t) {
        *o++ = (uint64_t) (uintptr_t) t->data;
        *o++ = (uint64_t) t->type;
        memcpy(o, t->ne, sizeof(t->ne)); o += GGML_MAX_DIMS;
        memcpy(o, t->nb, sizeof(t->nb)); o += GGML_MAX_DIMS;
    };
    static_assert(sizeof(((ggml_tensor *) nullptr)->ne) == GGML_MAX_DIMS * 8 && sizeof(((ggml_tensor *) nullptr)->nb) == GGML_MAX_DIMS * 8, "ne / nb are 64-bit");
    for (int i = 0; i < cgraph->n_nodes; i++) {
        const ggml_tensor * n = cgraph->nodes[i];
        if (is_noop(n)) continue;
        (*n_run)++;
        *o++ = (uint64_t) n->op | ((uint64_t) n->flags << 32);
        put_tensor(n);
        memcpy(o, n->op_params, GGML_MAX_OP_PARAMS); o += GGML_MAX_OP_PARAMS / sizeof(uint64_t);
        for (int k = 0; k < GGML_MAX_SRC; k++) {
            const ggml_tensor * s = n->src[k];
            if (!s) { *o++ = ~(uint64_t) 0; continue; }
            put_tensor(s);
            if (s->buffer && buft_is_split(s->buffer->buft)) *capturable = false;  // several streams and devices: launched directly
        }
    }
    return (size_t) (o - buf.data());
}

uint64_t hash_key(const uint64_t * key, size_t n) {   // four independent lanes: the multiply chains overlap
    uint64_t h[4] = { 0xcbf29ce484222325ull, 0x9e3779b97f4a7c15ull, 0xc2b2ae3d27d4eb4full, 0x165667b19e3779f9ull };
    size_t i = 0;
    for (; i + 4 <= n; i += 4) {
        for (int l = 0; l < 4; l++) { h[l] ^= key[i + l]; h[l] *= 0x100000001b3ull; h[l] ^= h[l] >> 29; }
    }
    for (; i < n; i++) { h[0] ^= key[i]; h[0] *= 0x100000001b3ull; h[0] ^= h[0] >> 29; }
    return (h[0] * 31 + h[1]) * 31 + (h[2] * 31 + h[3]) + n;
}

void evict_graphs(be_ctx * c) {  // bounded cache: drop the least recently used half
    std::vector<uint64_t> stamps;
    for (auto & kv : c->graphs) for (auto & g : kv.second) stamps.push_back(g.stamp);
    std::sort(stamps.begin(), stamps.end());
    const uint64_t cut = stamps[stamps.size() / 2];
    b200_ctx_synchronize(c->ctx);
    for (auto it = c->graphs.begin(); it != c->graphs.end();) {
        auto & v = it->second;
        for (size_t i = 0; i < v.size();) {
            if (v[i].stamp < cut) { b200_graph_destroy(v[i].exec); v.erase(v.begin() + (long) i); c->n_graphs--; }
            else i++;
        }
        it = v.empty() ? c->graphs.erase(it) : std::next(it);
    }
}

std::atomic<uint64_t> g_graph_replays{ 0 };  // proc address "ggml_backend_b200_graph_replays" (tests, bench.py)
uint64_t graph_replays() { return g_graph_replays.load(std::memory_order_relaxed); }

ggml_status be_graph_compute_impl(ggml_backend_t backend, ggml_cgraph * cgraph, uint64_t * key_ns);

uint64_t now_ns() { return (uint64_t) std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// GGML_B200_TIMING=1: host time spent keying / launching and device time between two events around the launches of every
// graph_compute (read back at the next call, when the caller has long synchronized)
void timing_collect(be_ctx * c) {
    if (!c->t_pending) return;
    float ms = 0;
    be_ctx::tstat & ts = c->t_pending_replay ? c->t_replay : c->t_direct;
    if (b200_event_synchronize(c->t_ev1) == B200_OK && b200_event_elapsed_ms(c->t_ev0, c->t_ev1, &ms) == B200_OK) { ts.gpu_ms += ms; ts.gpu_n++; }
    c->t_pending = false;
}

ggml_status be_graph_compute(ggml_backend_t backend, ggml_cgraph * cgraph) {
    be_ctx * c = (be_ctx *) backend->context;
    if (!c->timing) return be_graph_compute_impl(backend, cgraph, nullptr);
    timing_collect(c);
    if (!c->t_ev0) { CHECK(b200_event_create(c->device, &c->t_ev0)); CHECK(b200_event_create(c->device, &c->t_ev1)); }
    const uint64_t t0 = now_ns();
    uint64_t key_ns = 0;
    CHECK(b200_event_record(c->ctx, c->t_ev0));
    const uint64_t replays0 = c->graph_replays;
    const ggml_status st = be_graph_compute_impl(backend, cgraph, &key_ns);
    CHECK(b200_event_record(c->ctx, c->t_ev1));
    c->t_pending = true;
    c->t_pending_replay = c->graph_replays != replays0;
    be_ctx::tstat & ts = c->t_pending_replay ? c->t_replay : c->t_direct;
    ts.n++;
    ts.key_ns += key_ns;
    ts.run_ns += now_ns() - t0 - key_ns;
    return st;
}

ggml_status be_graph_compute_impl(ggml_backend_t backend, ggml_cgraph * cgraph, uint64_t * key_ns) {
    be_ctx * c = (be_ctx *) backend->context;
    if (!c->use_graphs) return run_nodes(c, cgraph);
    static thread_local std::vector<uint64_t> key;
    int  n_run = 0;
    bool capturable = true;
    const uint64_t tk0 = key_ns ? now_ns() : 0;
    const size_t n_key = graph_key(cgraph, key, &n_run, &capturable);
    if (!capturable || n_run < c->graph_min_nodes) return run_nodes(c, cgraph);
    auto same = [&](const be_ctx::cached_graph & e) { return e.key.size() == n_key && memcmp(e.key.data(), key.data(), n_key * sizeof(uint64_t)) == 0; };
    // decode repeats one cgraph token after token: the bucket that answered last time is compared first, and only a miss pays for the hash
    be_ctx::cached_graph * g = nullptr;
    if (c->last_hash_valid) {
        auto it = c->graphs.find(c->last_hash);
        if (it != c->graphs.end()) for (auto & e : it->second) if (same(e)) { g = &e; break; }
    }
    if (!g) {
        const uint64_t h = hash_key(key.data(), n_key);
        auto & bucket = c->graphs[h];
        for (auto & e : bucket) if (same(e)) { g = &e; break; }
        if (!g) {
            if (c->n_graphs >= 1024) {
                evict_graphs(c);
                return be_graph_compute_impl(backend, cgraph, key_ns);  // `bucket` may be gone
            }
            bucket.emplace_back();
            g = &bucket.back();
            g->key.assign(key.begin(), key.begin() + (long) n_key);
            c->n_graphs++;
        }
        c->last_hash = h;
        c->last_hash_valid = true;
    }
    if (key_ns) *key_ns = now_ns() - tk0;
    g->stamp = ++c->graph_clock;
    g->seen++;
    if (g->exec) {
        CHECK(b200_graph_launch(c->ctx, g->exec));
        g_graph_replays.fetch_add(1, std::memory_order_relaxed);
        c->graph_replays++;
        return GGML_STATUS_SUCCESS;
    }
    // first appearance: launched directly (this also sizes the scratch buffers and fills the weight-expansion cache, neither of
    // which can happen inside a capture); second appearance: captured, then launched
    if (g->seen < 2 || g->bad) return run_nodes(c, cgraph);
    if (b200_graph_capture_begin(c->ctx) != B200_OK) { g->bad = true; return run_nodes(c, cgraph); }
    c->capturing = true;
    ggml_status st = GGML_STATUS_SUCCESS;
    bool ok = true;
    try {
        st = run_nodes(c, cgraph);
    } catch (const capture_failed &) {
        ok = false;
    }
    c->capturing = false;
    b200_graph * exec = nullptr;
    if (b200_graph_capture_end(c->ctx, &exec) != B200_OK || !ok || st != GGML_STATUS_SUCCESS) {
        if (exec) b200_graph_destroy(exec);
        g->bad = true;
        GGML_LOG_DEBUG("ggml-b200: CUDA graph capture of a %d-node cgraph failed (%s); launching directly\n", n_run, b200_last_error());
        return run_nodes(c, cgraph);
    }
    g->exec = exec;
    CHECK(b200_graph_launch(c->ctx, g->exec));
    g_graph_replays.fetch_add(1, std::memory_order_relaxed);
    return GGML_STATUS_SUCCESS;
}

void be_event_record(ggml_backend_t backend, ggml_backend_event_t event) { CHECK(b200_event_record(((be_ctx *) backend->context)->ctx, event->context)); }
void be_event_wait(ggml_backend_t backend, ggml_backend_event_t event) { CHECK(b200_event_wait(((be_ctx *) backend->context)->ctx, event->context)); }

const ggml_backend_i be_iface = {
    /* .get_name           = */ be_name,
    /* .free               = */ be_free,
    /* .set_tensor_async   = */ be_set_tensor_async,
    /* .get_tensor_async   = */ be_get_tensor_async,
    /* .cpy_tensor_async   = */ be_cpy_tensor_async,
    /* .synchronize        = */ be_synchronize,
    /* .graph_plan_create  = */ nullptr,
    /* .graph_plan_free    = */ nullptr,
    /* .graph_plan_update  = */ nullptr,
    /* .graph_plan_compute = */ nullptr,
    /* .graph_compute      = */ be_graph_compute,
    /* .event_record       = */ be_event_record,
    /* .event_wait         = */ be_event_wait,
    /* .graph_optimize     = */ nullptr,
};

ggml_guid_t be_guid() {
    static ggml_guid guid = { 0xb2, 0x00, 0x51, 0x00, 0xa0, 0x67, 0x67, 0x6d, 0x6c, 0x2d, 0x62, 0x32, 0x30, 0x30, 0x00, 0x01 };
    return &guid;
}

// ------------------------------------------------------------------------------------------------ device

const char * dev_name(ggml_backend_dev_t dev) { return ((dev_ctx *) dev->context)->name.c_str(); }
const char * dev_description(ggml_backend_dev_t dev) { return ((dev_ctx *) dev->context)->description.c_str(); }
void dev_memory(ggml_backend_dev_t dev, size_t * free, size_t * total) {
    if (b200_device_memory(((dev_ctx *) dev->context)->device, free, total) != B200_OK) { *free = 0; *total = 0; }
}
enum ggml_backend_dev_type dev_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }
void dev_props(ggml_backend_dev_t dev, ggml_backend_dev_props * props) {
    dev_ctx * d = (dev_ctx *) dev->context;
    props->name        = d->name.c_str();
    props->description = d->description.c_str();
    props->type        = GGML_BACKEND_DEVICE_TYPE_GPU;
    props->device_id   = d->pci_id.c_str();
    dev_memory(dev, &props->memory_free, &props->memory_total);
    props->caps = { /* async */ true, /* host_buffer */ true, /* buffer_from_host_ptr */ false, /* events */ true };
}
ggml_backend_t dev_init_backend(ggml_backend_dev_t dev, const char *) {
    dev_ctx * d = (dev_ctx *) dev->context;
    b200_ctx * ctx = nullptr;
    if (b200_ctx_create(d->device, nullptr, &ctx) != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: cannot create a context on device %d: %s\n", d->device, b200_last_error());
        return nullptr;
    }
    // this backend owns its buffers and reports every write to them (drop_cached), so the int8 expansion of a weight
    // matrix used by the tcgen05 prefill kernel is built once and kept (GGML_B200_WEIGHT_CACHE=0 rebuilds it per call)
    const char * wc = getenv("GGML_B200_WEIGHT_CACHE");
    const bool weight_cache = !(wc && !atoi(wc));
    b200_ctx_set_weight_cache(ctx, weight_cache);
    // every src0 this backend sees lives in a buffer it allocated with b200_malloc and frees with b200_free (which drops the
    // learned next-kernel predictions), so the decode weight prefetch is safe to leave on
    b200_ctx_set_prefetch(ctx, 64);
    const char * nf = getenv("GGML_B200_DISABLE_FUSION");
    be_ctx * c = new be_ctx{ d->device, ctx, d->name, !(nf && atoi(nf)) };
    const char * ne = getenv("GGML_B200_DISABLE_EPILOGUE_FUSION");
    c->fuse_epilogue = c->fuse && !(ne && atoi(ne));
    c->weight_cache  = weight_cache;
    if (const char * tm = getenv("GGML_B200_TIMING")) c->timing = atoi(tm) != 0;
    if (const char * nf3 = getenv("GGML_B200_DISABLE_F3_FUSION")) c->fuse_f3 = atoi(nf3) == 0;
    const char * ng = getenv("GGML_B200_DISABLE_GRAPHS");
    c->use_graphs = !(ng && atoi(ng));
    if (const char * mn = getenv("GGML_B200_GRAPH_MIN_NODES")) c->graph_min_nodes = std::max(1, atoi(mn));
    return new ggml_backend{ be_guid(), be_iface, dev, c };
}
ggml_backend_buffer_type_t dev_buffer_type(ggml_backend_dev_t dev) { return &((dev_ctx *) dev->context)->buft; }
ggml_backend_buffer_type_t dev_host_buffer_type(ggml_backend_dev_t dev) { return &((dev_ctx *) dev->context)->host_buft; }

bool dev_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    return (buft_is_b200(buft) && buft->device == dev) || buft_is_split(buft);  // as ggml-cuda.cu:4587-4590
}

// GGML_B200_EXACT=1: validation mode - the MUL_MAT(_ID) kernels reproduce ggml-cpu's generic fp32 operation order bit for bit
// (b200_ctx_set_exact), and the backend declines GLU (its libm-accurate expf is not ggml-cpu's) so that everything around the
// matmuls is computed by the same ggml-cpu code as in a CPU-only run
bool exact_mode() {
    static const bool on = [] { const char * s = getenv("GGML_B200_EXACT"); return s && atoi(s) != 0; }();
    return on;
}

bool dev_supports_op(ggml_backend_dev_t dev, const ggml_tensor * op) {
    // answered for hypothetical tensors too: llama probes with 0-byte buffers of the candidate buft (src/llama-model.cpp:165-305)
    // and test-backend-ops asks for leaves and views (:1265-1272). Sources in foreign buffers are the scheduler's business
    // (it copies them in: offload_op, ggml-backend.cpp:750-772).
    switch (op->op) {
        case GGML_OP_NONE:
        case GGML_OP_RESHAPE:
        case GGML_OP_VIEW:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT: {
            b200_mul_mat_desc d;
            if (!fill_mul_mat(op, &d)) return false;
            const ggml_tensor * a = op->src[0];
            if (a->buffer && buft_is_split(a->buffer->buft)) {  // row slices: plain 2-D matrices only, computed on the main device's backend
                const split_buft_ctx * bc = (const split_buft_ctx *) a->buffer->buft->context;
                return d.ne02 == 1 && d.ne03 == 1 && d.ne12 == 1 && d.ne13 == 1 && bc->main_device == ((dev_ctx *) dev->context)->device;
            }
            return true;
        }
        case GGML_OP_MUL_MAT_ID: {
            b200_mul_mat_id_desc d;
            if (op->src[0] && op->src[0]->buffer && buft_is_split(op->src[0]->buffer->buft)) return false;  // as ggml-cuda.cu:4261-4268
            return fill_mul_mat_id(op, &d);
        }
        // K7: the nodes the fused mat-vec epilogue absorbs; owned stand-alone as well so that the scheduler keeps
        // {MUL_MAT, ADD, GLU} chains on this backend (the reference: ggml-cuda.cu:4331-4345 GLU, :4509-4516 ADD / ADD_ID)
        case GGML_OP_GLU:
            return !exact_mode() && glu_ok(op);
        case GGML_OP_ADD:
        case GGML_OP_MUL:  // same broadcasting kernel; the expert-weight scaling behind a fused MoE FFN
            return add_ok(op);
        case GGML_OP_ADD_ID:
            return add_id_ok(op);
        // f3: ggml-cuda.cu:4346-4618 answers for the same ops
        case GGML_OP_RMS_NORM:
        case GGML_OP_GET_ROWS:
        case GGML_OP_ROPE:
        case GGML_OP_SET_ROWS:
        case GGML_OP_CPY:
        case GGML_OP_CONT:
        case GGML_OP_DUP:
        case GGML_OP_FLASH_ATTN_EXT:
            return f3_ok(op);
        default:
            return false;
    }
}

bool dev_offload_op(ggml_backend_dev_t, const ggml_tensor * op) {
    const int min_batch = 32;  // ggml-cuda.cu:4641-4647
    return (op->op == GGML_OP_MUL_MAT && op->ne[1] >= min_batch) || (op->op == GGML_OP_MUL_MAT_ID && op->ne[2] >= min_batch);
}

ggml_backend_event_t dev_event_new(ggml_backend_dev_t dev) {
    void * ev = nullptr;
    if (b200_event_create(((dev_ctx *) dev->context)->device, &ev) != B200_OK) return nullptr;
    return new ggml_backend_event{ dev, ev };
}
void dev_event_free(ggml_backend_dev_t dev, ggml_backend_event_t event) {
    b200_event_destroy(((dev_ctx *) dev->context)->device, event->context);
    delete event;
}
void dev_event_synchronize(ggml_backend_dev_t, ggml_backend_event_t event) { CHECK(b200_event_synchronize(event->context)); }

const ggml_backend_device_i dev_iface = {
    /* .get_name             = */ dev_name,
    /* .get_description      = */ dev_description,
    /* .get_memory           = */ dev_memory,
    /* .get_type             = */ dev_type,
    /* .get_props            = */ dev_props,
    /* .init_backend         = */ dev_init_backend,
    /* .get_buffer_type      = */ dev_buffer_type,
    /* .get_host_buffer_type = */ dev_host_buffer_type,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ dev_supports_op,
    /* .supports_buft        = */ dev_supports_buft,
    /* .offload_op           = */ dev_offload_op,
    /* .event_new            = */ dev_event_new,
    /* .event_free           = */ dev_event_free,
    /* .event_synchronize    = */ dev_event_synchronize,
};

// ------------------------------------------------------------------------------------------------ reg

struct reg_ctx {
    std::vector<ggml_backend_device *> devices;
};

const char * reg_name(ggml_backend_reg_t) { return B200_NAME; }
size_t reg_dev_count(ggml_backend_reg_t reg) { return ((reg_ctx *) reg->context)->devices.size(); }
ggml_backend_dev_t reg_dev_get(ggml_backend_reg_t reg, size_t i) {
    reg_ctx * r = (reg_ctx *) reg->context;
    GGML_ASSERT(i < r->devices.size());
    return r->devices[i];
}

ggml_backend_feature * reg_features(ggml_backend_reg_t) {
    static ggml_backend_feature f[] = {
        { "ARCH", "sm_100a" }, { "TMA_BULK_MMVQ", "1" }, { "TCGEN05_I8_MMQ", "1" }, { "ACT_QUANT", "bit-exact q8_0/q8_1/q8_K" },
        { "CPU_FALLBACK", "0" }, { nullptr, nullptr },
    };
    return f;
}

// ggml_backend_split_buffer_type_t (ggml-backend.h:199-200): singletons per (main device, normalised split), mutex guarded
// because llama calls it from loader threads (ggml-cuda.cu:1062-1099)
ggml_backend_buffer_type_t split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    static std::map<std::pair<int, std::array<float, B200_MAX_DEVICES>>, ggml_backend_buffer_type> bufts;
    const int n = split_dev_count();
    if (main_device < 0 || main_device >= n) return nullptr;
    std::array<float, B200_MAX_DEVICES> ts = {};
    bool all_zero = tensor_split == nullptr;
    if (!all_zero) {
        all_zero = true;
        for (int i = 0; i < n; i++) all_zero = all_zero && tensor_split[i] == 0.0f;
    }
    if (all_zero) {  // default: equal shares (the reference weighs by device memory, ggml-cuda.cu:277-285; these are identical parts)
        for (int i = 0; i < n; i++) ts[i] = (float) i / (float) n;
    } else {         // cumulative sum of the requested proportions, normalised (:1070-1080)
        float sum = 0.0f;
        for (int i = 0; i < n; i++) { ts[i] = sum; sum += tensor_split[i]; }
        for (int i = 0; i < n; i++) ts[i] /= sum;
    }
    auto it = bufts.find({ main_device, ts });
    if (it != bufts.end()) return &it->second;
    split_buft_ctx * c = new split_buft_ctx{ main_device, ts, std::string(B200_NAME) + std::to_string(main_device) + "_Split" };
    ggml_backend_buffer_type buft = { split_buft_iface, ggml_backend_reg_dev_get(b200_reg(), (size_t) main_device), c };
    return &bufts.emplace(std::make_pair(main_device, ts), buft).first->second;
}

void * reg_proc_address(ggml_backend_reg_t, const char * name) {
    if (!strcmp(name, "ggml_backend_get_features")) return (void *) reg_features;
    if (!strcmp(name, "ggml_backend_split_buffer_type")) return (void *) split_buffer_type;
    if (!strcmp(name, "ggml_backend_b200_fused_epilogues")) return (void *) fused_epilogues;  // uint64_t (*)(void): K7 launches so far
    if (!strcmp(name, "ggml_backend_b200_graph_replays")) return (void *) graph_replays;      // uint64_t (*)(void): CUDA-graph launches so far
    // "ggml_backend_set_n_threads", "ggml_backend_dev_get_extra_bufts": not provided
    return nullptr;
}

const ggml_backend_reg_i reg_iface = { reg_name, reg_dev_count, reg_dev_get, reg_proc_address };

ggml_backend_reg_t b200_reg() {
    static ggml_backend_reg reg;
    static std::once_flag   once;
    std::call_once(once, [] {
        reg_ctx * r = new reg_ctx();
        reg = { GGML_BACKEND_API_VERSION, reg_iface, r };
        const int n = b200_device_count();
        for (int i = 0; i < n && i < B200_MAX_DEVICES; i++) {
            char name[128] = "", pci[32] = "";
            int  sms = 0, maj = 0, min = 0;
            if (b200_device_info(i, name, sizeof(name), pci, sizeof(pci), &sms, &maj, &min) != B200_OK) break;
            dev_ctx * d    = new dev_ctx();
            d->device      = i;
            d->name        = std::string(B200_NAME) + std::to_string(i);
            d->description = std::string(name) + " (sm_" + std::to_string(maj) + std::to_string(min) + "a, " + std::to_string(sms) + " SMs)";
            d->pci_id      = pci;
            ggml_backend_device * dev = new ggml_backend_device{ dev_iface, &reg, d };
            d->buft      = { buft_iface, dev, d };
            d->host_buft = { host_buft_iface, dev, d };
            r->devices.push_back(dev);
        }
        if (n == 0) GGML_LOG_WARN("ggml-b200: no sm_100 device found; the backend registers zero devices (there is no CPU fallback)\n");
    });
    return &reg;
}

} // namespace

extern "C" {
GGML_BACKEND_API ggml_backend_reg_t ggml_backend_init(void) { return b200_reg(); }
GGML_BACKEND_API int ggml_backend_score(void) { return b200_device_count() > 0 ? 100 : 0; }
// in-process registration for hosts that link the library directly (ggml_backend_register, ggml-backend.h:218)
GGML_BACKEND_API ggml_backend_reg_t ggml_backend_b200_reg(void) { return b200_reg(); }
}
