// api.cu — the C-ABI of include/ggml_b200.h: device/context/memory plumbing and the MUL_MAT dispatch that replaces
// ggml_cuda_mul_mat (/root/reference/ggml/src/ggml-cuda/ggml-cuda.cu:2178-2261).
// There is no CPU path in this file: every entry point either launches sm_100a kernels or fails.

#include "internal.h"

#include <atomic>

#include <mutex>
#include <vector>

using namespace b200;

// ------------------------------------------------------------------------------------------------ errors

static thread_local char g_err[512] = "";

b200_status b200_fail(b200_status st, const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return st;
}
#define fail b200_fail

const char * b200_last_error(void) { return g_err; }

b200_status b200_ensure(b200_ctx * ctx, void ** buf, size_t * cur, size_t need) {
    if (need <= *cur) return B200_OK;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(ctx->stream, &cs);
    if (cs != cudaStreamCaptureStatusNone) {
        return fail(B200_ERR_ARG, "scratch buffer would have to grow (%zu -> %zu) during stream capture; run once uncaptured first", *cur, need);
    }
    CU(cudaStreamSynchronize(ctx->stream));
    if (*buf) CU(cudaFree(*buf));
    *buf = nullptr;
    *cur = 0;
    const size_t sz = (need + (size_t) (1 << 20) - 1) & ~((size_t) (1 << 20) - 1);
    CU(cudaMalloc(buf, sz));
    *cur = sz;
    return B200_OK;
}
#define ensure b200_ensure

int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int ok = 0;
    for (int i = 0; i < n; i++) {
        int maj = 0;
        if (cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && maj == 10) ok++;
        else return ok;  // devices must be a homogeneous prefix; sm_100a images run on CC 10.0 only
    }
    return ok;
}

b200_status b200_device_info(int device, char * name, size_t name_len, char * pci, size_t pci_len, int * sms, int * maj, int * min) {
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (name && name_len) snprintf(name, name_len, "%s", prop.name);
    if (pci && pci_len) snprintf(pci, pci_len, "%04x:%02x:%02x.0", prop.pciDomainID, prop.pciBusID, prop.pciDeviceID);
    if (sms) *sms = prop.multiProcessorCount;
    if (maj) *maj = prop.major;
    if (min) *min = prop.minor;
    return B200_OK;
}

b200_status b200_device_memory(int device, size_t * free_b, size_t * total_b) {
    CU(cudaSetDevice(device));
    CU(cudaMemGetInfo(free_b, total_b));
    return B200_OK;
}

b200_status b200_ctx_create(int device, void * stream, b200_ctx ** out) {
    if (!out) return fail(B200_ERR_ARG, "out is NULL");
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(B200_ERR_CUDA, "no CUDA device: the B200 backend has no CPU fallback");
    }
    if (device < 0 || device >= n) return fail(B200_ERR_ARG, "device %d out of range (%d devices)", device, n);
    int maj = 0, sms = 0;
    CU(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, device));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (maj != 10) return fail(B200_ERR_CUDA, "device %d has compute capability %d.x; kernels are built for sm_100a only", device, maj);
    CU(cudaSetDevice(device));
    b200_ctx * c = new b200_ctx();
    c->device   = device;
    c->sm_count = sms;
    if (stream) {
        c->stream = (cudaStream_t) stream;
    } else {
        cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) {
            delete c;
            return fail(B200_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
        }
        c->own_stream = true;
    }
    if (const char * s = getenv("GGML_B200_FORCE")) {
        if (!strcmp(s, "mmvq")) c->force_path = B200_PATH_MMVQ;
        else if (!strcmp(s, "mmq")) c->force_path = B200_PATH_MMQ;
        else if (!strcmp(s, "simple")) c->force_path = B200_PATH_SIMPLE;
    }
    if (const char * s = getenv("GGML_B200_EXACT")) {
        if (atoi(s)) { c->exact = 1; c->force_path = B200_PATH_SIMPLE; }
    }
    if (const char * s = getenv("GGML_B200_MV_WARPS")) c->mv_warps = atoi(s);
    if (const char * s = getenv("GGML_B200_MV_STAGES")) c->mv_stages = atoi(s);
    if (const char * s = getenv("GGML_B200_PDL")) c->pdl = atoi(s) != 0;
    if (const char * s = getenv("GGML_B200_MMQ_MIN")) c->mmq_min_cols = atoi(s);
    if (const char * s = getenv("GGML_B200_WEIGHT_CACHE")) c->weight_cache = atoi(s) != 0;
    if (const char * s = getenv("GGML_B200_FUSE_QUANT")) c->fuse_quant = atoi(s) != 0;
    if (const char * s = getenv("GGML_B200_PREFETCH_KB")) c->prefetch_kb = atoi(s);
    if (const char * s = getenv("GGML_B200_FA_PREFETCH_KB")) c->fa_prefetch_kb = atoi(s);
    if (const char * s = getenv("GGML_B200_DBG")) c->dbg = atoi(s);
    if (const char * s = getenv("GGML_B200_MMA_MIN")) c->mma_min_cols = atoi(s);
    if (const char * s = getenv("GGML_B200_MMA_WARPS")) c->mma_warps = atoi(s);
    if (const char * s = getenv("GGML_B200_TC_MIN")) c->tc_min_cols = atoi(s);
    if (const char * s = getenv("GGML_B200_TC_MAX")) c->tc_max_cols = atoi(s);
    if (const char * s = getenv("GGML_B200_MMA_FUSE_UNITS")) c->mma_fuse_units = atoi(s);
    if (const char * s = getenv("GGML_B200_MOE_GROUP_MIN")) c->moe_group_min = atoi(s);
    if (const char * s = getenv("GGML_B200_MOE_TILE_MIN")) c->moe_tile_min = atoi(s);
    if (const char * s = getenv("GGML_B200_MMQ_F16")) c->mmq_f16 = atoi(s) != 0;
    *out = c;
    return B200_OK;
}

void b200_ctx_destroy(b200_ctx * ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->scratch) cudaFree(ctx->scratch);
    if (ctx->io) cudaFree(ctx->io);
    if (ctx->scratch2) cudaFree(ctx->scratch2);
    if (ctx->scratch3) cudaFree(ctx->scratch3);
    if (ctx->fa_ws) cudaFree(ctx->fa_ws);
    if (ctx->trace) cudaFree(ctx->trace);
    if (ctx->tc_turn) cudaFree(ctx->tc_turn);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int          b200_ctx_device(const b200_ctx * ctx) { return ctx->device; }
void *       b200_ctx_stream(const b200_ctx * ctx) { return (void *) ctx->stream; }
uint64_t     b200_ctx_launch_count(const b200_ctx * ctx) { return ctx->launches; }
void         b200_ctx_set_force_path(b200_ctx * ctx, int path) { ctx->force_path = ctx->exact ? B200_PATH_SIMPLE : path; }
void         b200_ctx_set_exact(b200_ctx * ctx, int enable) {
    ctx->exact      = enable != 0;
    ctx->force_path = enable ? B200_PATH_SIMPLE : B200_PATH_AUTO;
}
void         b200_ctx_set_weight_cache(b200_ctx * ctx, int enable) { ctx->weight_cache = enable != 0; }
void         b200_ctx_set_mmq_f16(b200_ctx * ctx, int enable) { ctx->mmq_f16 = enable != 0; }
b200_status  b200_ctx_trace(b200_ctx * ctx, int n_slots) {
    CU(cudaSetDevice(ctx->device));
    if (ctx->trace) {
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaFree(ctx->trace));
        ctx->trace = nullptr;
        ctx->trace_slots = 0;
    }
    if (n_slots <= 0) return B200_OK;
    const size_t bytes = (size_t) n_slots * ctx->sm_count * 64;
    CU(cudaMalloc((void **) &ctx->trace, bytes));
    CU(cudaMemset(ctx->trace, 0, bytes));
    ctx->trace_slots = n_slots;
    return B200_OK;
}
b200_status  b200_ctx_trace_read(b200_ctx * ctx, void * host, size_t bytes) {
    if (!ctx->trace) return fail(B200_ERR_ARG, "trace is off");
    const size_t have = (size_t) ctx->trace_slots * ctx->sm_count * 64;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaMemcpy(host, ctx->trace, bytes < have ? bytes : have, cudaMemcpyDeviceToHost));
    return B200_OK;
}
void         b200_ctx_set_prefetch(b200_ctx * ctx, int kb_per_cta) {
    if (getenv("GGML_B200_PREFETCH_KB")) return;  // the environment knob wins (experiments)
    ctx->prefetch_kb = kb_per_cta < 0 ? 0 : kb_per_cta;
    ctx->next_of.clear();
    ctx->last_key = nullptr;
}
b200_status  b200_ctx_synchronize(b200_ctx * ctx) {
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ row split

b200_status b200_ctx_set_row_split(b200_ctx * ctx, const b200_row_split * rs) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    if (b200_status st = b200_split_flush(ctx)) return st;
    ctx->split = SplitParams{};
    ctx->ll_on = 0;
    if (!rs || rs->n_ranks <= 1) return B200_OK;
    if (rs->n_ranks > SPLIT_MAX_RANKS || rs->rank < 0 || rs->rank >= rs->n_ranks || !rs->sync_words[rs->rank])
        return fail(B200_ERR_ARG, "bad row-split description (n_ranks=%d rank=%d)", rs->n_ranks, rs->rank);
    ctx->split.n_ranks = rs->n_ranks;
    ctx->split.rank    = rs->rank;
    for (int r = 0; r < rs->n_ranks; r++) {
        if (!rs->sync_words[r]) return fail(B200_ERR_ARG, "row split: rank %d has no sync words", r);
        ctx->split.dst_delta[r] = rs->dst_delta[r];
        ctx->split.sync_peer[r] = (uint32_t *) rs->sync_words[r];
    }
    ctx->split.sync_local = (uint32_t *) rs->sync_words[rs->rank];
    return B200_OK;
}

b200_status b200_ctx_set_row_split_ll(b200_ctx * ctx, const b200_row_split_ll * ll) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    if (b200_status st = b200_split_flush(ctx)) return st;
    ctx->ll_on = 0;
    if (!ll) return B200_OK;
    const int n = ctx->split.n_ranks;
    if (n <= 1) return fail(B200_ERR_ARG, "b200_ctx_set_row_split_ll: call b200_ctx_set_row_split first");
    if (!ll->dst_base || ll->region_bytes == 0 || ll->region_bytes % 8 || (uintptr_t) ll->dst_base % 8 || ll->rounding < 1 || ll->tensor_split[0] != 0.0)
        return fail(B200_ERR_ARG, "b200_ctx_set_row_split_ll: bad region / split description");
    for (int r = 0; r < n; r++) {
        if (!ll->stage[r] || (uintptr_t) ll->stage[r] % 8) return fail(B200_ERR_ARG, "b200_ctx_set_row_split_ll: rank %d has no (aligned) staging area", r);
        if (r > 0 && ll->tensor_split[r] < ll->tensor_split[r - 1]) return fail(B200_ERR_ARG, "b200_ctx_set_row_split_ll: tensor_split must not decrease");
        ctx->split.ll_stage[r]  = (char *) ll->stage[r];
        ctx->ll_tensor_split[r] = ll->tensor_split[r];
    }
    ctx->split.ll_base = (const char *) ll->dst_base;
    ctx->split.ll_half = 2 * ll->region_bytes;
    ctx->ll_region     = ll->region_bytes;
    ctx->ll_rounding   = ll->rounding;
    ctx->ll_on         = 1;
    return B200_OK;
}

bool b200_split_ll_geom(const b200_ctx * ctx, int n_mats, const int64_t * Ms, float * const * dsts, const int64_t * dst_cs, int ncols, SplitGeom * g) {
    const SplitParams & sp = ctx->split;
    if (sp.n_ranks <= 1 || !ctx->ll_on || ctx->split_deferred || n_mats > 4) return false;
    for (int i = 0; i < n_mats; i++) {
        const int64_t Mf = dst_cs[i];
        if (Mf < Ms[i] || Mf > INT32_MAX) return false;
        for (int r = 0; r < sp.n_ranks; r++) {  // get_row_split (ggml-cuda.cu:786-799): cumulative fractions, rounded down, the last rank takes the rest
            int64_t lo = r == 0 ? 0 : (int64_t) ((double) Mf * ctx->ll_tensor_split[r]);
            lo -= lo % ctx->ll_rounding;
            g->lo[i][r] = (int) lo;
        }
        g->lo[i][sp.n_ranks] = (int) Mf;
        for (int r = sp.n_ranks + 1; r <= SPLIT_MAX_RANKS; r++) g->lo[i][r] = (int) Mf;
        const int lo = g->lo[i][sp.rank], hi = g->lo[i][sp.rank + 1];
        if (hi - lo != Ms[i]) return false;
        const char * first = (const char *) (dsts[i] - lo), * last = (const char *) (dsts[i] - lo + (int64_t) (ncols - 1) * dst_cs[i] + Mf);
        if (first < sp.ll_base || last > sp.ll_base + ctx->ll_region) return false;
    }
    return true;
}

__global__ void split_wait_kernel(const SplitParams sp) { split_wait(sp); }

b200_status b200_split_flush(b200_ctx * ctx) {
    if (!ctx->split_pending || ctx->split.n_ranks <= 1) return B200_OK;
    CU(cudaSetDevice(ctx->device));
    split_wait_kernel<<<1, 32, 0, ctx->stream>>>(ctx->split);
    CU(cudaGetLastError());
    ctx->launches++;
    ctx->split_pending = 0;
    return B200_OK;
}

b200_status b200_ctx_set_row_split_deferred(b200_ctx * ctx, int enable) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    b200_status st = b200_split_flush(ctx);
    ctx->split_deferred = enable != 0;
    return st;
}
b200_status b200_ctx_row_split_wait(b200_ctx * ctx) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    return b200_split_flush(ctx);
}

// ------------------------------------------------------------------------------------------------ memory

b200_status b200_malloc(int device, size_t size, void ** ptr) {
    *ptr = nullptr;
    CU(cudaSetDevice(device));
    cudaError_t e = cudaMalloc(ptr, size ? size : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *ptr = nullptr;
        return fail(B200_ERR_ALLOC, "cudaMalloc(%zu) on device %d: %s", size, device, cudaGetErrorString(e));
    }
    return B200_OK;
}
static std::atomic<uint64_t> g_mem_gen{ 1 };
uint64_t b200_mem_generation() { return g_mem_gen.load(std::memory_order_relaxed); }
void     b200_mem_generation_bump() { g_mem_gen.fetch_add(1, std::memory_order_relaxed); }

b200_status b200_free(int device, void * ptr) {
    b200_mem_generation_bump();  // learned next-kernel predictions may point into this allocation
    CU(cudaSetDevice(device));
    CU(cudaFree(ptr));
    return B200_OK;
}
b200_status b200_host_malloc(size_t size, void ** ptr) {
    *ptr = nullptr;
    cudaError_t e = cudaMallocHost(ptr, size ? size : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        *ptr = nullptr;
        return fail(B200_ERR_ALLOC, "cudaMallocHost(%zu): %s", size, cudaGetErrorString(e));
    }
    return B200_OK;
}
b200_status b200_host_free(void * ptr) {
    CU(cudaFreeHost(ptr));
    return B200_OK;
}
b200_status b200_memcpy_h2d(int device, void * dst, const void * src, size_t size) {
    CU(cudaSetDevice(device));
    CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyHostToDevice, cudaStreamPerThread));
    CU(cudaStreamSynchronize(cudaStreamPerThread));
    return B200_OK;
}
b200_status b200_memcpy_d2h(int device, void * dst, const void * src, size_t size) {
    CU(cudaSetDevice(device));
    CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyDeviceToHost, cudaStreamPerThread));
    CU(cudaStreamSynchronize(cudaStreamPerThread));
    return B200_OK;
}
b200_status b200_memcpy_d2d(int device, void * dst, const void * src, size_t size) {
    CU(cudaSetDevice(device));
    CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    CU(cudaStreamSynchronize(cudaStreamPerThread));
    return B200_OK;
}
b200_status b200_memcpy_peer(int dst_device, void * dst, int src_device, const void * src, size_t size) {
    CU(cudaMemcpyPeer(dst, dst_device, src, src_device, size));
    return B200_OK;
}
b200_status b200_memset(int device, void * dst, int value, size_t size) {
    CU(cudaSetDevice(device));
    CU(cudaMemsetAsync(dst, value, size, cudaStreamPerThread));
    CU(cudaStreamSynchronize(cudaStreamPerThread));
    return B200_OK;
}
b200_status b200_memcpy_h2d_async(b200_ctx * ctx, void * dst, const void * src, size_t size) {
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyHostToDevice, ctx->stream));
    return B200_OK;
}
b200_status b200_memcpy_d2h_async(b200_ctx * ctx, void * dst, const void * src, size_t size) {
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyDeviceToHost, ctx->stream));
    return B200_OK;
}
b200_status b200_memcpy_d2d_async(b200_ctx * ctx, void * dst, const void * src, size_t size) {
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyDeviceToDevice, ctx->stream));
    return B200_OK;
}

b200_status b200_memcpy2d_async(b200_ctx * ctx, void * dst, size_t dpitch, const void * src, size_t spitch, size_t width, size_t height) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    if (width == 0 || height == 0) return B200_OK;
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, cudaMemcpyDefault, ctx->stream));  // unified addressing names the devices
    return B200_OK;
}
b200_status b200_memcpy_peer_async(b200_ctx * ctx, void * dst, int dst_device, const void * src, int src_device, size_t size) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    if (size == 0) return B200_OK;
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    if (dst_device == src_device) CU(cudaMemcpyAsync(dst, src, size, cudaMemcpyDeviceToDevice, ctx->stream));
    else CU(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, size, ctx->stream));
    return B200_OK;
}
b200_status b200_enable_peer_access(int device, int peer) {
    if (device == peer) return B200_OK;
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, device, peer));
    if (!can) return fail(B200_ERR_UNSUPPORTED, "device %d cannot access device %d directly", device, peer);
    CU(cudaSetDevice(device));
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(B200_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", device, peer, cudaGetErrorString(e));
    cudaGetLastError();
    return B200_OK;
}

b200_status b200_event_create(int device, void ** event) {
    CU(cudaSetDevice(device));
    cudaEvent_t ev;
    CU(cudaEventCreate(&ev));
    *event = (void *) ev;
    return B200_OK;
}
b200_status b200_event_destroy(int device, void * event) {
    CU(cudaSetDevice(device));
    CU(cudaEventDestroy((cudaEvent_t) event));
    return B200_OK;
}
b200_status b200_event_record(b200_ctx * ctx, void * event) {
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaEventRecord((cudaEvent_t) event, ctx->stream));
    return B200_OK;
}
b200_status b200_event_wait(b200_ctx * ctx, void * event) {
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamWaitEvent(ctx->stream, (cudaEvent_t) event, 0));
    return B200_OK;
}
b200_status b200_event_synchronize(void * event) {
    CU(cudaEventSynchronize((cudaEvent_t) event));
    return B200_OK;
}
b200_status b200_event_elapsed_ms(void * start, void * stop, float * ms) {
    CU(cudaEventElapsedTime(ms, (cudaEvent_t) start, (cudaEvent_t) stop));
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ CUDA graphs
//
// graph_compute of a decode step is a chain of short launches: ~2.5 us of host time each when launched one by one. The ggml
// backend captures the launches of a cgraph it has seen before into a CUDA graph and replays it (the reference does the same,
// ggml-cuda.cu:3679-3754); programmatic-dependent-launch edges between the mat-vec kernels survive the capture.

struct b200_graph {
    cudaGraphExec_t exec;
    uint64_t        launches;  // kernels inside
    int             device;
};

b200_status b200_graph_capture_begin(b200_ctx * ctx) {
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
    ctx->capture_launches0 = ctx->launches;
    return B200_OK;
}

b200_status b200_graph_capture_end(b200_ctx * ctx, b200_graph ** out) {
    if (!ctx || !out) return fail(B200_ERR_ARG, "null argument");
    *out = nullptr;
    CU(cudaSetDevice(ctx->device));
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(ctx->stream, &g);
    const uint64_t n = ctx->launches - ctx->capture_launches0;
    ctx->launches = ctx->capture_launches0;  // nothing has run yet: b200_graph_launch counts the kernels when they do
    if (e != cudaSuccess || !g) {
        cudaGetLastError();
        if (g) cudaGraphDestroy(g);
        return fail(B200_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
    }
    cudaGraphExec_t x = nullptr;
    e = cudaGraphInstantiate(&x, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e));
    }
    *out = new b200_graph{ x, n, ctx->device };
    return B200_OK;
}

b200_status b200_graph_launch(b200_ctx * ctx, b200_graph * g) {
    if (!ctx || !g) return fail(B200_ERR_ARG, "null argument");
    CU(cudaSetDevice(ctx->device));
    CU(cudaGraphLaunch(g->exec, ctx->stream));
    ctx->launches += g->launches;
    return B200_OK;
}

void b200_graph_destroy(b200_graph * g) {
    if (!g) return;
    cudaSetDevice(g->device);
    cudaGraphExecDestroy(g->exec);
    delete g;
}

uint64_t b200_graph_kernel_count(const b200_graph * g) { return g ? g->launches : 0; }

// ------------------------------------------------------------------------------------------------ K1

static b200_status quantize_ggml_api(b200_ctx * ctx, int val, int64_t blk, const float * x, int64_t k, int64_t nrows, size_t stride, void * y) {
    if (!ctx || !x || !y) return fail(B200_ERR_ARG, "null argument");
    if (k <= 0 || k % blk != 0) return fail(B200_ERR_UNSUPPORTED, "k=%lld is not a multiple of %lld", (long long) k, (long long) blk);
    if (stride % 4 != 0) return fail(B200_ERR_ARG, "row stride must be a multiple of 4 bytes");
    CU(cudaSetDevice(ctx->device));
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    int l = 0;
    CU(quantize_ggml(val, x, k, nrows, stride, y, ctx->stream, &l));
    ctx->launches += l;
    return B200_OK;
}
b200_status b200_quantize_q8_0(b200_ctx * ctx, const float * x, int64_t k, int64_t nrows, size_t s, void * y) { return quantize_ggml_api(ctx, 0, 32, x, k, nrows, s, y); }
b200_status b200_quantize_q8_1(b200_ctx * ctx, const float * x, int64_t k, int64_t nrows, size_t s, void * y) { return quantize_ggml_api(ctx, 2, 32, x, k, nrows, s, y); }
b200_status b200_quantize_q8_K(b200_ctx * ctx, const float * x, int64_t k, int64_t nrows, size_t s, void * y) { return quantize_ggml_api(ctx, 1, 256, x, k, nrows, s, y); }

b200_status b200_quantize_units(b200_ctx * ctx, int kind, const float * x, int64_t k, int64_t nrows, size_t stride, void * main, void * aux) {
    if (!ctx || !x || !main || !aux) return fail(B200_ERR_ARG, "null argument");
    if (k <= 0 || k % 256 != 0) return fail(B200_ERR_UNSUPPORTED, "k=%lld is not a multiple of 256", (long long) k);
    if (kind != ACT_KIND_Q8_0 && kind != ACT_KIND_Q8_K) return fail(B200_ERR_ARG, "kind must be 0 or 1");
    CU(cudaSetDevice(ctx->device));
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    int l = 0;
    CU(quantize_units(kind, x, k, nrows, stride, main, aux, ctx->stream, &l));
    ctx->launches += l;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ MUL_MAT

static bool type_ok(int t) { return t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K; }

static size_t row_bytes(int type, int64_t k) { return (size_t) (k / type_block_elems(type)) * type_block_bytes(type); }

int b200_mul_mat_supported(const b200_mul_mat_desc * d) {
    if (!d || !type_ok(d->type)) return 0;
    if (d->ne00 <= 0 || d->ne00 % type_block_elems(d->type) != 0) return 0;
    if (d->ne01 < 0 || d->ne02 <= 0 || d->ne03 <= 0 || d->ne11 < 0 || d->ne12 <= 0 || d->ne13 <= 0) return 0;
    if (d->ne12 % d->ne02 != 0 || d->ne13 % d->ne03 != 0) return 0;
    if (d->nb01 < row_bytes(d->type, d->ne00)) return 0;
    if (d->nb11 % 4 || d->nb12 % 4 || d->nb13 % 4 || d->nb1 % 4 || d->nb2 % 4 || d->nb3 % 4) return 0;
    if (d->ne01 > INT32_MAX || d->ne00 > INT32_MAX) return 0;
    return 1;
}

struct mv_cfg { int nw, ns; size_t smem; };

// Two CTAs of <= 113 KB fit one SM (228 KB, 1 KB reserved per CTA): with that budget the next mat-vec kernel of the stream
// becomes co-resident (programmatic dependent launch) and streams its first stages while this one computes. Configurations
// beyond it get the SM to themselves.
static const size_t MV_SMEM_CORESIDENT = 113 * 1024;
static const size_t MV_SMEM_MAX        = 227 * 1024;

static bool pick_mv_cfg(const b200_ctx * ctx, int type, int U, int nc, mv_cfg * out, bool lean = false) {
    if (lean && !(ctx->mv_warps > 0 && ctx->mv_stages > 0)) {
        // mv_lean_kernel consumes two stages at a time: a 4-deep ring keeps one pair in flight while one is being multiplied.
        // measured on the Llama-3-8B stack (profiles/r02): 11 x 4 beats 12 x 4, 10 x 4, 8 x 4, 8 x 6 and every 2-deep ring
        static const int lean_cfg[][2] = { { 11, 4 }, { 10, 4 }, { 8, 4 }, { 6, 4 }, { 4, 4 }, { 2, 4 }, { 1, 2 } };
        for (auto & c : lean_cfg) {
            const size_t s = mv_fast_smem(type, U, nc, c[0], c[1]);
            if (s <= MV_SMEM_MAX) { *out = { c[0], c[1], s }; return true; }
        }
        return false;
    }
    if (ctx->mv_warps > 0 && ctx->mv_stages > 0 && ctx->mv_warps <= (nc <= MV_WIDE_MAX_COLS ? MV_MAX_CONS : 8)) {
        const size_t s = mv_fast_smem(type, U, nc, ctx->mv_warps, ctx->mv_stages);
        if (s <= MV_SMEM_MAX) { *out = { ctx->mv_warps, ctx->mv_stages, s }; return true; }
    }
    // measured on the Llama-3-8B decode stack (profiles/): 16 consumer warps x 2 stages beat every 8- and 12-warp ring
    if (nc <= MV_WIDE_MAX_COLS) {
        const size_t s = mv_fast_smem(type, U, nc, 16, 2);
        if (s <= MV_SMEM_MAX) { *out = { 16, 2, s }; return true; }
    }
    static const int small[][2] = { { 8, 3 }, { 8, 2 } };
    if (nc <= 4) {  // the <= 4 column instantiations are compiled for two CTAs per SM (<= 112 registers)
        for (auto & c : small) {
            const size_t s = mv_fast_smem(type, U, nc, c[0], c[1]);
            if (s <= MV_SMEM_CORESIDENT) { *out = { c[0], c[1], s }; return true; }
        }
    }
    static const int big[][2] = { { 8, 4 }, { 8, 3 }, { 8, 2 }, { 6, 2 }, { 4, 2 }, { 2, 2 }, { 1, 2 } };
    for (auto & c : big) {
        const size_t s = mv_fast_smem(type, U, nc, c[0], c[1]);
        if (s <= MV_SMEM_MAX) { *out = { c[0], c[1], s }; return true; }
    }
    return false;
}

static void mv_geometry(int U, MvParams * p) {
    if (U <= 32) {
        int L = 1;
        while (L < U) L <<= 1;
        p->L = L; p->rpg = 32 / L; p->slices = 1;
    } else {
        p->L = 32; p->rpg = 1; p->slices = (U + 31) / 32;
    }
    p->U = U;
}

static bool fast_layout_ok(int type, int64_t K, const void * w, size_t nb01, size_t nb02, size_t nb03) {
    if (type == B200_TYPE_Q6_K && K % 2048 != 0) return false;  // 210-byte units: a row slice is a whole number of 16-byte words only 8 units at a time
    return K % 256 == 0 && (uintptr_t) w % 16 == 0 && nb01 % 16 == 0 && nb02 % 16 == 0 && nb03 % 16 == 0;
}

// the mat-vec kernel quantizes src1 itself (no K1 launch, no scratch round trip) while the per-CTA redundancy is small
static bool mv_fuse_quant(const b200_ctx * ctx, int U, int64_t ncols) { return ctx->fuse_quant && ncols <= MV_MAX_COLS && ncols * U <= 64; }

// run `n_ops` matrices (same type, K) against `ncols` activation columns: either F32 columns x (+ c*x_cs bytes), quantized
// inside the kernel, or already-quantized unit-record columns (act_main / act_aux)
// K7 epilogue of a one-column launch: bias rows per matrix; glu >= 0: ws[0] = gate, ws[1] = up, dsts[0] = the GLU's output
struct mv_epilogue { int glu; float alpha, limit; const float * bias[MV_MAX_MATS]; };

static b200_status mv_fast_run(b200_ctx * ctx, int type, int64_t K, int n_ops, const uint8_t * const * ws, const size_t * nb01s,
                               const int64_t * Ms, float * const * dsts, const int64_t * dst_cs, int ncols,
                               const float * x, size_t x_cs, const uint8_t * act_main, const uint8_t * act_aux, const mv_epilogue * ep = nullptr) {
    const int U = (int) (K / 256);
    for (int c0 = 0; c0 < ncols;) {
        int     nc = ncols - c0 < MV_MAX_COLS ? ncols - c0 : MV_MAX_COLS;
        if (!ep && ctx->mma_min_cols > 0 && nc >= ctx->mma_min_cols && mv_mma_usable(type) && mv_mma_fits(type, K)) {  // 3..8 columns: tensor-core mat-vec (mvmma.cu)
            float * ds[MV_MAX_MATS * 8];
            for (int i = 0; i < n_ops; i++) ds[i] = dsts[i] + (int64_t) c0 * dst_cs[i];
            if (x && ctx->split.n_ranks > 1 && ctx->split_deferred) {
                if (b200_status fst = b200_split_flush(ctx)) return fst;  // this kernel reads src1 without a split_wait prologue
            }
            b200_status st = x ? b200_mv_mma_run(ctx, type, K, n_ops, ws, nb01s, Ms, ds, dst_cs, nc, nullptr, nullptr,
                                                 (const float *) ((const char *) x + (size_t) c0 * x_cs), x_cs)
                               : b200_mv_mma_run(ctx, type, K, n_ops, ws, nb01s, Ms, ds, dst_cs, nc, act_main + (size_t) c0 * U * ACT_MAIN,
                                                 act_aux + (size_t) c0 * U * ACT_AUX);
            if (st != B200_OK) return st;
            c0 += nc;
            continue;
        }
        mv_cfg  cfg;
        // the lean kernel takes the K7 epilogues (bias rows, GLU pairs) of one device
        const bool lean = (!ep || (ctx->split.n_ranks <= 1 && nc == 1)) && (ctx->split.n_ranks <= 1 || (ctx->ll_on && !ctx->split_deferred)) && !(ctx->dbg & 1) && mv_lean_type(type);
        while (nc >= 1 && !pick_mv_cfg(ctx, type, U, nc, &cfg, lean && nc <= MV_LEAN_MAX_COLS)) nc--;
        if (nc < 1) return fail(B200_ERR_UNSUPPORTED, "K=%lld: activations do not fit in shared memory", (long long) K);
        for (int o0 = 0; o0 < n_ops; o0 += MV_MAX_MATS) {
            MvParams p;
            memset(&p, 0, sizeof(p));
            mv_geometry(U, &p);
            p.n_mats = n_ops - o0 < MV_MAX_MATS ? n_ops - o0 : MV_MAX_MATS;
            int g = 0;
            for (int i = 0; i < p.n_mats; i++) {
                p.mat[i].w      = ws[o0 + i];
                p.mat[i].nb01   = (int64_t) nb01s[o0 + i];
                p.mat[i].dst    = dsts[o0 + i] + (int64_t) c0 * dst_cs[o0 + i];
                p.mat[i].dst_cs = dst_cs[o0 + i];
                p.mat[i].M      = (int) Ms[o0 + i];
                p.mat[i].group0 = g;
                g += (int) ((Ms[o0 + i] + p.rpg - 1) / p.rpg);
            }
            for (int i = p.n_mats; i < MV_MAX_MATS; i++) p.mat[i].group0 = INT32_MAX;
            if (ep) {
                for (int i = 0; i < p.n_mats; i++) p.mat[i].bias = ep->bias[o0 + i];
                if (ep->glu >= 0) {  // gate rows and up rows of a group run back to back on one warp (kernels.h MvEpi)
                    p.mat[1].group0 = 0;
                    g               = (int) ((Ms[0] + p.rpg - 1) / p.rpg);
                    p.epi.glu       = ep->glu + 1;
                    p.epi.alpha     = ep->alpha;
                    p.epi.limit     = ep->limit;
                }
            }
            p.total_groups = g;
            if (g == 0) continue;
            p.n_cons   = cfg.nw;
            p.n_stages = cfg.ns;
            p.pdl      = ctx->pdl;
            p.split    = ctx->split;
            if (b200_split_ll_geom(ctx, p.n_mats, Ms + o0, dsts + o0, dst_cs + o0, c0 + nc, &p.geom)) {
                p.split.ll = 1;  // flagged-payload exchange: no flag round trip, nothing to flush
            } else if (ctx->split.n_ranks > 1) {
                if (ctx->split_deferred && x) {  // src1 is read inside the kernel, behind its split_wait prologue
                    p.split.deferred   = 1;
                    ctx->split_pending = 1;
                } else if (b200_status st = b200_split_flush(ctx)) {
                    return st;
                }
            }
            if (x) {
                p.fuse_quant = 1;
                p.x          = (const float *) ((const char *) x + (size_t) c0 * x_cs);
                p.x_cs       = (int64_t) x_cs;
                p.x_vec      = (uintptr_t) p.x % 16 == 0 && x_cs % 16 == 0;
            } else {
                p.act_main = act_main + (size_t) c0 * U * ACT_MAIN;
                p.act_aux  = act_aux + (size_t) c0 * U * ACT_AUX;
            }
            if (ctx->norm.on && x && ncols == 1 && !(ctx->dbg & 4) && p.x_vec && p.split.n_ranks <= 1 && (uintptr_t) ctx->norm.w % 16 == 0) {
                // MvNorm: the row's RMS_NORM (+ MUL) in the K1 prologue - a GLU pair on the EPI kernel, or a plain launch on the lean kernel
                const bool glu = ep && ep->glu >= 0, bias_only = ep && ep->glu < 0, lean_run = !bias_only && lean && mv_lean_takes(type, 1, p, 1);
                const int  nq  = lean_run ? std::min(MV_MAX_CONS, std::max(cfg.nw, U)) : cfg.nw;
                if ((glu || lean_run) && U <= (lean_run ? 2 : 4) * nq) { p.norm = ctx->norm; ctx->norm_used = true; }   // units a warp holds: mmvq.cu NB
            }
            if (ctx->rope.on && x && ncols == 1 && n_ops == 3 && p.n_mats == 3 && !ep && !(p.rpg & 1) && p.split.n_ranks <= 1 && lean &&
                mv_lean_takes(type, 1, p, 1)) {
                p.rope = ctx->rope;   // MvRope: the rows leave the kernel rotated / as KV-cache rows
                ctx->rope_used = true;
            }
            const int n_ctas = g < ctx->sm_count ? g : ctx->sm_count;
            if (ctx->prefetch_kb > 0) {
                if (ctx->mem_gen != b200_mem_generation()) {  // an allocation went away: forget what was learned
                    ctx->next_of.clear();
                    ctx->last_key = nullptr;
                    ctx->mem_gen  = b200_mem_generation();
                }
                const void * key = p.mat[0].w;
                auto it = ctx->next_of.find(key);
                if (it != ctx->next_of.end()) p.next = it->second;
                if (ctx->last_key) {  // this launch is what followed `last_key`
                    if (ctx->next_of.size() > 65536) ctx->next_of.clear();
                    MvNext & nx = ctx->next_of[ctx->last_key];
                    memset(&nx, 0, sizeof(nx));
                    for (int i = 0; i < p.n_mats; i++) {
                        nx.w[i] = p.mat[i].w; nx.nb01[i] = p.mat[i].nb01; nx.M[i] = p.mat[i].M; nx.group0[i] = p.mat[i].group0;
                    }
                    nx.n_mats         = p.n_mats;
                    nx.total_groups   = p.total_groups;
                    nx.n_ctas         = n_ctas;
                    nx.rpg            = p.rpg;
                    nx.row_bytes      = U * unit_bytes(type);
                    const int64_t gb  = (int64_t) p.rpg * nx.row_bytes;
                    nx.groups_per_cta = (int) (((int64_t) ctx->prefetch_kb * 1024 + gb - 1) / gb);
                    if (p.epi.glu) {  // a group streams rows of both matrices: half as many groups for the same bytes
                        nx.glu            = 1;
                        nx.groups_per_cta = (nx.groups_per_cta + 1) / 2;
                    }
                }
                ctx->last_key = key;
            }
            p.dbg = ctx->dbg;
            if (ctx->trace) p.trace = ctx->trace + (size_t) (ctx->launches % ctx->trace_slots) * ctx->sm_count * 8;
            CU(mv_fast_launch(type, nc, p, n_ctas, 1, cfg.smem, ctx->stream));
            ctx->launches++;
        }
        c0 += nc;
    }
    return B200_OK;
}

static b200_status mul_mat_impl(b200_ctx * ctx, int n_ops, const b200_mul_mat_desc * descs, const void * const * src0s,
                                const float * src1, float * const * dsts) {
    const b200_mul_mat_desc & d = descs[0];
    const int     type = d.type;
    const int64_t K = d.ne00, N = d.ne11;
    if (N == 0 || d.ne01 == 0) return B200_OK;
    CU(cudaSetDevice(ctx->device));

    // 4..32 columns of q4_K: the tcgen05 mat-vec that expands the raw blocks in shared memory (mvtc.cuh)
    if (ctx->force_path == B200_PATH_AUTO && ctx->split.n_ranks <= 1) {
        bool tc = true;
        for (int i = 0; i < n_ops && tc; i++) tc = b200_mvtc_usable(ctx, type, K, N, src0s[i], descs[i].nb01);
        if (tc) {
            bool first = true;
            for (int64_t i13 = 0; i13 < d.ne13; i13++)
                for (int64_t i12 = 0; i12 < d.ne12; i12++) {
                    for (int i = 0; i < n_ops; i++) {
                        const b200_mul_mat_desc & di = descs[i];
                        const int64_t r2 = di.ne12 / di.ne02, r3 = di.ne13 / di.ne03;
                        b200_status st = b200_mvtc_run(ctx, (const uint8_t *) src0s[i] + (i12 / r2) * di.nb02 + (i13 / r3) * di.nb03, di.nb01, di.ne01, K,
                                                       (const float *) ((const char *) src1 + i12 * di.nb12 + i13 * di.nb13), di.nb11, N,
                                                       (float *) ((char *) dsts[i] + i12 * di.nb2 + i13 * di.nb3), (int64_t) (di.nb1 / 4),
                                                       /* act_shared = */ !first && i > 0);
                        if (st != B200_OK) return st;
                        first = false;
                    }
                    first = true;
                }
            return B200_OK;
        }
    }
    // more columns than the mat-vec kernels hold in registers: the tcgen05 int8 tile kernel (K4)
    const int64_t mmq_min = ctx->mmq_min_cols > 0 ? ctx->mmq_min_cols : (ctx->weight_cache ? 9 : 33);
    bool mmq = ctx->force_path == B200_PATH_MMQ || (ctx->force_path == B200_PATH_AUTO && N >= mmq_min);
    for (int i = 0; i < n_ops && mmq; i++) mmq = b200_mmq_usable(type, K, src0s[i], descs[i].nb01);
    if (mmq) {
        if (b200_status fst = b200_split_flush(ctx)) return fst;
        if (n_ops > 1 && d.ne12 == 1 && d.ne13 == 1) {  // q/k/v, gate/up: one launch for the whole group when every expansion is cached
            const void * ws[MV_MAX_MATS]; size_t nb01s[MV_MAX_MATS]; int64_t Ms[MV_MAX_MATS], ldcs[MV_MAX_MATS];
            bool plain = n_ops <= MV_MAX_MATS;
            for (int i = 0; i < n_ops && plain; i++) {
                plain = descs[i].ne12 == 1 && descs[i].ne13 == 1 && descs[i].ne02 == 1 && descs[i].ne03 == 1;
                ws[i] = src0s[i]; nb01s[i] = descs[i].nb01; Ms[i] = descs[i].ne01; ldcs[i] = (int64_t) (descs[i].nb1 / 4);
            }
            if (plain) {
                const b200_status gst = b200_mmq_run_group(ctx, type, n_ops, ws, nb01s, Ms, K, src1, d.nb11, N, dsts, ldcs);
                if (gst != B200_ERR_UNSUPPORTED) return gst;
            }
        }
        for (int i = 0; i < n_ops; i++) {
            const b200_mul_mat_desc & di = descs[i];
            const int64_t r2 = di.ne12 / di.ne02, r3 = di.ne13 / di.ne03;
            for (int64_t i13 = 0; i13 < di.ne13; i13++)
                for (int64_t i12 = 0; i12 < di.ne12; i12++) {
                    b200_status st = b200_mmq_run(ctx, type, (const uint8_t *) src0s[i] + (i12 / r2) * di.nb02 + (i13 / r3) * di.nb03, di.nb01,
                                                  di.ne01, K, (const float *) ((const char *) src1 + i12 * di.nb12 + i13 * di.nb13), di.nb11, N,
                                                  (float *) ((char *) dsts[i] + i12 * di.nb2 + i13 * di.nb3), (int64_t) (di.nb1 / 4),
                                                  /* act_shared = */ i > 0 && di.ne12 == 1 && di.ne13 == 1);
                    if (st != B200_OK) return st;
                }
        }
        return B200_OK;
    }

    bool fast = ctx->force_path != B200_PATH_SIMPLE;
    for (int i = 0; i < n_ops; i++) fast = fast && fast_layout_ok(type, K, src0s[i], descs[i].nb01, descs[i].nb02, descs[i].nb03);

    const int64_t n_batch = d.ne12 * d.ne13;
    const int64_t n_rows  = N * n_batch;  // activation rows to quantize

    if (fast) {
        const int  U    = (int) (K / 256);
        const bool mma  = ctx->mma_min_cols > 0 && N >= ctx->mma_min_cols && mv_mma_usable(type);  // the IMMA kernel reads pre-quantized columns
        // src1 is quantized inside the mat-vec kernel while the per-CTA redundancy is small (no quantizer launch); the IMMA kernel
        // has its own bound (it is used for more columns) and needs the row to fit its 8 columns in shared memory
        const bool fuse = mma ? (ctx->fuse_quant && N <= MV_MAX_COLS && N * U <= ctx->mma_fuse_units && mv_mma_fits(type, K)) : mv_fuse_quant(ctx, U, N);
        uint8_t * am = nullptr, * ax = nullptr;
        if (!fuse) {
            if (b200_status fst = b200_split_flush(ctx)) return fst;  // the quantizer launch reads src1 without a split_wait prologue
            const size_t main_b = (size_t) n_rows * U * ACT_MAIN, aux_b = (size_t) n_rows * U * ACT_AUX;
            const size_t aux_o  = (main_b + 255) & ~(size_t) 255;
            b200_status  st     = ensure(ctx, &ctx->scratch, &ctx->scratch_size, aux_o + aux_b);
            if (st != B200_OK) return st;
            am = (uint8_t *) ctx->scratch;
            ax = am + aux_o;
            const int kind = act_kind_for(type);
            for (int64_t i13 = 0; i13 < d.ne13; i13++)
                for (int64_t i12 = 0; i12 < d.ne12; i12++) {
                    const int64_t b = i13 * d.ne12 + i12;
                    int l = 0;
                    CU(quantize_units(kind, (const float *) ((const char *) src1 + i12 * d.nb12 + i13 * d.nb13), K, N, d.nb11,
                                      am + (size_t) b * N * U * ACT_MAIN, ax + (size_t) b * N * U * ACT_AUX, ctx->stream, &l, ctx->pdl));
                    ctx->launches += l;
                }
        }
        const uint8_t * ws[MV_MAX_MATS * 8];
        size_t          nb01s[MV_MAX_MATS * 8];
        int64_t         Ms[MV_MAX_MATS * 8], dcs[MV_MAX_MATS * 8];
        float *         ds[MV_MAX_MATS * 8];
        if (n_ops > MV_MAX_MATS * 8) return fail(B200_ERR_ARG, "too many fused ops");
        for (int64_t i13 = 0; i13 < d.ne13; i13++)
            for (int64_t i12 = 0; i12 < d.ne12; i12++) {
                const int64_t b = i13 * d.ne12 + i12;
                for (int i = 0; i < n_ops; i++) {
                    const b200_mul_mat_desc & di = descs[i];
                    const int64_t r2 = di.ne12 / di.ne02, r3 = di.ne13 / di.ne03;
                    ws[i]    = (const uint8_t *) src0s[i] + (i12 / r2) * di.nb02 + (i13 / r3) * di.nb03;
                    nb01s[i] = di.nb01;
                    Ms[i]    = di.ne01;
                    ds[i]    = (float *) ((char *) dsts[i] + i12 * di.nb2 + i13 * di.nb3);
                    dcs[i]   = (int64_t) (di.nb1 / 4);
                }
                b200_status st = mv_fast_run(ctx, type, K, n_ops, ws, nb01s, Ms, ds, dcs, (int) N,
                                             fuse ? (const float *) ((const char *) src1 + i12 * d.nb12 + i13 * d.nb13) : nullptr, d.nb11,
                                             fuse ? nullptr : am + (size_t) b * N * U * ACT_MAIN, fuse ? nullptr : ax + (size_t) b * N * U * ACT_AUX);
                if (st != B200_OK) return st;
            }
        return B200_OK;
    }

    if (b200_status fst = b200_split_flush(ctx)) return fst;
    // generic path: reference block layout for the activations, one warp per row
    const int    val    = type_is_k(type) ? 1 : 0;
    const size_t act_rs = type_is_k(type) ? (size_t) (K / QK_K) * SZ_Q8_K : (size_t) (K / QK) * SZ_Q8_0;
    b200_status  st     = ensure(ctx, &ctx->scratch, &ctx->scratch_size, (size_t) n_rows * act_rs + 16);
    if (st != B200_OK) return st;
    uint8_t * act = (uint8_t *) ctx->scratch;
    for (int64_t i13 = 0; i13 < d.ne13; i13++)
        for (int64_t i12 = 0; i12 < d.ne12; i12++) {
            const int64_t b = i13 * d.ne12 + i12;
            int l = 0;
            CU(quantize_ggml(val, (const float *) ((const char *) src1 + i12 * d.nb12 + i13 * d.nb13), K, N, d.nb11,
                             act + (size_t) b * N * act_rs, ctx->stream, &l));
            ctx->launches += l;
            for (int i = 0; i < n_ops; i++) {
                const b200_mul_mat_desc & di = descs[i];
                const int64_t r2 = di.ne12 / di.ne02, r3 = di.ne13 / di.ne03;
                for (int64_t c0 = 0; c0 < N; c0 += 65535) {
                    MvSimpleParams sp;
                    memset(&sp, 0, sizeof(sp));
                    sp.w      = (const uint8_t *) src0s[i] + (i12 / r2) * di.nb02 + (i13 / r3) * di.nb03;
                    sp.nb01   = (int64_t) di.nb01;
                    sp.act    = act + (size_t) (b * N + c0) * act_rs;
                    sp.act_rs = (int64_t) act_rs;
                    sp.dst    = (float *) ((char *) dsts[i] + i12 * di.nb2 + i13 * di.nb3) + c0 * (int64_t) (di.nb1 / 4);
                    sp.dst_cs = (int64_t) (di.nb1 / 4);
                    sp.M      = (int) di.ne01;
                    sp.K      = (int) K;
                    sp.ncols  = (int) (N - c0 < 65535 ? N - c0 : 65535);
                    sp.exact  = ctx->exact;
                    CU(mv_simple_launch(type, sp, ctx->stream));
                    ctx->launches++;
                }
            }
        }
    return B200_OK;
}

// Row split: a rank whose slice of a MUL_MAT is EMPTY (get_row_split rounds boundaries down: M = 512 over 8 ranks with rounding 128
// leaves four ranks without rows) still takes part in the exchange - both protocols count launches in the sync words, and in the
// flagged-payload protocol the rank must collect the peers' rows. One small launch: no weights, no dot products.
__global__ void __launch_bounds__(256) split_empty_kernel(const __grid_constant__ MvParams p, int ncols) {
    if (p.split.ll) {
        const uint32_t ep = ll_epoch(p.split);
        ll_receive(p.split, ep, p.geom, p.mat, p.n_mats, ncols, (int) (blockIdx.x * blockDim.x + threadIdx.x), (int) (gridDim.x * blockDim.x));
        __syncthreads();
        if (threadIdx.x == 0) ll_finish(p.split, ep, gridDim.x);
    } else if (threadIdx.x < 32) {
        split_complete(p.split, gridDim.x);
    }
}

static b200_status split_empty_participant(b200_ctx * ctx, int n_ops, const b200_mul_mat_desc * descs, float * const * dsts) {
    CU(cudaSetDevice(ctx->device));
    for (int o0 = 0; o0 < n_ops; o0 += MV_MAX_MATS) {
        MvParams p;
        memset(&p, 0, sizeof(p));
        p.n_mats = n_ops - o0 < MV_MAX_MATS ? n_ops - o0 : MV_MAX_MATS;
        int64_t Ms[MV_MAX_MATS], dcs[MV_MAX_MATS];
        const int ncols = (int) descs[0].ne11;
        for (int i = 0; i < p.n_mats; i++) {
            p.mat[i].dst    = dsts[o0 + i];
            p.mat[i].dst_cs = (int64_t) (descs[o0 + i].nb1 / 4);
            Ms[i] = 0; dcs[i] = p.mat[i].dst_cs;
        }
        p.split = ctx->split;
        // the decode kernels (at most 8 columns, plain 2-D MUL_MAT) use the flagged-payload exchange when it is configured: the
        // peers will send records, so this rank has to receive them; everything else runs the flag protocol
        const bool ll = ncols <= MV_MAX_COLS && descs[0].ne12 == 1 && descs[0].ne13 == 1 &&
                        b200_split_ll_geom(ctx, p.n_mats, Ms, dsts + o0, dcs, ncols, &p.geom);
        if (ll) p.split.ll = 1;
        else if (b200_status st = b200_split_flush(ctx)) return st;
        split_empty_kernel<<<ll ? 8 : 1, 256, 0, ctx->stream>>>(p, ncols);
        CU(cudaGetLastError());
        ctx->launches++;
    }
    return B200_OK;
}

b200_status b200_mul_mat(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * src0, const float * src1, float * dst) {
    if (!ctx || !d) return fail(B200_ERR_ARG, "null argument");
    if (!b200_mul_mat_supported(d)) return fail(B200_ERR_UNSUPPORTED, "MUL_MAT geometry/type not supported");
    if (d->ne11 > 0 && d->ne01 == 0 && ctx->split.n_ranks > 1 && dst) {  // an empty row slice still takes part in the exchange
        float * ds0[1] = { dst };
        return split_empty_participant(ctx, 1, d, ds0);
    }
    if (d->ne11 == 0 || d->ne01 == 0) return B200_OK;  // empty tensors carry no storage (ggml_is_empty nodes are skipped upstream too)
    if (!src0 || !src1 || !dst) return fail(B200_ERR_ARG, "null argument");
    const void * s0[1] = { src0 };
    float *      ds[1] = { dst };
    return mul_mat_impl(ctx, 1, d, s0, src1, ds);
}

b200_status b200_mul_mat_shared_src1(b200_ctx * ctx, int n_ops, const b200_mul_mat_desc * descs, const void * const * src0s,
                                     const float * src1, float * const * dsts) {
    if (!ctx || !descs || !src0s || !src1 || !dsts || n_ops < 1) return fail(B200_ERR_ARG, "bad argument");
    for (int i = 0; i < n_ops; i++) {
        if (!b200_mul_mat_supported(&descs[i])) return fail(B200_ERR_UNSUPPORTED, "MUL_MAT %d geometry/type not supported", i);
        const b200_mul_mat_desc & a = descs[0], & b = descs[i];
        if (a.type != b.type || a.ne00 != b.ne00 || a.ne11 != b.ne11 || a.ne12 != b.ne12 || a.ne13 != b.ne13 || a.nb11 != b.nb11 ||
            a.nb12 != b.nb12 || a.nb13 != b.nb13) {
            return fail(B200_ERR_ARG, "fused MUL_MATs must share type, K and the src1 geometry");
        }
    }
    if (ctx->split.n_ranks > 1 && descs[0].ne11 > 0) {
        bool all_empty = true;
        for (int i = 0; i < n_ops; i++) all_empty = all_empty && descs[i].ne01 == 0;
        if (all_empty) return split_empty_participant(ctx, n_ops, descs, dsts);
    }
    return mul_mat_impl(ctx, n_ops, descs, src0s, src1, dsts);
}

// RMS_NORM (+ MUL) in front of the MUL_MATs: the mat-vec kernel normalizes the one activation column in its own prologue (MvNorm);
// whenever the call does not end up on such a kernel (more columns, another path) the norm runs as its own launch into norm->y first
static b200_status norm_then(b200_ctx * ctx, const b200_mul_mat_desc & d, const float * x, const b200_norm * nm) {
    b200_tensor tx = {}, tw = {}, ty = {};
    tx.data = (void *) x; tx.type = B200_TYPE_F32;
    tx.ne[0] = d.ne00; tx.ne[1] = d.ne11; tx.ne[2] = d.ne12; tx.ne[3] = d.ne13;
    tx.nb[0] = 4; tx.nb[1] = d.nb11; tx.nb[2] = d.nb12; tx.nb[3] = d.nb13;
    ty = tx; ty.data = nm->y;
    ty.nb[1] = nm->y_nb1; ty.nb[2] = nm->y_nb1 * (size_t) d.ne11; ty.nb[3] = ty.nb[2] * (size_t) d.ne12;
    if (!nm->w) return b200_rms_norm(ctx, &tx, &ty, nm->eps);
    tw.data = (void *) nm->w; tw.type = B200_TYPE_F32;
    tw.ne[0] = d.ne00; tw.ne[1] = tw.ne[2] = tw.ne[3] = 1;
    tw.nb[0] = 4; tw.nb[1] = tw.nb[2] = tw.nb[3] = (size_t) d.ne00 * 4;
    return b200_rms_norm_mul(ctx, &tx, &tw, &ty, nm->eps);
}
static bool norm_worth_trying(const b200_ctx * ctx, const b200_mul_mat_desc & d) {
    return d.ne11 == 1 && d.ne12 == 1 && d.ne13 == 1 && d.ne00 % 256 == 0 && d.ne00 <= 16384 && ctx->force_path == B200_PATH_AUTO && !ctx->exact &&
           ctx->split.n_ranks <= 1 && ctx->fuse_quant;
}
static b200_mul_mat_desc with_src1_rows(b200_mul_mat_desc d, size_t nb1) {
    d.nb11 = nb1; d.nb12 = nb1 * (size_t) d.ne11; d.nb13 = d.nb12 * (size_t) d.ne12;
    return d;
}

b200_status b200_mul_mat_shared_src1_norm(b200_ctx * ctx, int n_ops, const b200_mul_mat_desc * descs, const void * const * src0s, const float * x,
                                          const b200_norm * nm, float * const * dsts) {
    if (!ctx || !descs || !nm || !nm->y || n_ops < 1 || n_ops > 8 || nm->y_nb1 % 4) return fail(B200_ERR_ARG, "bad argument");
    if (norm_worth_trying(ctx, descs[0])) {
        ctx->norm = b200::MvNorm{ nm->w, nm->eps, 1 };
        ctx->norm_used = false;
        const b200_status st = b200_mul_mat_shared_src1(ctx, n_ops, descs, src0s, x, dsts);
        ctx->norm.on = 0;
        if (st != B200_OK || ctx->norm_used) return st;
    }
    if (b200_status st = norm_then(ctx, descs[0], x, nm)) return st;
    b200_mul_mat_desc dn[8];
    for (int i = 0; i < n_ops; i++) dn[i] = with_src1_rows(descs[i], nm->y_nb1);
    return b200_mul_mat_shared_src1(ctx, n_ops, dn, src0s, nm->y, dsts);
}

b200_status b200_mul_mat_fused_norm(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * up, const void * gate, const float * x, const b200_norm * nm,
                                    const b200_fusion * f, float * dst) {
    if (!ctx || !d || !nm || !nm->y || nm->y_nb1 % 4) return fail(B200_ERR_ARG, "bad argument");
    if (norm_worth_trying(ctx, *d)) {
        ctx->norm = b200::MvNorm{ nm->w, nm->eps, 1 };
        ctx->norm_used = false;
        const b200_status st = b200_mul_mat_fused(ctx, d, up, gate, x, f, dst);
        ctx->norm.on = 0;
        if (st != B200_OK || ctx->norm_used) return st;
    }
    if (b200_status st = norm_then(ctx, *d, x, nm)) return st;
    const b200_mul_mat_desc dn = with_src1_rows(*d, nm->y_nb1);
    return b200_mul_mat_fused(ctx, &dn, up, gate, nm->y, f, dst);
}

// q / k / v of one token with ROPE(q), ROPE(k), SET_ROWS(k cache), SET_ROWS(v cache) folded into the mat-vec launch (MvRope). When
// the launch cannot take them (another kernel, another layout) the projections go to dsts[] as usual and b200_attn_prep follows.
b200_status b200_qkv_rope(b200_ctx * ctx, const b200_mul_mat_desc * descs, const void * const * src0s, const float * x, const b200_norm * nm,
                          float * const * dsts, const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                          const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx,
                          const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache) {
    if (!ctx || !descs || !src0s || !x || !dsts) return fail(B200_ERR_ARG, "null argument");
    if (!b200_attn_prep_supported(q, q_dst, k, k_dst, pos, freq_factors, rp, k_idx, k_cache, v, v_idx, v_cache))
        return fail(B200_ERR_UNSUPPORTED, "qkv_rope: the ROPE / SET_ROWS nodes of one layer (see b200_attn_prep)");
    const int64_t D = q->ne[0];
    const bool one_token = descs[0].ne11 == 1 && q->ne[2] == 1 && q->ne[3] == 1 && k->ne[2] == 1 && v->ne[1] == 1 && v->ne[2] == 1 && v->ne[3] == 1;
    const bool fold = one_token && rp->mode == B200_ROPE_NORMAL && rp->n_dims == D && k_cache->type == B200_TYPE_F16 && v_cache->type == B200_TYPE_F16 &&
                      q_dst->nb[1] == (size_t) D * 4 && k_dst->nb[1] == (size_t) D * 4 && descs[0].ne01 == D * q->ne[1] && descs[1].ne01 == D * k->ne[1] &&
                      descs[2].ne01 == v->ne[0] && k_cache->ne[0] == descs[1].ne01 && v_cache->ne[0] == descs[2].ne01 && k_idx->ne[0] == 1 && v_idx->ne[0] == 1 &&
                      q->data == dsts[0] && k->data == dsts[1] && v->data == dsts[2];
    if (fold) {
        b200::MvRope r;
        memset(&r, 0, sizeof(r));
        r.on = 1; r.head_dim = (int) D;
        r.rp = b200_rope_consts(rp);
        r.pos = (const int32_t *) pos->data; r.ff = freq_factors ? (const float *) freq_factors->data : nullptr;
        r.q_dst = (float *) q_dst->data; r.k_dst = (float *) k_dst->data;
        r.k_idx = (const char *) k_idx->data; r.v_idx = (const char *) v_idx->data;
        r.k_i64 = k_idx->type == B200_TYPE_I64; r.v_i64 = v_idx->type == B200_TYPE_I64;
        r.kc = (char *) k_cache->data; r.vc = (char *) v_cache->data;
        r.kc_nb1 = (int64_t) k_cache->nb[1]; r.vc_nb1 = (int64_t) v_cache->nb[1];
        r.kc_rows = k_cache->ne[1]; r.vc_rows = v_cache->ne[1];
        ctx->rope = r;
        ctx->rope_used = false;
    }
    const b200_status st = nm ? b200_mul_mat_shared_src1_norm(ctx, 3, descs, src0s, x, nm, dsts) : b200_mul_mat_shared_src1(ctx, 3, descs, src0s, x, dsts);
    ctx->rope.on = 0;
    if (st != B200_OK) return st;
    if (fold && ctx->rope_used) return B200_OK;
    return b200_attn_prep(ctx, q, q_dst, k, k_dst, pos, freq_factors, rp, k_idx, k_cache, v, v_idx, v_cache);
}

// ------------------------------------------------------------------------------------------------ K7: fused epilogues

int b200_mul_mat_fused_supported(const b200_ctx * ctx, const b200_mul_mat_desc * d) {
    if (!ctx || !d || !b200_mul_mat_supported(d) || d->ne11 != 1) return 0;
    if (ctx->split.n_ranks > 1 || ctx->force_path == B200_PATH_SIMPLE || ctx->force_path == B200_PATH_MMQ) return 0;
    if (d->ne00 % 256 != 0 || d->nb01 % 16 || d->nb02 % 16 || d->nb03 % 16 || (d->type == B200_TYPE_Q6_K && d->ne00 % 2048 != 0)) return 0;
    mv_cfg cfg;
    return pick_mv_cfg(ctx, d->type, (int) (d->ne00 / 256), 1, &cfg) ? 1 : 0;
}

b200_status b200_mul_mat_fused(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * up, const void * gate, const float * src1,
                               const b200_fusion * f, float * dst) {
    if (!ctx || !d || !f) return fail(B200_ERR_ARG, "null argument");
    const bool glu = f->glu_op != B200_GLU_NONE;
    if (glu ? (f->glu_op < 0 || f->glu_op > B200_GLU_GEGLU_QUICK || !gate) : (gate != nullptr || f->gate_bias != nullptr))
        return fail(B200_ERR_ARG, "fused MUL_MAT: glu_op / gate / bias arguments do not agree");
    if (f->bias_nb2 % 4 || f->bias_nb3 % 4) return fail(B200_ERR_ARG, "fused MUL_MAT: bias strides");
    if (!b200_mul_mat_fused_supported(ctx, d) || !fast_layout_ok(d->type, d->ne00, up, d->nb01, d->nb02, d->nb03) ||
        (gate && (uintptr_t) gate % 16 != 0))
        return fail(B200_ERR_UNSUPPORTED, "fused MUL_MAT: one column on the TMA-staged kernel only (run the nodes one by one)");
    if (d->ne01 == 0) return B200_OK;
    if (!up || !src1 || !dst) return fail(B200_ERR_ARG, "null argument");
    CU(cudaSetDevice(ctx->device));
    const int64_t K = d->ne00, r2 = d->ne12 / d->ne02, r3 = d->ne13 / d->ne03;
    const int     U = (int) (K / 256);
    const bool    fuse = mv_fuse_quant(ctx, U, 1);
    for (int64_t i13 = 0; i13 < d->ne13; i13++)
        for (int64_t i12 = 0; i12 < d->ne12; i12++) {
            const float * x = (const float *) ((const char *) src1 + i12 * d->nb12 + i13 * d->nb13);
            uint8_t * am = nullptr, * ax = nullptr;
            if (!fuse) {  // long rows: one quantizer launch in front (as in mul_mat_impl)
                const size_t aux_o = ((size_t) U * ACT_MAIN + 255) & ~(size_t) 255;
                if (b200_status st = ensure(ctx, &ctx->scratch, &ctx->scratch_size, aux_o + (size_t) U * ACT_AUX)) return st;
                am = (uint8_t *) ctx->scratch;
                ax = am + aux_o;
                int l = 0;
                CU(quantize_units(act_kind_for(d->type), x, K, 1, d->nb11, am, ax, ctx->stream, &l, ctx->pdl));
                ctx->launches += l;
            }
            const size_t    w_off = (size_t) (i12 / r2) * d->nb02 + (size_t) (i13 / r3) * d->nb03;
            const size_t    b_off = (size_t) i12 * f->bias_nb2 + (size_t) i13 * f->bias_nb3;
            const uint8_t * ws[2]; size_t nb01s[2] = { d->nb01, d->nb01 }; int64_t Ms[2] = { d->ne01, d->ne01 }, dcs[2] = { 0, 0 };
            float *         ds[2];
            mv_epilogue     ep = {};
            ep.glu = glu ? f->glu_op : -1; ep.alpha = f->alpha; ep.limit = f->limit;
            const float * ub = f->up_bias ? (const float *) ((const char *) f->up_bias + b_off) : nullptr;
            const float * gb = f->gate_bias ? (const float *) ((const char *) f->gate_bias + b_off) : nullptr;
            ds[0] = ds[1] = (float *) ((char *) dst + i12 * d->nb2 + i13 * d->nb3);
            if (glu) {
                ws[0] = (const uint8_t *) gate + w_off; ws[1] = (const uint8_t *) up + w_off;
                ep.bias[0] = gb; ep.bias[1] = ub;
            } else {
                ws[0] = (const uint8_t *) up + w_off;
                ep.bias[0] = ub;
            }
            if (b200_status st = mv_fast_run(ctx, d->type, K, glu ? 2 : 1, ws, nb01s, Ms, ds, dcs, 1, fuse ? x : nullptr, d->nb11, am, ax, &ep)) return st;
        }
    return B200_OK;
}

b200_status b200_mul_mat_host(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * src0, int w_on_device, const float * src1_host,
                              float * dst_host) {
    if (!ctx || !d || !src0 || !src1_host || !dst_host) return fail(B200_ERR_ARG, "null argument");
    if (!b200_mul_mat_supported(d)) return fail(B200_ERR_UNSUPPORTED, "MUL_MAT geometry/type not supported");
    CU(cudaSetDevice(ctx->device));
    // host tensors are taken as contiguous in the order the strides describe (last row end = total size)
    const size_t x_bytes = (size_t) (d->ne13 - 1) * d->nb13 + (size_t) (d->ne12 - 1) * d->nb12 + (size_t) (d->ne11 - 1) * d->nb11 + (size_t) d->ne00 * 4;
    const size_t y_bytes = (size_t) (d->ne13 - 1) * d->nb3 + (size_t) (d->ne12 - 1) * d->nb2 + (size_t) (d->ne11 - 1) * d->nb1 + (size_t) d->ne01 * 4;
    const size_t w_bytes = w_on_device ? 0 : (size_t) (d->ne03 - 1) * d->nb03 + (size_t) (d->ne02 - 1) * d->nb02 + (size_t) d->ne01 * d->nb01;
    const size_t xo = 0, yo = (x_bytes + 255) & ~(size_t) 255, wo = yo + ((y_bytes + 255) & ~(size_t) 255);
    b200_status  st = ensure(ctx, &ctx->io, &ctx->io_size, wo + w_bytes + 256);
    if (st != B200_OK) return st;
    uint8_t * io = (uint8_t *) ctx->io;
    CU(cudaMemcpyAsync(io + xo, src1_host, x_bytes, cudaMemcpyHostToDevice, ctx->stream));
    const void * w = src0;
    if (!w_on_device) {
        CU(cudaMemcpyAsync(io + wo, src0, w_bytes, cudaMemcpyHostToDevice, ctx->stream));
        w = io + wo;
    }
    st = b200_mul_mat(ctx, d, w, (const float *) (io + xo), (float *) (io + yo));
    if (st != B200_OK) return st;
    CU(cudaMemcpyAsync(dst_host, io + yo, y_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}


// ------------------------------------------------------------------------------------------------ MUL_MAT without allocations
//
// b200_mul_mat_noalloc: the decode mat-vec for callers that may not allocate device memory (code under CUDA-graph capture,
// tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// It is the "pre-quantized activations" path of b200_mul_mat - K1 writes src1's unit records to scratch, the
// TMA-staged mat-vec kernels read them - with the scratch in a file-scope __device__ array instead of the context's grow-on-demand
// cudaMalloc buffer. Every GPU that loads the library holds the array (B200_NOALLOC_SCRATCH_BYTES) until exit.
// The array is shared by every call on a device and keeps the last call's records: K1 rewrites every byte the mat-vec kernels
// read (256 int8 + 16 bytes of scales / block sums + the 16-byte aux record per unit), so nothing has to start at zero, but two
// calls that overlap on the GPU (different streams or contexts of one device) race on it.

static __device__ __align__(256) uint8_t g_noalloc_scratch[B200_NOALLOC_MAX_UNITS * (ACT_MAIN + ACT_AUX)];

// host code: the symbol's name is a host-side stand-in, not a device address - ask the runtime, once per device, with it current
static b200_status noalloc_scratch_ptr(int device, uint8_t ** out) {
    static std::mutex mutex;
    static uint8_t *  ptrs[64] = {};
    if (device < 0 || device >= 64) return fail(B200_ERR_ARG, "device %d out of range", device);
    std::lock_guard<std::mutex> lock(mutex);
    if (!ptrs[device]) {
        CU(cudaSetDevice(device));
        void * p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p || (uintptr_t) p % 256 != 0) return fail(B200_ERR_CUDA, "cudaGetSymbolAddress returned an unusable address %p", p);
        ptrs[device] = (uint8_t *) p;
    }
    *out = ptrs[device];
    return B200_OK;
}

size_t b200_mul_mat_noalloc_max_units(void) { return B200_NOALLOC_MAX_UNITS; }

b200_status b200_mul_mat_noalloc(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * src0, const float * src1, float * dst) {
    if (!ctx || !d) return fail(B200_ERR_ARG, "null argument");
    if (!b200_mul_mat_supported(d)) return fail(B200_ERR_UNSUPPORTED, "MUL_MAT geometry/type not supported");
    if (d->ne02 != 1 || d->ne03 != 1 || d->ne12 != 1 || d->ne13 != 1) return fail(B200_ERR_UNSUPPORTED, "b200_mul_mat_noalloc: plain 2-D MUL_MAT only");
    if (d->ne11 < 1 || d->ne11 > MV_MAX_COLS) return fail(B200_ERR_UNSUPPORTED, "b200_mul_mat_noalloc: 1..%d activation columns (got %lld)", MV_MAX_COLS, (long long) d->ne11);
    if (d->ne00 % 256 != 0) return fail(B200_ERR_UNSUPPORTED, "b200_mul_mat_noalloc: K=%lld is not a multiple of 256", (long long) d->ne00);
    const int64_t U = d->ne00 / 256, N = d->ne11;
    if (N * U > (int64_t) B200_NOALLOC_MAX_UNITS)
        return fail(B200_ERR_UNSUPPORTED, "b200_mul_mat_noalloc: %lld columns x K=%lld need %lld activation units; the static scratch holds %d (columns x K <= %lld)",
                    (long long) N, (long long) d->ne00, (long long) (N * U), B200_NOALLOC_MAX_UNITS, (long long) B200_NOALLOC_MAX_UNITS * 256);
    if (ctx->split.n_ranks > 1) return fail(B200_ERR_UNSUPPORTED, "b200_mul_mat_noalloc: not in row-split mode");
    if (d->ne01 == 0) return B200_OK;
    if (!src0 || !src1 || !dst) return fail(B200_ERR_ARG, "null argument");
    if (!fast_layout_ok(d->type, d->ne00, src0, d->nb01, d->nb02, d->nb03))
        return fail(B200_ERR_UNSUPPORTED, "b200_mul_mat_noalloc: weights must be 16-byte aligned rows of whole 256-weight units (q6_K: K %% 2048 == 0)");
    if (N >= 3 && mv_mma_usable(d->type) && !mv_mma_fits(d->type, d->ne00)) {
        mv_cfg cfg;
        if (!pick_mv_cfg(ctx, d->type, (int) U, 1, &cfg)) return fail(B200_ERR_UNSUPPORTED, "K=%lld: activations do not fit in shared memory", (long long) d->ne00);
    }
    uint8_t * scratch = nullptr;
    if (b200_status st = noalloc_scratch_ptr(ctx->device, &scratch)) return st;
    CU(cudaSetDevice(ctx->device));
    uint8_t * am = scratch, * ax = scratch + (size_t) B200_NOALLOC_MAX_UNITS * ACT_MAIN;
    int l = 0;
    CU(quantize_units(act_kind_for(d->type), src1, d->ne00, N, d->nb11, am, ax, ctx->stream, &l, ctx->pdl));
    ctx->launches += l;
    const uint8_t * ws[1] = { (const uint8_t *) src0 };
    size_t          nb01s[1] = { d->nb01 };
    int64_t         Ms[1] = { d->ne01 }, dcs[1] = { (int64_t) (d->nb1 / 4) };
    float *         ds[1] = { dst };
    return mv_fast_run(ctx, d->type, d->ne00, 1, ws, nb01s, Ms, ds, dcs, (int) N, nullptr, d->nb11, am, ax);
}
