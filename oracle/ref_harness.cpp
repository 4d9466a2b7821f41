// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (not product code).
//
// A small C-ABI driver over the reference's PUBLIC ggml API (ggml.h, ggml-backend.h, ggml-alloc.h)
// so that Python tests / bench.py can run one GGML_OP_MUL_MAT or GGML_OP_MUL_MAT_ID graph on any
// registered ggml backend device: the reference's ggml-cpu variants (the oracle and the timed CPU
// baseline) or our own libggml-b200.so loaded through ggml_backend_load() (the drop-in check).
// It mirrors what tests/test-backend-ops.cpp does per test case (build 1-op graph, allocate on the
// backend, ggml_backend_tensor_set inputs, ggml_backend_graph_compute, ggml_backend_tensor_get).
//
// Compiled against the reference's headers where they lie (/root/reference/ggml/include) and linked
// to oracle/_ref/libggml.so + libggml-base.so; output oracle/_ref/libref_harness.so.

#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

typedef ggml_threadpool_t (*threadpool_new_t)(struct ggml_threadpool_params *);
typedef void (*threadpool_free_t)(ggml_threadpool_t);
typedef void (*set_threadpool_t)(ggml_backend_t, ggml_threadpool_t);

struct graph_run {
    ggml_context *        ctx   = nullptr;
    ggml_backend_t        be    = nullptr;
    ggml_backend_buffer_t buf   = nullptr;
    ggml_cgraph *         graph = nullptr;
    ggml_threadpool_t     tp    = nullptr;
    threadpool_free_t     tp_free = nullptr;
    ~graph_run() {
        if (buf) ggml_backend_buffer_free(buf);
        if (be) ggml_backend_free(be);
        if (tp && tp_free) tp_free(tp);
        if (ctx) ggml_free(ctx);
    }
};

std::vector<ggml_backend_reg_t> g_regs;

// A persistent thread pool for the CPU backend, attached the way llama.cpp's own tools do (common/common.cpp,
// tools/llama-bench): without one, ggml-cpu creates and joins its worker threads inside every graph_compute.
void attach_threadpool(graph_run & g, int reg_index, int n_threads) {
    if (n_threads <= 1 || reg_index < 0 || reg_index >= (int) g_regs.size() || !g.be) return;
    ggml_backend_reg_t reg = g_regs[reg_index];
    auto tp_new  = (threadpool_new_t) ggml_backend_reg_get_proc_address(reg, "ggml_threadpool_new");
    auto tp_free = (threadpool_free_t) ggml_backend_reg_get_proc_address(reg, "ggml_threadpool_free");
    auto tp_set  = (set_threadpool_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_cpu_set_threadpool");
    if (!tp_new || !tp_free || !tp_set) return;
    ggml_threadpool_params tpp = ggml_threadpool_params_default(n_threads);
    g.tp = tp_new(&tpp);
    g.tp_free = tp_free;
    if (g.tp) tp_set(g.be, g.tp);
}

ggml_backend_t open_backend(int reg_index, int dev_index, int n_threads) {
    if (reg_index < 0 || reg_index >= (int) g_regs.size()) return nullptr;
    ggml_backend_reg_t reg = g_regs[reg_index];
    if (dev_index < 0 || dev_index >= (int) ggml_backend_reg_dev_count(reg)) return nullptr;
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, dev_index);
    ggml_backend_t be = ggml_backend_dev_init(dev, nullptr);
    if (!be) return nullptr;
    if (n_threads > 0) {
        auto set_n = (ggml_backend_set_n_threads_t) ggml_backend_reg_get_proc_address(reg, "ggml_backend_set_n_threads");
        if (set_n) set_n(be, n_threads);
    }
    return be;
}

bool make_ctx(graph_run & g) {
    ggml_init_params ip = { ggml_tensor_overhead() * 64 + ggml_graph_overhead(), nullptr, true };
    g.ctx = ggml_init(ip);
    return g.ctx != nullptr;
}

bool finish(graph_run & g, ggml_tensor * out) {
    g.graph = ggml_new_graph(g.ctx);
    ggml_build_forward_expand(g.graph, out);
    g.buf = ggml_backend_alloc_ctx_tensors(g.ctx, g.be);
    return g.buf != nullptr;
}

double time_graph(graph_run & g, double seconds, int * runs_out) {
    // warm-up once, then repeat for >= `seconds` (tests/test-backend-ops.cpp:1499-1515 pattern)
    ggml_backend_graph_compute(g.be, g.graph);
    ggml_backend_synchronize(g.be);
    int runs = 0;
    auto t0 = std::chrono::steady_clock::now();
    double el = 0;
    do {
        ggml_backend_graph_compute(g.be, g.graph);
        ggml_backend_synchronize(g.be);
        runs++;
        el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    } while (el < seconds);
    if (runs_out) *runs_out = runs;
    return el / runs * 1e6;
}

} // namespace

extern "C" {

// load a backend shared library (a ggml-cpu variant or libggml-b200.so); returns a handle >= 0, or -1
int refh_load_backend(const char * path) {
    ggml_backend_reg_t reg = ggml_backend_load(path);
    if (!reg) return -1;
    g_regs.push_back(reg);
    return (int) g_regs.size() - 1;
}

const char * refh_reg_name(int reg) { return reg >= 0 && reg < (int) g_regs.size() ? ggml_backend_reg_name(g_regs[reg]) : ""; }
int refh_reg_dev_count(int reg) { return reg >= 0 && reg < (int) g_regs.size() ? (int) ggml_backend_reg_dev_count(g_regs[reg]) : 0; }
const char * refh_dev_description(int reg, int dev) {
    if (reg < 0 || reg >= (int) g_regs.size() || dev < 0 || dev >= (int) ggml_backend_reg_dev_count(g_regs[reg])) return "";
    return ggml_backend_dev_description(ggml_backend_reg_dev_get(g_regs[reg], dev));
}

// dst[M, N, ne12, ne13] = MUL_MAT(W[K, M, ne02, ne03] (type), X[K, N, ne12, ne13] f32); contiguous tensors.
// seconds > 0: also time it, *us_per_run receives the mean. returns 0 on success.
int refh_mul_mat(int reg, int dev, int n_threads, int type, int64_t M, int64_t K, int64_t N,
                 int64_t ne02, int64_t ne03, int64_t ne12, int64_t ne13,
                 const void * W, const float * X, float * dst, double seconds, double * us_per_run) {
    graph_run g;
    g.be = open_backend(reg, dev, n_threads);
    if (!g.be || !make_ctx(g)) return 1;
    if (seconds > 0) attach_threadpool(g, reg, n_threads);
    ggml_tensor * w = ggml_new_tensor_4d(g.ctx, (ggml_type) type, K, M, ne02, ne03);
    ggml_tensor * x = ggml_new_tensor_4d(g.ctx, GGML_TYPE_F32, K, N, ne12, ne13);
    ggml_tensor * o = ggml_mul_mat(g.ctx, w, x);
    if (!ggml_backend_supports_op(g.be, o)) return 2;
    if (!finish(g, o)) return 3;
    ggml_backend_tensor_set(w, W, 0, ggml_nbytes(w));
    ggml_backend_tensor_set(x, X, 0, ggml_nbytes(x));
    if (ggml_backend_graph_compute(g.be, g.graph) != GGML_STATUS_SUCCESS) return 4;
    ggml_backend_synchronize(g.be);
    if (dst) ggml_backend_tensor_get(o, dst, 0, ggml_nbytes(o));
    if (seconds > 0 && us_per_run) *us_per_run = time_graph(g, seconds, nullptr);
    return 0;
}

// The same MUL_MAT with W in the backend's SPLIT buffer type (ggml_backend_split_buffer_type, ggml-backend.h:199-200: weight rows
// spread over the devices of the registry; how llama places weights with --split-mode row, src/llama-model.cpp:385-403).
// tensor_split: n_split proportions (nullptr / all zero = the backend's default). W_back (optional) receives the weight bytes read
// back through the split buffer. returns 0 on success, 5 if the registry has no split buffer type.
int refh_mul_mat_split(int reg, int dev, int type, int64_t M, int64_t K, int64_t N, const void * W, const float * X, float * dst,
                       const float * tensor_split, int n_split, void * W_back, int repeats) {
    if (reg < 0 || reg >= (int) g_regs.size()) return 1;
    auto split_fn = (ggml_backend_split_buffer_type_t) ggml_backend_reg_get_proc_address(g_regs[reg], "ggml_backend_split_buffer_type");
    if (!split_fn) return 5;
    float ts[64] = {};
    for (int i = 0; i < n_split && i < 64; i++) ts[i] = tensor_split ? tensor_split[i] : 0.0f;
    ggml_backend_buffer_type_t sbuft = split_fn(dev, ts);
    if (!sbuft) return 5;
    graph_run g;
    g.be = open_backend(reg, dev, 0);
    if (!g.be) return 1;
    // weights in their own context: allocated from the split buffer type; activations and result from the device's default one
    ggml_init_params ipw = { ggml_tensor_overhead() * 4, nullptr, true };
    ggml_context * wctx = ggml_init(ipw);
    if (!wctx || !make_ctx(g)) return 1;
    ggml_tensor * w = ggml_new_tensor_2d(wctx, (ggml_type) type, K, M);
    ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors_from_buft(wctx, sbuft);
    int rc = 0;
    if (!wbuf) {
        rc = 3;
    } else {
        ggml_tensor * x = ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, K, N);
        ggml_tensor * o = ggml_mul_mat(g.ctx, w, x);
        if (!ggml_backend_supports_op(g.be, o)) rc = 2;
        else if (!finish(g, o)) rc = 3;
        else {
            ggml_backend_tensor_set(w, W, 0, ggml_nbytes(w));
            if (W_back) ggml_backend_tensor_get(w, W_back, 0, ggml_nbytes(w));
            ggml_backend_tensor_set(x, X, 0, ggml_nbytes(x));
            for (int r = 0; r < (repeats > 0 ? repeats : 1) && rc == 0; r++) {
                if (ggml_backend_graph_compute(g.be, g.graph) != GGML_STATUS_SUCCESS) rc = 4;
            }
            ggml_backend_synchronize(g.be);
            if (rc == 0 && dst) ggml_backend_tensor_get(o, dst, 0, ggml_nbytes(o));
        }
    }
    if (wbuf) ggml_backend_buffer_free(wbuf);
    ggml_free(wctx);
    return rc;
}

// dst[M, n_used, n_tok] = MUL_MAT_ID(as[K, M, n_as] (type), b[K, ne11, n_tok] f32, ids[n_used, n_tok] i32)
int refh_mul_mat_id(int reg, int dev, int n_threads, int type, int64_t M, int64_t K, int64_t n_as,
                    int64_t n_used, int64_t ne11, int64_t n_tok,
                    const void * W, const float * B, const int32_t * ids, float * dst,
                    double seconds, double * us_per_run) {
    graph_run g;
    g.be = open_backend(reg, dev, n_threads);
    if (!g.be || !make_ctx(g)) return 1;
    if (seconds > 0) attach_threadpool(g, reg, n_threads);
    ggml_tensor * w = ggml_new_tensor_3d(g.ctx, (ggml_type) type, K, M, n_as);
    ggml_tensor * b = ggml_new_tensor_3d(g.ctx, GGML_TYPE_F32, K, ne11, n_tok);
    ggml_tensor * i = ggml_new_tensor_2d(g.ctx, GGML_TYPE_I32, n_used, n_tok);
    ggml_tensor * o = ggml_mul_mat_id(g.ctx, w, b, i);
    if (!ggml_backend_supports_op(g.be, o)) return 2;
    if (!finish(g, o)) return 3;
    ggml_backend_tensor_set(w, W, 0, ggml_nbytes(w));
    ggml_backend_tensor_set(b, B, 0, ggml_nbytes(b));
    ggml_backend_tensor_set(i, ids, 0, ggml_nbytes(i));
    if (ggml_backend_graph_compute(g.be, g.graph) != GGML_STATUS_SUCCESS) return 4;
    ggml_backend_synchronize(g.be);
    if (dst) ggml_backend_tensor_get(o, dst, 0, ggml_nbytes(o));
    if (seconds > 0 && us_per_run) *us_per_run = time_graph(g, seconds, nullptr);
    return 0;
}

// A stack of `n_mats` independent MUL_MATs in ONE graph (distinct weight tensors, one shared or per-op
// activation), the CPU-baseline counterpart of bench.py's layer-stack step. Ms/Ks give each matrix' shape.
// Weights: W_all is the concatenation of the matrices' bytes. Returns us per graph run in *us_per_run.
// `reps` repeats the whole stack inside the graph (fresh result tensors over the same inputs), as eval_perf duplicates its
// node (tests/test-backend-ops.cpp:1449-1515), so that one graph_compute is long against the thread pool's wake-up cost;
// *us_per_run is per repetition.
int refh_mul_mat_stack(int reg, int dev, int n_threads, int type, int n_mats, const int64_t * Ms, const int64_t * Ks,
                       int64_t N, const void * W_all, const float * X_all, double seconds, int reps, double * us_per_run, int * runs) {
    if (reps < 1) reps = 1;
    graph_run g;
    g.be = open_backend(reg, dev, n_threads);
    if (!g.be) return 1;
    attach_threadpool(g, reg, n_threads);
    const size_t n_nodes = (size_t) (2 + reps) * n_mats + 16;
    ggml_init_params ip = { ggml_tensor_overhead() * n_nodes + ggml_graph_overhead_custom(n_nodes, false), nullptr, true };
    g.ctx = ggml_init(ip);
    if (!g.ctx) return 1;
    std::vector<ggml_tensor *> ws(n_mats), xs(n_mats), os(n_mats);
    g.graph = ggml_new_graph_custom(g.ctx, n_nodes, false);
    for (int i = 0; i < n_mats; i++) {
        ws[i] = ggml_new_tensor_2d(g.ctx, (ggml_type) type, Ks[i], Ms[i]);
        xs[i] = ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, Ks[i], N);
    }
    for (int r = 0; r < reps; r++)
        for (int i = 0; i < n_mats; i++) {
            os[i] = ggml_mul_mat(g.ctx, ws[i], xs[i]);
            ggml_build_forward_expand(g.graph, os[i]);
        }
    g.buf = ggml_backend_alloc_ctx_tensors(g.ctx, g.be);
    if (!g.buf) return 3;
    const char *  wp = (const char *) W_all;
    const float * xp = X_all;
    for (int i = 0; i < n_mats; i++) {
        ggml_backend_tensor_set(ws[i], wp, 0, ggml_nbytes(ws[i]));
        ggml_backend_tensor_set(xs[i], xp, 0, ggml_nbytes(xs[i]));
        wp += ggml_nbytes(ws[i]);
        xp += Ks[i] * N;
    }
    *us_per_run = time_graph(g, seconds, runs) / reps;
    if (runs) *runs *= reps;
    return 0;
}

// The MUL_MAT stack of one decode / prefill step as ONE ggml_cgraph on any registered backend, driven the way a host drives a
// backend: activations from (pinned) host memory with ggml_backend_tensor_set_async, ggml_backend_graph_compute_async, the step's
// results back with ggml_backend_tensor_get_async, ggml_backend_synchronize. Matrix i multiplies activation tensor x_of[i]
// (q/k/v and gate/up share theirs, as in the llama graph, src/llama-graph.cpp:612-660). The dst tensors are placed back to back
// in one buffer (ggml_backend_tensor_alloc), so the results of a step are one copy. Wall clock:
//   *us_compute  graph_compute + synchronize only
//   *us_e2e      set_async of every activation + graph_compute + get_async of the results + synchronize
// out_all (optional): the concatenated results of the last step. returns 0 on success.
int refh_stack_e2e(int reg, int dev, int n_threads, int type, int n_mats, const int64_t * Ms, const int64_t * Ks, const int32_t * x_of,
                   int n_x, const int64_t * x_K, int64_t N, const void * W_all, const float * X_all, int warmup, int steps,
                   float * out_all, double * us_compute, double * us_e2e) {
    graph_run g;
    g.be = open_backend(reg, dev, n_threads);
    if (!g.be) return 1;
    attach_threadpool(g, reg, n_threads);
    ggml_backend_dev_t device = ggml_backend_get_device(g.be);
    const size_t n_nodes = (size_t) 3 * n_mats + n_x + 16;
    ggml_init_params ip  = { ggml_tensor_overhead() * n_nodes + ggml_graph_overhead_custom(n_nodes, false), nullptr, true };
    ggml_init_params ipo = { ggml_tensor_overhead() * ((size_t) n_mats + 8) + ggml_graph_overhead_custom(n_nodes, false), nullptr, true };  // results + pool + the cgraph
    g.ctx = ggml_init(ip);
    ggml_context * octx = ggml_init(ipo);
    if (!g.ctx || !octx) return 1;
    std::vector<ggml_tensor *> ws(n_mats), xs(n_x), os(n_mats);
    for (int i = 0; i < n_mats; i++) ws[i] = ggml_new_tensor_2d(g.ctx, (ggml_type) type, Ks[i], Ms[i]);
    for (int j = 0; j < n_x; j++) xs[j] = ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, x_K[j], N);
    g.buf = ggml_backend_alloc_ctx_tensors(g.ctx, g.be);
    if (!g.buf) { ggml_free(octx); return 3; }
    g.graph = ggml_new_graph_custom(octx, n_nodes, false);
    size_t out_floats = 0;
    for (int i = 0; i < n_mats; i++) {
        os[i] = ggml_mul_mat(octx, ws[i], xs[x_of[i]]);
        if (!ggml_backend_supports_op(g.be, os[i])) { ggml_free(octx); return 2; }
        ggml_build_forward_expand(g.graph, os[i]);
        out_floats += (size_t) Ms[i] * N;
    }
    ggml_tensor * pool = ggml_new_tensor_1d(octx, GGML_TYPE_F32, (int64_t) out_floats);
    ggml_backend_buffer_t obuf = ggml_backend_alloc_buffer(g.be, out_floats * sizeof(float) + 256);
    if (!obuf) { ggml_free(octx); return 3; }
    char * base = (char *) ggml_backend_buffer_get_base(obuf);
    size_t off = 0;
    for (int i = 0; i < n_mats; i++) {
        ggml_backend_tensor_alloc(obuf, os[i], base + off);
        off += ggml_nbytes(os[i]);
    }
    ggml_backend_tensor_alloc(obuf, pool, base);
    const char * wp = (const char *) W_all;
    for (int i = 0; i < n_mats; i++) {
        ggml_backend_tensor_set(ws[i], wp, 0, ggml_nbytes(ws[i]));
        wp += ggml_nbytes(ws[i]);
    }
    // host side of the copies: the device's pinned host buffer type where it has one (llama's input / output buffers)
    size_t x_bytes = 0;
    for (int j = 0; j < n_x; j++) x_bytes += ggml_nbytes(xs[j]);
    ggml_backend_buffer_type_t hbuft = ggml_backend_dev_host_buffer_type(device);
    ggml_backend_buffer_t hbuf = hbuft ? ggml_backend_buft_alloc_buffer(hbuft, x_bytes + out_floats * sizeof(float) + 512) : nullptr;
    std::vector<char> pageable;
    char * hx = nullptr;
    if (hbuf) hx = (char *) ggml_backend_buffer_get_base(hbuf);
    else { pageable.resize(x_bytes + out_floats * sizeof(float) + 512); hx = pageable.data(); }
    char * hy = hx + ((x_bytes + 255) & ~(size_t) 255);
    memcpy(hx, X_all, x_bytes);

    auto step_e2e = [&]() {
        size_t o = 0;
        for (int j = 0; j < n_x; j++) {
            ggml_backend_tensor_set_async(g.be, xs[j], hx + o, 0, ggml_nbytes(xs[j]));
            o += ggml_nbytes(xs[j]);
        }
        ggml_backend_graph_compute_async(g.be, g.graph);
        ggml_backend_tensor_get_async(g.be, pool, hy, 0, out_floats * sizeof(float));
        ggml_backend_synchronize(g.be);
    };
    int rc = 0;
    for (int i = 0; i < warmup + 2; i++) step_e2e();
    auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) {
        if (ggml_backend_graph_compute_async(g.be, g.graph) != GGML_STATUS_SUCCESS) rc = 4;
        ggml_backend_synchronize(g.be);
    }
    if (us_compute) *us_compute = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / steps * 1e6;
    t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) step_e2e();
    if (us_e2e) *us_e2e = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / steps * 1e6;
    if (out_all) memcpy(out_all, hy, out_floats * sizeof(float));
    if (hbuf) ggml_backend_buffer_free(hbuf);
    ggml_backend_buffer_free(obuf);
    ggml_free(octx);
    return rc;
}

// calls a `uint64_t (*)(void)` the registry exposes through get_proc_address (e.g. a backend's own counters); ~0 if absent
uint64_t refh_proc_u64(int reg, const char * name) {
    if (reg < 0 || reg >= (int) g_regs.size()) return ~(uint64_t) 0;
    auto fn = (uint64_t (*)(void)) ggml_backend_reg_get_proc_address(g_regs[reg], name);
    return fn ? fn() : ~(uint64_t) 0;
}

// K7 patterns (test_mul_mat_vec_fusion, tests/test-backend-ops.cpp:5120-5249) as ONE graph on any registered backend:
//   glu_op >= 0: out = GLU(W_gate . X [+ gate_bias], W_up . X [+ up_bias])   (ggml_glu_split / ggml_swiglu_oai)
//   glu_op <  0: out = W_up . X + up_bias
// W_*: [K, M] of `type`; X: F32 [K, N]; biases: F32 [M] (added to every column; nullptr = none).
// returns 0 on success, 2 if the backend does not support one of the nodes.
int refh_mul_mat_glu(int reg, int dev, int n_threads, int type, int64_t M, int64_t K, int64_t N, const void * W_up, const void * W_gate,
                     const float * X, const float * up_bias, const float * gate_bias, int glu_op, float alpha, float limit, float * dst) {
    graph_run g;
    g.be = open_backend(reg, dev, n_threads);
    if (!g.be || !make_ctx(g)) return 1;
    ggml_tensor * x  = ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, K, N);
    ggml_tensor * wu = ggml_new_tensor_2d(g.ctx, (ggml_type) type, K, M);
    ggml_tensor * wg = glu_op >= 0 ? ggml_new_tensor_2d(g.ctx, (ggml_type) type, K, M) : nullptr;
    ggml_tensor * bu = up_bias ? ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, M, 1) : nullptr;
    ggml_tensor * bg = gate_bias && wg ? ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, M, 1) : nullptr;
    ggml_tensor * up = ggml_mul_mat(g.ctx, wu, x);
    if (bu) up = ggml_add(g.ctx, up, bu);
    ggml_tensor * out = up;
    if (wg) {
        ggml_tensor * gate = ggml_mul_mat(g.ctx, wg, x);
        if (bg) gate = ggml_add(g.ctx, gate, bg);
        out = glu_op == GGML_GLU_OP_SWIGLU_OAI ? ggml_swiglu_oai(g.ctx, gate, up, alpha, limit) : ggml_glu_split(g.ctx, gate, up, (ggml_glu_op) glu_op);
    }
    if (!finish(g, out)) return 3;
    for (int i = 0; i < ggml_graph_n_nodes(g.graph); i++)
        if (!ggml_backend_supports_op(g.be, ggml_graph_node(g.graph, i))) return 2;
    ggml_backend_tensor_set(x, X, 0, ggml_nbytes(x));
    ggml_backend_tensor_set(wu, W_up, 0, ggml_nbytes(wu));
    if (wg) ggml_backend_tensor_set(wg, W_gate, 0, ggml_nbytes(wg));
    if (bu) ggml_backend_tensor_set(bu, up_bias, 0, ggml_nbytes(bu));
    if (bg) ggml_backend_tensor_set(bg, gate_bias, 0, ggml_nbytes(bg));
    if (ggml_backend_graph_compute(g.be, g.graph) != GGML_STATUS_SUCCESS) return 4;
    ggml_backend_synchronize(g.be);
    ggml_backend_tensor_get(out, dst, 0, ggml_nbytes(out));
    return 0;
}

// the MUL_MAT_ID form: experts W_*: [K, M, n_as]; B: F32 [K, ne11, n_tok]; ids: I32 [n_used, n_tok]; biases: F32 [M, n_as] (ADD_ID)
int refh_mul_mat_id_glu(int reg, int dev, int n_threads, int type, int64_t M, int64_t K, int64_t n_as, int64_t n_used, int64_t ne11, int64_t n_tok,
                        const void * W_up, const void * W_gate, const float * B, const int32_t * ids, const float * up_bias, const float * gate_bias,
                        int glu_op, float alpha, float limit, float * dst) {
    graph_run g;
    g.be = open_backend(reg, dev, n_threads);
    if (!g.be || !make_ctx(g)) return 1;
    ggml_tensor * b  = ggml_new_tensor_3d(g.ctx, GGML_TYPE_F32, K, ne11, n_tok);
    ggml_tensor * id = ggml_new_tensor_2d(g.ctx, GGML_TYPE_I32, n_used, n_tok);
    ggml_tensor * wu = ggml_new_tensor_3d(g.ctx, (ggml_type) type, K, M, n_as);
    ggml_tensor * wg = glu_op >= 0 ? ggml_new_tensor_3d(g.ctx, (ggml_type) type, K, M, n_as) : nullptr;
    ggml_tensor * bu = up_bias ? ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, M, n_as) : nullptr;
    ggml_tensor * bg = gate_bias && wg ? ggml_new_tensor_2d(g.ctx, GGML_TYPE_F32, M, n_as) : nullptr;
    ggml_tensor * up = ggml_mul_mat_id(g.ctx, wu, b, id);
    if (bu) up = ggml_add_id(g.ctx, up, bu, id);
    ggml_tensor * out = up;
    if (wg) {
        ggml_tensor * gate = ggml_mul_mat_id(g.ctx, wg, b, id);
        if (bg) gate = ggml_add_id(g.ctx, gate, bg, id);
        out = glu_op == GGML_GLU_OP_SWIGLU_OAI ? ggml_swiglu_oai(g.ctx, gate, up, alpha, limit) : ggml_glu_split(g.ctx, gate, up, (ggml_glu_op) glu_op);
    }
    if (!finish(g, out)) return 3;
    for (int i = 0; i < ggml_graph_n_nodes(g.graph); i++)
        if (!ggml_backend_supports_op(g.be, ggml_graph_node(g.graph, i))) return 2;
    ggml_backend_tensor_set(b, B, 0, ggml_nbytes(b));
    ggml_backend_tensor_set(id, ids, 0, ggml_nbytes(id));
    ggml_backend_tensor_set(wu, W_up, 0, ggml_nbytes(wu));
    if (wg) ggml_backend_tensor_set(wg, W_gate, 0, ggml_nbytes(wg));
    if (bu) ggml_backend_tensor_set(bu, up_bias, 0, ggml_nbytes(bu));
    if (bg) ggml_backend_tensor_set(bg, gate_bias, 0, ggml_nbytes(bg));
    if (ggml_backend_graph_compute(g.be, g.graph) != GGML_STATUS_SUCCESS) return 4;
    ggml_backend_synchronize(g.be);
    ggml_backend_tensor_get(out, dst, 0, ggml_nbytes(out));
    return 0;
}

} // extern "C"
